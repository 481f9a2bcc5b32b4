"""Reference module path ``openflash.meem_problem``."""
from .model import MEEMProblem  # noqa: F401
