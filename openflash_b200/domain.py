"""Reference module path ``openflash.domain``."""
from .model import Domain  # noqa: F401
