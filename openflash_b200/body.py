"""Reference module path ``openflash.body``."""
from .model import Body, SteppedBody, CoordinateBody  # noqa: F401
