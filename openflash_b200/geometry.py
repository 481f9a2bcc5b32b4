"""Reference module path ``openflash.geometry``."""
from .model import BodyArrangement, ConcentricBodyGroup, Geometry  # noqa: F401
