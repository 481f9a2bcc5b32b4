"""Reference module path ``openflash.basic_region_geometry``."""
from .model import BasicRegionGeometry  # noqa: F401
