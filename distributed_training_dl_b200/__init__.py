"""B200-native data-parallel training framework (see README.md / DESIGN.md)."""
__version__ = "0.1.0"

from . import _native  # noqa: F401
from . import models, ops  # noqa: F401
