"""B200-native pieces behind the template's plug-in surface."""
from . import native  # noqa: F401
