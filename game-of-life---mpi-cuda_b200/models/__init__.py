"""Model layer: the Life stepper and its initial patterns."""
from .life import Life
from .patterns import PATTERNS, pattern_bytes

__all__ = ["Life", "PATTERNS", "pattern_bytes"]
