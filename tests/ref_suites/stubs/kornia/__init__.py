"""Stand-in for kornia 0.6.12 (absent from this image) backed by oracle/ — see ../README.md.  UNPINNED."""
from . import color, losses  # noqa: F401

__version__ = "0.6.12+oracle-restatement"
