"""Stand-in for torchmetrics 1.2.0 (absent from this image) backed by oracle/ — see ../README.md.  UNPINNED."""
from . import image  # noqa: F401

__version__ = "1.2.0+oracle-restatement"
