"""kornia.color.{rgb_to_lab, lab_to_rgb} restated in oracle/colour.py (unpinned)."""
from oracle.colour import lab_to_rgb, rgb_to_lab  # noqa: F401
