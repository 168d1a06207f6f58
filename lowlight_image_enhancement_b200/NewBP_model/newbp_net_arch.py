"""Mirror of the reference's `NewBP_model/newbp_net_arch.py` (Scenario B: the PSF lives in the loss branch only).

`create_crosstalk_psf` (newbp_net_arch.py:88-99) is on the hot path's boundary and builds the sm_100a
`CrosstalkPSF`.  `create_newbp_net` (newbp_net_arch.py:31-85) returns the backbone, which stays stock PyTorch: here
the plain `nafnet_stock.NAFNet` (the reference imports BasicSR's NAFNet; same constructor keywords).
"""
from __future__ import annotations

from typing import Any, Dict, Optional, Sequence

from .nafnet_stock import NAFNet
from .newbp_layer import CrosstalkPSF, build_psf_kernels, create_crosstalk_psf  # noqa: F401

__all__ = ["NAFNet", "CrosstalkPSF", "build_psf_kernels", "create_crosstalk_psf", "create_newbp_net"]


def create_newbp_net(in_channels: int = 3, kernel_type: str = "panchromatic", kernel_spec: str = "P2",
                     width: Optional[int] = None, enc_blk_nums: Optional[Sequence[int]] = None,
                     middle_blk_num: Optional[int] = None, dec_blk_nums: Optional[Sequence[int]] = None,
                     nafnet_params: Optional[Dict[str, Any]] = None, **nafnet_kwargs: Any) -> NAFNet:
    """Backbone only; `kernel_type` / `kernel_spec` are accepted and ignored exactly as upstream (no input-side K)."""
    if nafnet_params is not None and not isinstance(nafnet_params, dict):
        raise TypeError("nafnet_params must be a dictionary if provided.")
    cfg: Dict[str, Any] = dict(nafnet_params or {})
    cfg.update(nafnet_kwargs)
    cfg["img_channel"] = in_channels
    for key, val in (("width", width), ("middle_blk_num", middle_blk_num)):
        if val is not None:
            cfg[key] = val
    for key, val in (("enc_blk_nums", enc_blk_nums), ("dec_blk_nums", dec_blk_nums)):
        if val is not None:
            cfg[key] = list(val)
    return NAFNet(**cfg)
