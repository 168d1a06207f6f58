"""Oracle (TEST INFRASTRUCTURE ONLY): HybridLossPlus restricted to the terms on the hot path
(L1 + DeltaE00-loss + DSSIM + physics), i.e. NewBP_model/losses.py:318-372 with the VGG
perceptual / LPIPS terms left out (those stay stock PyTorch).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import colour, psf_ops, quality


def hybrid_loss(*, Bhat_raw, B_raw, A_raw, expo_ratio, Bhat_srgb01, B_srgb01, A_srgb01=None,
                w_l1=1.0, w_de=0.02, w_ssim=0.05, w_phys=0.10, physics_kernel=None, psf_kernel=None):
    """Returns (total, logs) like losses.py:329-372.  `physics_kernel` selects the RAW physics term
    (losses.py:360-363), else `psf_kernel` (normalised CrosstalkPSF buffer) + A_srgb01 selects the sRGB
    one (losses.py:364-369)."""
    logs = {}
    l1 = F.l1_loss(Bhat_raw, B_raw)
    logs["L1_raw"] = l1.detach()
    total = w_l1 * l1
    de = colour.de00_loss(Bhat_srgb01, B_srgb01)
    logs["DeltaE"] = de.detach()
    total = total + w_de * de
    ss = quality.ssim_loss_kornia(Bhat_srgb01, B_srgb01)
    logs["SSIM"] = ss.detach()
    total = total + w_ssim * ss
    if physics_kernel is not None:
        ph = psf_ops.phys_loss_raw(Bhat_raw, A_raw, expo_ratio, physics_kernel)
        logs["Phys"] = ph.detach()
        total = total + w_phys * ph
    elif psf_kernel is not None and A_srgb01 is not None:
        ph = psf_ops.phys_loss_srgb(Bhat_srgb01, A_srgb01, expo_ratio, psf_kernel)
        logs["Phys"] = ph.detach()
        total = total + w_phys * ph
    logs["Total"] = total.detach()
    return total, logs
