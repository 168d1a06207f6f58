"""Oracle (TEST INFRASTRUCTURE ONLY): NewBP crosstalk-PSF operator and the
physics-consistency projector, restated with CPU torch ops.

Follows (paths relative to /root/reference):
  * NewBP_model/newbp_layer.py:7-21    NewBPFunction  (conv2d fwd, conv_transpose2d bwd)
  * NewBP_model/newbp_layer.py:88-126  CrosstalkPSF   (per-kernel sum normalisation, groups=3, pad 1)
  * NewBP_model/newbp_layer.py:129-173 build_psf_kernels (P2 / B2 constants)
  * metrics/phys_consistency.py:75-255 _prepare_psf / _apply_psf / _expand_exposure /
                                        _crop_valid / _phys_cons_core
  * NewBP_model/losses.py:158-192      PhysicsConsistencyLoss (replicate pad, clamp(A*rho))
  * NewBP_model/losses.py:195-220      align_exposure_srgb + PhysicalConsistencyLossSRGB
"""
from __future__ import annotations

import warnings

import torch
import torch.nn.functional as F

P2 = [[0.01, 0.02, 0.01], [0.02, 0.88, 0.02], [0.01, 0.02, 0.01]]
B2_R = [[0.0117, 0.0233, 0.0117], [0.0233, 0.86, 0.0233], [0.0117, 0.0233, 0.0117]]
B2_B = [[0.0083, 0.0167, 0.0083], [0.0167, 0.90, 0.0167], [0.0083, 0.0167, 0.0083]]


def canonical_kernels(mode: str, spec: str | None = None) -> torch.Tensor:
    """newbp_layer.py:129-173 — P2 mono [1,1,3,3] or B2 rgb [3,1,3,3]."""
    if mode == "mono":
        return torch.tensor(P2, dtype=torch.float32).view(1, 1, 3, 3)
    if mode == "rgb":
        return torch.tensor([B2_R, P2, B2_B], dtype=torch.float32).view(3, 1, 3, 3)
    raise ValueError(mode)


def stressed_kernels(stress: float = 2.0) -> torch.Tensor:
    """Bench-defined 'stressed R>G>B' family (the reference names the
    family in README.md:24,64,214 but ships no numbers): K_c = delta + s*alpha_c*offcentre(P2),
    alpha=(1.3,1.0,0.7), renormalised to sum 1."""
    p2 = torch.tensor(P2, dtype=torch.float64)
    off = p2.clone()
    off[1, 1] = 0.0
    ks = []
    for alpha in (1.3, 1.0, 0.7):
        k = off * (stress * alpha)
        k[1, 1] = 1.0
        ks.append(k / k.sum())
    return torch.stack(ks).to(torch.float32).view(3, 1, 3, 3)


def normalise_crosstalk(kernels: torch.Tensor) -> torch.Tensor:
    """newbp_layer.py:102-106 — K / clamp_min(sum K, 1e-12) per leading index."""
    s = kernels.reshape(kernels.shape[0], -1).sum(dim=1, keepdim=True).clamp_min(1e-12)
    return kernels / s.view(-1, 1, 1, 1)


def crosstalk_forward(x: torch.Tensor, kernel: torch.Tensor) -> torch.Tensor:
    """newbp_layer.py:117-126 — depthwise zero-padded 'same' 3x3 on 3 channels.
    `kernel` is the already-normalised buffer, [1,1,3,3] (mono) or [3,1,3,3]."""
    k = kernel.expand(3, 1, 3, 3) if kernel.shape[0] == 1 else kernel
    return F.conv2d(x, k, bias=None, stride=1, padding=1, groups=3)


def newbp_forward(x, kernel, padding, groups):
    """newbp_layer.py:9-13."""
    return F.conv2d(x, kernel, padding=padding, groups=groups)


def newbp_adjoint(gy, kernel, padding, groups):
    """newbp_layer.py:16-21 — the explicit NewBP backward."""
    return F.conv_transpose2d(gy, kernel, padding=padding, groups=groups)


# ----------------------------------------------------------------------------
# metric convention (phys_cons_raw / phys_cons_srgb): rho scales and clips K*pred
# ----------------------------------------------------------------------------
def prepare_dense_psf(psf, c_in, c_out, dtype, normalize, nonneg, eps):
    """phys_consistency.py:75-127 (dense [C_out,C_in,kh,kw] only, as in the reference)."""
    if psf.ndim == 2:
        psf = psf[None, None]
    if psf.ndim != 4:
        raise ValueError("psf must be 2-D or 4-D")
    co, ci, kh, kw = psf.shape
    if co != c_out or ci != c_in:
        raise ValueError("psf channel mismatch")
    if kh % 2 == 0 or kw % 2 == 0:
        raise ValueError("psf must be odd-sized")
    psf = psf.to(dtype=dtype)
    if nonneg:
        psf = psf.clamp_min(0)
    if normalize:
        sums = psf.reshape(co, -1).sum(dim=1)
        dead = sums.abs() < eps
        if bool(dead.any()):
            warnings.warn("PSF channel sums near zero", RuntimeWarning)
        psf = psf / torch.where(dead, torch.ones_like(sums), sums).view(co, 1, 1, 1)
    return psf


def expand_rho(rho, ref):
    """phys_consistency.py:160-190,230-232."""
    r = rho.to(dtype=ref.dtype) if torch.is_tensor(rho) else torch.tensor(float(rho), dtype=ref.dtype, device=ref.device)
    n, c = ref.shape[:2]
    if r.ndim == 0:
        r = r.view(1, 1, 1, 1).expand(n, 1, 1, 1)
    elif r.ndim == 1:
        if r.shape[0] != n:
            raise ValueError("rho length")
        r = r.view(n, 1, 1, 1)
    elif r.ndim == 4:
        if r.shape[0] != n or r.shape[1] not in (1, c):
            raise ValueError("rho shape")
    else:
        raise ValueError("rho ndim")
    if r.shape[1] == 1 and c != 1:
        r = r.expand(n, c, r.shape[2], r.shape[3])
    return r


def phys_metric(pred, obs, psf, rho, *, reduction="mean", padding="reflect", normalize_psf=True,
                enforce_nonnegative=False, crop="valid", robust="none", clamp01=False,
                return_map=False, eps=1e-12, keep_graph=False):
    """phys_consistency.py:193-255 with the wrappers' fp32 cast (:300-301, :349-350).
    keep_graph=True skips the reference's detach so autograd of the same ops gives d/d pred (used as
    the gradient oracle and as the CPU fwd+bwd baseline)."""
    if pred.ndim == 3:
        pred, obs = pred[None], obs[None]
    pred = (pred if keep_graph else pred.detach()).to(torch.float32)
    obs = obs.detach().to(torch.float32)
    k = prepare_dense_psf(psf, pred.shape[1], obs.shape[1], pred.dtype, normalize_psf,
                          enforce_nonnegative, eps)
    kh, kw = k.shape[-2:]
    ph, pw = kh // 2, kw // 2
    if padding == "zeros":
        y = F.conv2d(pred, k, padding=(ph, pw))
    else:
        y = F.conv2d(F.pad(pred, (pw, pw, ph, ph), mode=padding), k)
    y = y * expand_rho(rho, y)
    if clamp01:
        y = y.clamp(0.0, 1.0)
    if crop == "valid":
        if ph:
            y, obs = y[..., ph:-ph, :], obs[..., ph:-ph, :]
        if pw:
            y, obs = y[..., :, pw:-pw], obs[..., :, pw:-pw]
    d = y - obs
    mag = d.abs()
    lossmap = torch.sqrt(d * d + eps * eps) if robust == "charbonnier" else mag
    per = lossmap.flatten(1).mean(dim=1)
    out = per if reduction == "none" else (per.mean(0) if reduction == "mean" else per.sum(0))
    return (out, mag) if return_map else out


# ----------------------------------------------------------------------------
# loss convention (training): rho scales and clips A, global mean
# ----------------------------------------------------------------------------
def phys_loss_raw(bhat, a, rho, kernel, clamp_align=True):
    """losses.py:173-192 — replicate pad, depthwise (or C_in-expanded dense) valid conv, L1 mean."""
    if rho.dim() == 1:
        rho = rho.view(-1, 1, 1, 1)
    target = a * rho
    if clamp_align:
        target = target.clamp(0.0, 1.0)
    kh, kw = kernel.shape[-2:]
    x = F.pad(bhat, (kw // 2, kw // 2, kh // 2, kh // 2), mode="replicate")
    c = bhat.shape[1]
    k = kernel
    if k.shape[0] == 1 and c > 1:
        k = k.expand(c, 1, kh, kw)
    groups = c if k.shape[0] == c else 1
    if groups == 1 and k.shape[1] == 1 and c != 1:
        k = k.expand(k.shape[0], c, kh, kw)
    return F.l1_loss(F.conv2d(x, k, groups=groups), target)


def align_exposure(a, rho):
    """losses.py:195-203."""
    if not torch.is_tensor(rho):
        rho = torch.tensor(rho, dtype=a.dtype, device=a.device)
    if rho.dim() == 0:
        rho = rho.view(1)
    if rho.dim() == 1:
        rho = rho.view(-1, 1, 1, 1)
    return (a * rho).clamp(0.0, 1.0)


def phys_loss_srgb(bhat, a, rho, kernel):
    """losses.py:217-220 — L1 mean of CrosstalkPSF(bhat) - align(a, rho)."""
    return F.l1_loss(crosstalk_forward(bhat, kernel), align_exposure(a, rho))


def embed_dense(kernel: torch.Tensor, channels: int = 3) -> torch.Tensor:
    """Block-diagonal dense [C,C,kh,kw] embedding of a mono/depthwise kernel — the only form
    the reference's phys_cons_* accepts."""
    kh, kw = kernel.shape[-2:]
    k = kernel.expand(channels, 1, kh, kw) if kernel.shape[0] == 1 else kernel
    dense = torch.zeros(channels, channels, kh, kw, dtype=kernel.dtype)
    for c in range(channels):
        dense[c, c] = k[c, 0]
    return dense
