"""Training losses — drop-in for the hot-path part of the reference `NewBP_model/losses.py`.

Implemented on sm_100a kernels (fused forward + analytic backward):
    PhysicsConsistencyLoss        losses.py:158-192   | K * Bhat_raw (replicate pad) - clamp(A*rho) |_1
    PhysicalConsistencyLossSRGB   losses.py:206-220   | CrosstalkPSF(Bhat) - clamp(A*rho, 0, 1) |_1
    align_exposure_srgb           losses.py:195-203
    DeltaE00Loss                  losses.py:92-143    mean CIEDE2000 (loss variant) on clamp(.,0,1) sRGB
    SSIMLoss                      losses.py:146-155   kornia-policy DSSIM, no kornia dependency
    HybridLossPlus                losses.py:223-372   L1 + dE00 + DSSIM + physics (+ injected extras)

Out of the kernel scope (stock PyTorch in the reference too): the VGG19
`PerceptualLoss` and LPIPS.  `PerceptualLoss` below is the same stock torchvision wrapper (losses.py:32-69), and
`HybridLoss` / `HybridLossPlus` build it by default exactly as the reference does (losses.py:84, 250) so that a
drop-in swap keeps the training objective.  When the IMAGENET1K_V1 weights can be neither found in the torch hub
cache nor downloaded, construction emits ONE `RuntimeWarning` naming the missing term (never a silent 0); an
explicit `perceptual_module=` (any (gen, tgt) -> scalar module) overrides the default, `perceptual_module=False`
switches the term off on purpose.
"""
from __future__ import annotations

import warnings
from typing import Dict, Optional, Union

import torch
import torch.nn as nn
from torch import Tensor

from .. import _functional as Fn
from .. import _native as nat
from .newbp_layer import CrosstalkPSF


def _as_rho(expo_ratio, n: int, dev: torch.device):
    """[N] / [N,1,1,1] / 0-d / python scalar -> (tensor|None, scalar, strides)."""
    if not torch.is_tensor(expo_ratio):
        return None, float(expo_ratio), (0, 0, 0, 0)
    r = nat.stage(expo_ratio.detach(), dev, torch.float32)
    if r.numel() == 1:
        return r.reshape(1), 0.0, (0, 0, 0, 0)
    if r.numel() == n and (r.ndim == 1 or tuple(r.shape[1:]) == (1, 1, 1)):
        return r.reshape(n), 0.0, (1, 0, 0, 0)
    raise ValueError(f"expo_ratio of shape {tuple(expo_ratio.shape)} is not a per-sample ratio for batch {n}")


class PhysicsConsistencyLoss(nn.Module):
    """|K*Bhat_raw - align(A_raw)|_1 with a depthwise K ([C,1,kh,kw] or shared [1,1,kh,kw]).

    Border handling is replicate padding, `align(A) = clamp(A * expo_ratio, 0, 1)` when
    `clamp_align`; the value is the global mean over N*C*H*W (losses.py:173-192).  Forward and the
    gradient w.r.t. `Bhat_raw` are each a single fused kernel (csrc/phys.cu).
    """

    def __init__(self, K_kernel: Tensor, device: str = "cuda", clamp_align: bool = True):
        super().__init__()
        if K_kernel is None:
            raise ValueError("K_kernel must be provided for PhysicsConsistencyLoss")
        self.register_buffer("K", K_kernel.to(device))
        self.K: Tensor
        self.clamp_align = clamp_align
        self._prepared = None

    def _psf(self, c: int, dev: torch.device) -> Fn.PreparedPSF:
        k = self.K
        key = (id(k), k._version, tuple(k.shape), c, str(dev))
        if self._prepared is None or self._prepared[0] != key:
            kh, kw = k.shape[-2:]
            kk = k.detach().to(torch.float32)
            if kk.shape[0] == 1 and c > 1:
                kk = kk.expand(c, 1, kh, kw)
            if kk.shape[0] == c and kk.shape[1] == 1:
                layout = Fn.PSF_DEPTHWISE
            else:  # groups == 1 path of the reference: a [Co,1,kh,kw] bank is expanded along C_in
                if kk.shape[1] == 1 and c != 1:
                    kk = kk.expand(kk.shape[0], c, kh, kw)
                layout = Fn.PSF_DENSE
            self._prepared = (key, Fn.PreparedPSF(nat.stage(kk, dev), layout, kh, kw, False))
        return self._prepared[1]

    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(self, Bhat_raw: Tensor, A_raw: Tensor, expo_ratio: Tensor) -> Tensor:
        dev = nat.cuda_device_for(Bhat_raw, A_raw)
        home = Bhat_raw.device
        x = nat.stage(Bhat_raw, dev, torch.float32)
        a = nat.stage(A_raw.detach(), dev, torch.float32)
        n, c = x.shape[:2]
        psf = self._psf(c, dev)
        rho_t, rho_s, rho_st = _as_rho(expo_ratio, n, dev)
        cfg = Fn.PhysCfg(nat.PAD["replicate"], False, False, 1e-12, 1, False, bool(self.clamp_align))
        vec = Fn.PhysProjector.apply(x, a, psf, rho_t, rho_s, rho_st, cfg)[0]
        out = vec[n]
        return out if home == dev else out.to(home)


def align_exposure_srgb(a_srgb: Tensor, ratio: Tensor | float) -> Tensor:
    """clamp(a_srgb * ratio, 0, 1) with ratio a scalar, 0-d, [N] or [N,1,1,1] tensor (losses.py:195-203)."""
    if not torch.is_tensor(ratio):
        ratio = torch.tensor(ratio, dtype=a_srgb.dtype, device=a_srgb.device)
    if ratio.dim() == 0:
        ratio = ratio.view(1)
    if ratio.dim() == 1:
        ratio = ratio.view(-1, 1, 1, 1)
    return (a_srgb * ratio).clamp(0.0, 1.0)


class PhysicalConsistencyLossSRGB(nn.Module):
    """L_phys = || PSF(Bhat_srgb) - Align(A_srgb; ratio) ||_1 (global mean), PSF used only here.

    With a `CrosstalkPSF` module the whole term (depthwise zero-padded 3x3, exposure alignment,
    clamp, |.|, mean) is one fused kernel forward and one backward.  Any other module — or a
    CrosstalkPSF that carries forward hooks (the reference's call-count test) — is called as is and
    only the residual / reduction is fused (1x1 identity stencil).
    """

    def __init__(self, psf_module: nn.Module):
        super().__init__()
        self.psf = psf_module
        self._ident = None

    def _fused_psf(self, dev) -> Optional[Fn.PreparedPSF]:
        m = self.psf
        if not isinstance(m, CrosstalkPSF) or m._forward_hooks or m._forward_pre_hooks:
            return None
        k = m.kernel
        if m.mode == "mono":
            assert k.shape == (1, 1, 3, 3), "mono mode expects kernels of shape [1,1,3,3]"
            k = k.expand(3, 1, 3, 3)
        else:
            assert k.shape == (3, 1, 3, 3), "rgb mode expects kernels of shape [3,1,3,3]"
        return Fn.PreparedPSF(nat.stage(k.detach(), dev, torch.float32), Fn.PSF_DEPTHWISE, 3, 3, False)

    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(self, bhat_srgb: Tensor, a_srgb: Tensor, ratio: Tensor | float) -> Tensor:
        dev = nat.cuda_device_for(bhat_srgb, a_srgb)
        home = bhat_srgb.device
        a = nat.stage(a_srgb.detach(), dev, torch.float32)
        n = a.shape[0]
        rho_t, rho_s, rho_st = _as_rho(ratio, n, dev)
        psf = self._fused_psf(dev)
        if psf is not None:
            assert bhat_srgb.shape[1] == 3, "CrosstalkPSF expects sRGB inputs (3 channels)."
            x = nat.stage(bhat_srgb, dev, torch.float32)
        else:
            x = nat.stage(self.psf(bhat_srgb), dev, torch.float32)
            c = x.shape[1]
            psf = Fn.PreparedPSF(torch.ones(c, 1, 1, 1, dtype=torch.float32, device=dev), Fn.PSF_DEPTHWISE, 1, 1, False)
        cfg = Fn.PhysCfg(nat.PAD["zeros"], False, False, 1e-12, 1, False, True)
        out = Fn.PhysProjector.apply(x, a, psf, rho_t, rho_s, rho_st, cfg)[0][n]
        return out if home == dev else out.to(home)


def _resolve_device(device: Union[str, torch.device]) -> torch.device:
    """losses.py:20-29 ('auto' -> cuda when available)."""
    if isinstance(device, torch.device):
        return device
    if isinstance(device, str):
        if device == "auto":
            return torch.device("cuda" if torch.cuda.is_available() else "cpu")
        return torch.device(device)
    raise TypeError(f"Unsupported device spec: {device!r}")


def _pair_on_device(gen: Tensor, tgt: Tensor):
    if gen.shape != tgt.shape or gen.ndim != 4:
        raise ValueError(f"expected two [N,C,H,W] tensors of equal shape, got {tuple(gen.shape)} vs {tuple(tgt.shape)}")
    dev = nat.cuda_device_for(gen, tgt)
    return nat.stage(gen, dev, torch.float32), nat.stage(tgt.detach(), dev, torch.float32), gen.device, dev


class DeltaE00Loss(nn.Module):
    """Mean CIEDE2000 (the reference's *loss* formula: eps inside every sqrt, wrapped hue, sin(d_theta)) between
    two sRGB [0,1] images; one fused kernel yields the value and the gradient w.r.t. `gen_srgb01`
    (losses.py:92-143).  Identical images give sqrt(eps) = 1e-3, not 0, exactly as the reference."""

    def __init__(self, eps: float = 1e-6):
        super().__init__()
        self.eps = eps

    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(self, gen_srgb01: Tensor, tgt_srgb01: Tensor) -> Tensor:
        x, y, home, dev = _pair_on_device(gen_srgb01, tgt_srgb01)
        if x.shape[1] != 3:
            raise ValueError(f"DeltaE00Loss expects 3-channel sRGB inputs, got {x.shape[1]} channels")
        out = Fn.DeltaE00LossFn.apply(x, y, float(self.eps))
        return out if home == dev else out.to(home)


class SSIMLoss(nn.Module):
    """DSSIM = reduction(clamp((1 - SSIM) / 2, 0, 1)) with kornia 0.6.12's conventions (11-tap sigma-1.5
    Gaussian, reflect border, C1 = (0.01 max_val)^2, C2 = (0.03 max_val)^2, eps 1e-12); inputs are clamped
    to [0,1] (losses.py:146-155).  Forward is one separable-filter kernel, backward one adjoint-filter
    kernel (csrc/ssim.cu)."""

    def __init__(self, window_size: int = 11, max_val: float = 1.0, reduction: str = "mean"):
        super().__init__()
        if window_size <= 0 or window_size % 2 == 0:
            raise ValueError(f"window_size must be a positive odd integer; received {window_size}.")
        if window_size > nat.SSIM_MAX_K:
            raise NotImplementedError(f"the CUDA SSIM kernels take odd windows up to {nat.SSIM_MAX_K}")
        if reduction not in ("mean", "sum", "none"):
            raise ValueError(f"Invalid reduction option: {reduction}")
        self.window_size, self.max_val, self.reduction = int(window_size), float(max_val), reduction

    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(self, gen_srgb01: Tensor, tgt_srgb01: Tensor) -> Tensor:
        x, y, home, dev = _pair_on_device(gen_srgb01, tgt_srgb01)
        p = self.window_size // 2
        if p >= x.shape[2] or p >= x.shape[3] or self.window_size > x.shape[2] or self.window_size > x.shape[3]:
            raise RuntimeError("Padding size should be less than the corresponding input dimension")
        out = Fn.SSIMLossFn.apply(x, y, self.window_size, self.max_val, self.reduction)
        return out if home == dev else out.to(home)


class PerceptualLoss(nn.Module):
    """VGG19 feature distance on sRGB [0,1] inputs (losses.py:32-69): clamp, ImageNet mean / std normalisation,
    the first 36 layers of torchvision's `vgg19(IMAGENET1K_V1).features`, then MSE (or L1) between the two feature
    maps.  Stock PyTorch — this term is not part of the kernel path; it is here so that `HybridLoss(Plus)` keeps
    the reference's objective.  Raises whatever torchvision raises when the weights cannot be obtained."""

    def __init__(self, device: Union[str, torch.device] = "cuda", use_mse: bool = True, reduction: str = "mean"):
        super().__init__()
        from torchvision import models
        feats = models.vgg19(weights=models.VGG19_Weights.IMAGENET1K_V1).features
        self.vgg = nn.Sequential(*list(feats.children())[:36]).eval()
        for p in self.vgg.parameters():
            p.requires_grad = False
        self.register_buffer("mean", torch.tensor([0.485, 0.456, 0.406], dtype=torch.float32).view(1, 3, 1, 1),
                             persistent=False)
        self.register_buffer("std", torch.tensor([0.229, 0.224, 0.225], dtype=torch.float32).view(1, 3, 1, 1),
                             persistent=False)
        self.loss_fn = nn.functional.mse_loss if use_mse else nn.functional.l1_loss
        self.reduction = reduction
        self.to(_resolve_device(device))

    def forward(self, generated_img: Tensor, target_img: Tensor) -> Tensor:
        dev = generated_img.device
        target_img = target_img.to(dev)
        if next(self.vgg.parameters()).device != dev:
            self.vgg = self.vgg.to(dev)
        mean, std = self.mean.to(dev), self.std.to(dev)
        fg = self.vgg(((generated_img.clamp(0, 1) - mean) / std).float())
        ft = self.vgg(((target_img.clamp(0, 1) - mean) / std).float())
        return self.loss_fn(fg, ft, reduction=self.reduction)


_perc_warned = False
_perc_failure: Optional[BaseException] = None   # the attempt is made once per process (it builds a 143 M-parameter net)


def _default_perceptual(device, owner: str, weight_desc: str) -> Optional[nn.Module]:
    """The reference's default VGG term (losses.py:84, 250), or None with ONE RuntimeWarning when its pretrained
    weights are unobtainable (offline box): the caller's objective then lacks the term and the warning says so."""
    global _perc_warned, _perc_failure
    import socket
    old = socket.getdefaulttimeout()
    try:
        if _perc_failure is not None:
            raise _perc_failure
        socket.setdefaulttimeout(10.0)        # an offline box must not hang in the download attempt
        return PerceptualLoss(device=device)
    except Exception as exc:  # URLError / OSError / RuntimeError from torch.hub, ImportError without torchvision
        _perc_failure = exc
        if not _perc_warned:
            _perc_warned = True
            warnings.warn(
                f"{owner}: the VGG19 perceptual term ({weight_desc}) is DISABLED because its IMAGENET1K_V1 weights "
                f"could not be loaded ({type(exc).__name__}: {exc}). The reference always adds this term "
                "(NewBP_model/losses.py:84, 250, 337): the objective differs until the weights are in the torch hub "
                "cache or a module is passed as perceptual_module=. Pass perceptual_module=False to silence this.",
                RuntimeWarning, stacklevel=3)
        return None
    finally:
        socket.setdefaulttimeout(old)


class _L1Mean(nn.Module):
    """nn.L1Loss() (global mean) with its sign/numel gradient from one streaming pass (losses.py:249, 332)."""

    def forward(self, a: Tensor, b: Tensor) -> Tensor:
        x, y, home, dev = _pair_on_device(a, b)
        out = Fn.L1MeanFn.apply(x, y)
        return out if home == dev else out.to(home)


class HybridLoss(nn.Module):
    """Legacy L1 + perceptual loss (losses.py:72-89): `forward(generated, target) -> (total, l1, perceptual)`.
    The L1 term runs on the one-pass value + gradient kernel; the VGG19 perceptual term is the stock
    `PerceptualLoss` built by default as upstream (`perceptual_module=` overrides, `False` disables; a
    RuntimeWarning — never a silent 0 — when the pretrained weights are unobtainable)."""

    def __init__(self, lambda_l1: float = 1.0, lambda_perceptual: float = 0.1, device: str = "cuda", *,
                 perceptual_module: Union[nn.Module, None, bool] = None):
        super().__init__()
        self.device = _resolve_device(device)
        self.lambda_l1 = lambda_l1
        self.lambda_perceptual = lambda_perceptual
        self.l1_loss = _L1Mean()
        if perceptual_module is None:
            perceptual_module = _default_perceptual(self.device, "HybridLoss", f"lambda_perceptual={lambda_perceptual}")
        self.perceptual_loss = perceptual_module if isinstance(perceptual_module, nn.Module) else None

    def forward(self, generated_img: Tensor, target_img: Tensor):
        l1_val = self.l1_loss(generated_img, target_img)
        if self.perceptual_loss is not None:
            perceptual_val = self.perceptual_loss(generated_img, target_img)
        else:
            perceptual_val = torch.zeros((), dtype=l1_val.dtype, device=l1_val.device)
        total = self.lambda_l1 * l1_val + self.lambda_perceptual * perceptual_val
        return total, l1_val, perceptual_val


class HybridLossPlus(nn.Module):
    """Pluggable hybrid loss, same constructor / kwargs-only forward / `(total, logs)` contract as
    losses.py:223-372.  L1, DeltaE00, DSSIM and the physics term run on the sm_100a kernels.  The VGG perceptual
    term is the stock `PerceptualLoss`, built by default exactly as upstream (losses.py:250; `perceptual_module=`
    overrides it, `perceptual_module=False` disables it; when the pretrained weights are unobtainable the
    constructor warns once instead of silently dropping the term).  LPIPS is injected as `lpips_module`
    (any nn.Module mapping (gen, tgt) -> per-sample tensor; upstream builds `lpips.LPIPS` only if `use_lpips`).

    The reference checks every term for finiteness as it goes (5-7 host syncs per step); here the terms are
    checked together after the last one (a single sync), raising the same RuntimeError that names the first
    offending term.  `check_finite=False` removes that sync (CUDA-graph capture).

    With fixed weights the four kernel terms run as ONE autograd node (`Fn.HybridFusedFn`): every gradient kernel
    runs in forward with its weight folded in and accumulates into one buffer per input, so backward is a scaling.
    `fused=False`, `use_uncertainty=True` or a physics module that is not a plain `CrosstalkPSF` select the
    per-term modules (same values, more launches)."""

    def __init__(self, device: str = "cuda", w_l1_raw: float = 1.0, w_perc: float = 0.02, w_lpips: float = 0.0,
                 w_deltaE: float = 0.02, w_ssim: float = 0.05, w_phys: float = 0.10, use_deltaE: bool = True,
                 use_ssim: bool = True, use_lpips: bool = False, use_phys: bool = True, use_uncertainty: bool = False,
                 physics_kernel: Optional[Tensor] = None, physics_psf_module: Optional[nn.Module] = None, *,
                 perceptual_module: Union[nn.Module, None, bool] = None, lpips_module: Optional[nn.Module] = None,
                 check_finite: bool = True, fused: bool = True):
        super().__init__()
        device = _resolve_device(device)
        if perceptual_module is None:
            if use_uncertainty or w_perc != 0:
                perceptual_module = _default_perceptual(device, "HybridLossPlus", f"w_perc={w_perc}")
        if not isinstance(perceptual_module, nn.Module):
            perceptual_module = None
        if use_lpips and lpips_module is None:
            warnings.warn("HybridLossPlus(use_lpips=True) without lpips_module=: the LPIPS term is not built here "
                          "(pretrained weights; upstream losses.py:266-279) and is DISABLED.", RuntimeWarning,
                          stacklevel=2)
        self.register_buffer("_zero", torch.tensor(0.0), persistent=False)
        self._zero: Tensor
        self._zero_cache: Dict[torch.device, Tensor] = {}
        self.l1_raw = _L1Mean()
        self.perc = perceptual_module
        self.deltaE = DeltaE00Loss() if use_deltaE else None
        self.ssim = SSIMLoss() if use_ssim else None
        self.lpips = lpips_module if use_lpips else None
        self.phys = (PhysicsConsistencyLoss(physics_kernel, device=device)
                     if use_phys and physics_kernel is not None else None)
        if use_phys and physics_psf_module is not None:
            self.phys_srgb = PhysicalConsistencyLossSRGB(physics_psf_module.to(device))
        else:
            self.phys_srgb = None
        self.use_uncertainty = use_uncertainty
        self.check_finite = check_finite
        self.fused = fused          # one autograd node for the kernel terms (falls back to per-term modules when it cannot)
        if use_uncertainty:
            self.log_sigma = nn.ParameterDict({k: nn.Parameter(torch.zeros(()))
                                               for k in ("l1", "perc", "lpips", "de", "ssim", "phys")})
        else:
            self.w = dict(l1=w_l1_raw, perc=w_perc, lpips=w_lpips, de=w_deltaE, ssim=w_ssim, phys=w_phys)

    def _zero_on(self, dev: torch.device) -> Tensor:
        """A fresh 0-d zero on `dev` without a host -> device copy per step: the buffer lives wherever the module
        was left (CPU unless `.to()` was called), and a pageable H2D copy is a full stream synchronisation."""
        z = self._zero_cache.get(dev)
        if z is None:
            z = self._zero_cache[dev] = torch.zeros((), dtype=self._zero.dtype, device=dev)
        return z.clone()

    @staticmethod
    def _ensure_finite(name: str, value: Tensor):
        """losses.py:298-306."""
        if not torch.isfinite(value).all():
            finite_vals = value[torch.isfinite(value)]
            stats = ""
            if finite_vals.numel() > 0:
                stats = f" (finite min={finite_vals.min().item():.4e}, max={finite_vals.max().item():.4e})"
            raise RuntimeError(f"HybridLossPlus detected non-finite values in term '{name}'.{stats}")

    def _weighted(self, name: str, val: Optional[Tensor]):
        """losses.py:308-316."""
        if val is None:
            zero = self._zero.clone()
            return zero, zero.detach()
        if self.use_uncertainty:
            s = self.log_sigma[name]
            return (val * torch.exp(-2 * s.to(val.device)) + s.to(val.device)), val.detach()
        return self.w[name] * val, val.detach()

    def _fused_spec(self, dev, Bhat_raw, A_raw, expo_ratio, Bhat_srgb01, A_srgb01):
        """Arguments of Fn.HybridFusedFn for this call, or None when the modular path has to run (learned
        uncertainty weights, a non-CrosstalkPSF / hooked physics module, unexpected shapes)."""
        if self.use_uncertainty or Bhat_raw.ndim != 4 or Bhat_srgb01.ndim != 4 or Bhat_srgb01.shape[1] != 3:
            return None
        # rebuilt every call: `self.w` is read per forward like upstream (losses.py:308-316), so weight schedules
        # that edit the dict take effect, and no per-call tensor is kept on the module
        spec = dict(w={k: float(v) for k, v in self.w.items()}, use_de=self.deltaE is not None,
                    de_eps=self.deltaE.eps if self.deltaE is not None else 1e-6,
                    use_ssim=self.ssim is not None,
                    ssim_window=self.ssim.window_size if self.ssim is not None else 11,
                    ssim_max_val=self.ssim.max_val if self.ssim is not None else 1.0)
        spec["phys"] = None
        if self.phys is not None:
            n, c = Bhat_raw.shape[:2]
            rho_t, rho_s, rho_st = _as_rho(expo_ratio, n, dev)
            cfg = Fn.PhysCfg(nat.PAD["replicate"], False, False, 1e-12, 1, False, bool(self.phys.clamp_align))
            spec["phys"] = dict(srgb=False, a=nat.stage(A_raw.detach(), dev, torch.float32), psf=self.phys._psf(c, dev),
                                rho_t=rho_t, rho_s=rho_s, rho_st=rho_st, cfg=cfg)
        elif self.phys_srgb is not None and A_srgb01 is not None:
            psf = self.phys_srgb._fused_psf(dev)
            if psf is None:
                return None
            rho_t, rho_s, rho_st = _as_rho(expo_ratio, Bhat_srgb01.shape[0], dev)
            cfg = Fn.PhysCfg(nat.PAD["zeros"], False, False, 1e-12, 1, False, True)
            spec["phys"] = dict(srgb=True, a=nat.stage(A_srgb01.detach(), dev, torch.float32), psf=psf, rho_t=rho_t,
                                rho_s=rho_s, rho_st=rho_st, cfg=cfg)
        return spec

    def forward(self, *, Bhat_raw: Tensor, B_raw: Tensor, A_raw: Tensor, expo_ratio: Tensor, Bhat_srgb01: Tensor,
                B_srgb01: Tensor, A_srgb01: Optional[Tensor] = None):
        if self.fused:
            dev = nat.cuda_device_for(Bhat_raw, Bhat_srgb01)
            spec = self._fused_spec(dev, Bhat_raw, A_raw, expo_ratio, Bhat_srgb01, A_srgb01)
            if spec is not None:
                return self._forward_fused(dev, spec, Bhat_raw, B_raw, Bhat_srgb01, B_srgb01)
        return self._forward_modular(Bhat_raw=Bhat_raw, B_raw=B_raw, A_raw=A_raw, expo_ratio=expo_ratio,
                                     Bhat_srgb01=Bhat_srgb01, B_srgb01=B_srgb01, A_srgb01=A_srgb01)

    def _forward_fused(self, dev, spec, Bhat_raw, B_raw, Bhat_srgb01, B_srgb01):
        """One autograd node for L1 + dE00 + DSSIM + physics (weights folded into the gradient kernels), the
        injected perceptual / LPIPS modules on top; same `(total, logs)` as the modular path."""
        home = Bhat_raw.device
        xr = nat.stage(Bhat_raw, dev, torch.float32)
        xs = nat.stage(Bhat_srgb01, dev, torch.float32)
        br = nat.stage(B_raw.detach(), dev, torch.float32)
        bs = nat.stage(B_srgb01.detach(), dev, torch.float32)
        if xr.shape != br.shape or xs.shape != bs.shape:
            raise ValueError("HybridLossPlus: prediction and target shapes differ")
        total, terms = Fn.HybridFusedFn.apply(xr, xs, br, bs, spec)
        logs: Dict[str, Tensor] = {"L1_raw": terms[0]}
        checks = [("L1_raw", terms[0])]
        if self.perc is not None:
            lp = self.perc(Bhat_srgb01, B_srgb01)
            checks.append(("Perc", lp))
            lw, logs["Perc"] = self._weighted("perc", lp)
            total = total + lw.to(dev)
        else:
            logs["Perc"] = self._zero_on(home)
        if self.lpips is not None:
            ll = self.lpips(Bhat_srgb01, B_srgb01).mean()
            checks.append(("LPIPS", ll))
            lw, logs["LPIPS"] = self._weighted("lpips", ll)
            total = total + lw.to(dev)
        if spec["use_de"]:
            logs["DeltaE"] = terms[1]
            checks.append(("DeltaE", terms[1]))
        if spec["use_ssim"]:
            logs["SSIM"] = terms[2]
            checks.append(("SSIM", terms[2]))
        if spec["phys"] is not None:
            logs["Phys"] = terms[3]
            checks.append(("Phys_srgb" if spec["phys"]["srgb"] else "Phys_raw", terms[3]))
        if self.check_finite:
            vals = torch.stack([v.detach().reshape(()).float().to(dev) for _, v in checks])
            if not bool(torch.isfinite(vals).all()):                      # the one host sync of the step
                for name, v in checks:
                    self._ensure_finite(name, v.detach())
        if home != dev:
            total = total.to(home)
            logs = {k: v.to(home) for k, v in logs.items()}
        logs["Total"] = total.detach()
        return total, logs

    def _forward_modular(self, *, Bhat_raw: Tensor, B_raw: Tensor, A_raw: Tensor, expo_ratio: Tensor,
                         Bhat_srgb01: Tensor, B_srgb01: Tensor, A_srgb01: Optional[Tensor] = None):
        logs: Dict[str, Tensor] = {}
        terms = []          # (check name, value) in the reference's order
        total = 0.0

        def add(check_name, key, log_name, val):
            nonlocal total
            terms.append((check_name, val))
            lw, logs[log_name] = self._weighted(key, val)
            total = total + lw

        add("L1_raw", "l1", "L1_raw", self.l1_raw(Bhat_raw, B_raw))
        if self.perc is not None:
            add("Perc", "perc", "Perc", self.perc(Bhat_srgb01, B_srgb01))
        else:
            logs["Perc"] = self._zero_on(Bhat_raw.device)
        if self.lpips is not None:
            add("LPIPS", "lpips", "LPIPS", self.lpips(Bhat_srgb01, B_srgb01).mean())
        if self.deltaE is not None:
            add("DeltaE", "de", "DeltaE", self.deltaE(Bhat_srgb01, B_srgb01))
        if self.ssim is not None:
            add("SSIM", "ssim", "SSIM", self.ssim(Bhat_srgb01, B_srgb01))
        if self.phys is not None:
            add("Phys_raw", "phys", "Phys", self.phys(Bhat_raw, A_raw, expo_ratio))
        elif self.phys_srgb is not None and A_srgb01 is not None:
            add("Phys_srgb", "phys", "Phys", self.phys_srgb(Bhat_srgb01, A_srgb01, expo_ratio))
        if self.check_finite:
            vals = torch.stack([v.detach().reshape(()).float() for _, v in terms])
            if not bool(torch.isfinite(vals).all()):                      # the one host sync of the step
                for (name, v) in terms:
                    self._ensure_finite(name, v.detach())
        logs["Total"] = total.detach()
        return total, logs
