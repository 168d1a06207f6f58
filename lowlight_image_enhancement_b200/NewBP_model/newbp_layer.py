"""NewBP operator and crosstalk PSF — drop-in for the reference `NewBP_model/newbp_layer.py`.

* `NewBPFunction.apply(input, kernel, padding, groups)`  forward = zero-padded cross-correlation,
  backward = the explicit conv_transpose2d adjoint (newbp_layer.py:7-21); both are CUDA kernels of
  csrc/conv.cu (float32 / float64), no gradient for kernel / padding / groups.
* `CrosstalkPSF(mode, kernels)`  fixed, energy-normalised 3x3 PSF kept as the persistent buffer
  `kernel` (so checkpoints round-trip, newbp_layer.py:101); `forward` is the depthwise groups=3,
  padding=1 application used ONLY in the loss graph (Scenario B).
* `build_psf_kernels(mode, kernel_spec)`  canonical P2 (mono) / B2 (rgb) weights.
* `NewBPLayer` is kept for API compatibility; like the reference it refuses to run unless
  `deprecated=False`.
"""
from __future__ import annotations

import torch
from torch.autograd.function import once_differentiable
import torch.nn as nn

from .. import _functional as Fn
from .. import _native as nat

_P2 = ((0.0100, 0.0200, 0.0100), (0.0200, 0.8800, 0.0200), (0.0100, 0.0200, 0.0100))
_B2_RED = ((0.0117, 0.0233, 0.0117), (0.0233, 0.8600, 0.0233), (0.0117, 0.0233, 0.0117))
_B2_BLUE = ((0.0083, 0.0167, 0.0083), (0.0167, 0.9000, 0.0167), (0.0083, 0.0167, 0.0083))


def _k(rows) -> torch.Tensor:
    return torch.tensor(rows, dtype=torch.float32).view(1, 1, 3, 3)


class NewBPFunction(torch.autograd.Function):
    """y = conv2d(x, K, padding, groups);  dL/dx = conv_transpose2d(dL/dy, K, padding, groups)."""

    @staticmethod
    def forward(ctx, input, kernel, padding, groups):
        # dtype contract of F.conv2d under autocast (newbp_layer.py:13 runs inside the caller's autocast region): CUDA
        # operands are rounded to the autocast dtype, products are accumulated in fp32, the result is returned in the
        # autocast dtype.  The kernel below computes in fp32 on the rounded operands, i.e. the same arithmetic.
        lowp = None
        ctx.in_dtype = input.dtype
        if input.is_cuda and input.is_floating_point() and torch.is_autocast_enabled("cuda") \
                and input.dtype != torch.float64:
            lowp = torch.get_autocast_dtype("cuda")
            input = input.to(lowp)
            kernel = kernel.to(lowp)
        ctx.lowp = lowp
        if input.dtype not in (torch.float32, torch.float64):
            input = input.float()
        dev = nat.cuda_device_for(input, kernel)
        ctx.home = input.device
        x = nat.stage(input.detach(), dev)
        k = nat.stage(kernel.detach(), dev, x.dtype)
        if x.ndim != 4 or k.ndim != 4:
            raise RuntimeError("NewBPFunction expects a 4-D input [N,C,H,W] and a 4-D kernel")
        ctx.save_for_backward(k)
        ctx.padding = padding
        ctx.groups = groups
        ctx.x_shape = tuple(x.shape)
        y = Fn.newbp_conv(x, k, padding, groups)
        if lowp is not None:
            y = y.to(lowp)
        return y if ctx.home == dev else y.to(ctx.home)

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_outputs):
        (k,) = ctx.saved_tensors
        gy = nat.stage(grad_outputs, k.device, k.dtype)
        gx = Fn.newbp_adjoint(gy, k, ctx.padding, ctx.groups, ctx.x_shape)
        if ctx.lowp is not None:
            gx = gx.to(ctx.lowp).to(ctx.in_dtype)   # rounded like the half conv_transpose2d, in the caller's dtype
        return (gx if ctx.home == k.device else gx.to(ctx.home)), None, None, None


class NewBPLayer(nn.Module):
    """Input-side NewBP layer (deprecated upstream in favour of the loss-side CrosstalkPSF)."""

    def __init__(self, in_channels=3, kernel_type="panchromatic", kernel_spec="P2", *, deprecated: bool = True):
        super().__init__()
        if kernel_spec not in {"P2", "B2"}:
            raise ValueError(f"Unsupported kernel_spec '{kernel_spec}'. Expected 'P2' or 'B2'.")
        if kernel_type not in {"panchromatic", "rgb"}:
            raise ValueError(f"Unsupported kernel_type '{kernel_type}'. Expected 'panchromatic' or 'rgb'.")
        if kernel_type == "rgb" and in_channels != 3:
            raise ValueError("kernel_type 'rgb' requires in_channels to be 3.")
        if kernel_type == "panchromatic" and kernel_spec != "P2":
            raise ValueError("kernel_type 'panchromatic' requires kernel_spec 'P2'.")
        if kernel_type == "rgb" and kernel_spec != "B2":
            raise ValueError("kernel_type 'rgb' requires kernel_spec 'B2'.")
        self.in_channels = in_channels
        self.kernel_type = kernel_type
        self.kernel_spec = kernel_spec
        self.deprecated = deprecated
        bank = (build_psf_kernels("mono", "P2").repeat(in_channels, 1, 1, 1)
                if kernel_type == "panchromatic" else build_psf_kernels("rgb", "B2"))
        self.kernel = nn.Parameter(bank, requires_grad=False)

    def forward(self, x):
        if self.deprecated:
            raise RuntimeError("Deprecated: use CrosstalkPSF in loss path (Scenario B).")
        return NewBPFunction.apply(x, self.kernel, (self.kernel.shape[-1] - 1) // 2, self.in_channels)


class CrosstalkPSF(nn.Module):
    """Fixed PSF for the output-side consistency loss.

    mode='mono': `kernels` [1,1,3,3] shared by R, G, B;  mode='rgb': [3,1,3,3], one kernel per channel.
    Each kernel is divided by max(sum, 1e-12) at construction (energy preservation).
    """

    def __init__(self, mode: str, kernels: torch.Tensor):
        super().__init__()
        assert mode in {"mono", "rgb"}
        self.mode = mode
        with torch.no_grad():
            k = kernels.clone()
            total = k.reshape(k.shape[0], -1).sum(dim=1, keepdim=True).clamp_min(1e-12)
            k = k / total.view(-1, 1, 1, 1)
        self.register_buffer("kernel", k, persistent=True)
        self.kernel: torch.Tensor

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        assert x.shape[1] == 3, "CrosstalkPSF expects sRGB inputs (3 channels)."
        k = self.kernel
        if self.mode == "mono":
            assert k.shape == (1, 1, 3, 3), "mono mode expects kernels of shape [1,1,3,3]"
            k = k.expand(3, 1, 3, 3)
        else:
            assert k.shape == (3, 1, 3, 3), "rgb mode expects kernels of shape [3,1,3,3]"
        return NewBPFunction.apply(x, k, 1, 3)


def build_psf_kernels(mode: str, kernel_spec: str = "P2") -> torch.Tensor:
    """Canonical kernels: mono/P2 -> [1,1,3,3]; rgb/B2 -> [3,1,3,3] (R, G=P2, B)."""
    if mode not in {"mono", "rgb"}:
        raise ValueError("mode must be 'mono' or 'rgb'")
    if mode == "mono":
        if kernel_spec != "P2":
            raise ValueError("mono mode expects kernel_spec 'P2'")
        return _k(_P2)
    if kernel_spec != "B2":
        raise ValueError("rgb mode expects kernel_spec 'B2'")
    return torch.cat((_k(_B2_RED), _k(_P2), _k(_B2_BLUE)), dim=0)


def create_crosstalk_psf(psf_mode: str = "mono", kernel_spec: str = "P2") -> CrosstalkPSF:
    """Same contract as the reference NewBP_model/newbp_net_arch.py:88-99."""
    return CrosstalkPSF(psf_mode, build_psf_kernels(psf_mode, kernel_spec))
