"""ctypes binding of the C ABI in include/llie_b200.h (the only way the package reaches the GPU).

There is NO CPU implementation behind these functions: if the shared library is missing, or no
CUDA device is present, every entry point raises — loudly, by design (no fallback path).
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
# LLIE_B200_LIB selects an alternative build of the SAME library (A/B variants made by tools/); default in-tree
LIB_PATH = os.environ.get("LLIE_B200_LIB") or os.path.join(_PKG, "_lib", "libllie_b200.so")

_lib = None
_lock = threading.Lock()


class NativeLibraryError(RuntimeError):
    pass


class PhysDesc(C.Structure):
    _fields_ = [
        ("n", C.c_int32), ("c_in", C.c_int32), ("c_out", C.c_int32), ("h", C.c_int32), ("w", C.c_int32),
        ("kh", C.c_int32), ("kw", C.c_int32), ("psf_layout", C.c_int32), ("pad_mode", C.c_int32),
        ("crop_valid", C.c_int32), ("robust", C.c_int32), ("rho_side", C.c_int32), ("clip_y", C.c_int32),
        ("clip_a", C.c_int32), ("stream_overlap", C.c_int32), ("eps", C.c_float), ("rho_scalar", C.c_float),
        ("rho_sn", C.c_int64), ("rho_sc", C.c_int64), ("rho_sh", C.c_int64), ("rho_sw", C.c_int64),
    ]


class ConvDesc(C.Structure):
    _fields_ = [
        ("n", C.c_int32), ("c_in", C.c_int32), ("c_out", C.c_int32), ("h", C.c_int32), ("w", C.c_int32),
        ("kh", C.c_int32), ("kw", C.c_int32), ("pad_h", C.c_int32), ("pad_w", C.c_int32),
        ("depthwise", C.c_int32), ("dtype", C.c_int32),
    ]


class PsnrDesc(C.Structure):
    _fields_ = [
        ("n", C.c_int32), ("c", C.c_int32), ("plane", C.c_int64), ("dtype", C.c_int32),
        ("cast_first", C.c_int32), ("clamp", C.c_int32), ("clamp_hi", C.c_double),
        ("data_range", C.c_double), ("eps", C.c_double),
    ]


class SsimDesc(C.Structure):
    _fields_ = [(k, C.c_int32) for k in ("n", "c", "h", "w", "ksize", "pad_mode", "luma", "clamp_in", "clamp_var",
                                         "interior", "loss_map")] + [(k, C.c_float) for k in ("c1", "c2", "eps")]


class DeltaEDesc(C.Structure):
    _fields_ = [(k, C.c_int32) for k in ("n", "h", "w", "variant", "input_lab", "clamp_in")] + \
               [(k, C.c_float) for k in ("kL", "kC", "kH", "eps")]


class DecodeDesc(C.Structure):
    _fields_ = [("n", C.c_int32), ("h", C.c_int32), ("w", C.c_int32), ("c", C.c_int32), ("top", C.c_int32),
                ("left", C.c_int32), ("crop_h", C.c_int32), ("crop_w", C.c_int32), ("swap_rb", C.c_int32)]


class EvalDesc(C.Structure):
    _fields_ = [("n", C.c_int32), ("c", C.c_int32), ("hw", C.c_int64), ("ssim_c", C.c_int32),
                ("ssim_count", C.c_int64), ("data_range", C.c_double), ("k", C.c_int32), ("slot_psnr", C.c_int32),
                ("slot_ssim", C.c_int32), ("slot_de_mean", C.c_int32), ("slot_de_p95", C.c_int32),
                ("slot_edge", C.c_int32), ("slot_cpsnr", C.c_int32), ("slot_rgb_psnr", C.c_int32 * 4),
                ("slot_rgb_ssim", C.c_int32 * 4)]


class SelectDesc(C.Structure):
    _fields_ = [("images", C.c_int32), ("per_image", C.c_int64), ("nq", C.c_int32), ("joint", C.c_int32)]


SSIM_MAX_K = 63  # LLIE_SSIM_MAX_K (include/llie_b200.h)
PAD = {"zeros": 0, "constant": 0, "reflect": 1, "replicate": 2, "circular": 3}

_VP = C.c_void_p
_SIGNATURES = {
    "llie_abi_version": (C.c_int, []),
    "llie_last_error": (C.c_char_p, []),
    "llie_workspace_bytes": (C.c_size_t, []),
    "llie_check_device": (C.c_int, []),
    "llie_tune": (C.c_int, [C.c_int, C.c_int]),
    "llie_phys_forward": (C.c_int, [C.POINTER(PhysDesc), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "llie_phys_backward": (C.c_int, [C.POINTER(PhysDesc), _VP, _VP, _VP, _VP, _VP, C.c_int64, C.c_float,
                                     _VP, _VP, _VP]),
    "llie_phys_value_and_grad": (C.c_int, [C.POINTER(PhysDesc), _VP, _VP, _VP, _VP, _VP, C.c_int64, C.c_float, _VP, _VP,
                                          _VP, _VP, _VP, _VP, _VP]),
    "llie_phys_backward_needs_scratch": (C.c_int, [C.POINTER(PhysDesc), _VP, _VP, _VP, _VP]),
    "llie_newbp_forward": (C.c_int, [C.POINTER(ConvDesc), _VP, _VP, _VP, _VP]),
    "llie_newbp_adjoint": (C.c_int, [C.POINTER(ConvDesc), _VP, _VP, _VP, _VP]),
    "llie_psnr_stats": (C.c_int, [C.POINTER(PsnrDesc), _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "llie_psnr_stats_static": (C.c_int, [C.POINTER(PsnrDesc), _VP, _VP, _VP, _VP, _VP, _VP]),
    "llie_ssim_stats": (C.c_int, [C.POINTER(SsimDesc), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "llie_ssim_stats_dual_eligible": (C.c_int, [C.POINTER(SsimDesc), _VP, _VP, _VP]),
    "llie_ssim_stats_dual": (C.c_int, [C.POINTER(SsimDesc), C.c_float, C.c_float, C.c_double, _VP, _VP, _VP, _VP, _VP,
                                       _VP, _VP, _VP, _VP]),
    "llie_deltae": (C.c_int, [C.POINTER(DeltaEDesc), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "llie_deltae_sqerr": (C.c_int, [C.POINTER(DeltaEDesc), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "llie_deltae_sqerr_rgb": (C.c_int, [C.POINTER(DeltaEDesc), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "llie_deltae_loss_grad": (C.c_int, [C.POINTER(DeltaEDesc), _VP, _VP, C.c_float, _VP, _VP, _VP, _VP, _VP]),
    "llie_ssim_loss_forward": (C.c_int, [C.POINTER(SsimDesc), _VP, _VP, _VP, _VP, C.c_float, _VP, _VP, _VP, _VP, _VP,
                                         _VP]),
    "llie_ssim_loss_backward": (C.c_int, [C.POINTER(SsimDesc), _VP, _VP, _VP, _VP, _VP, _VP, C.c_int32, C.c_int32,
                                          _VP, _VP]),
    "llie_l1_grad": (C.c_int, [_VP, _VP, C.c_int64, C.c_float, C.c_int32, _VP, _VP, _VP, _VP, _VP]),
    "llie_pixel_loss_grad": (C.c_int, [C.c_int32, _VP, _VP, _VP, C.c_int32, _VP, C.c_int32, C.c_int32, C.c_int64, C.c_float,
                                       C.c_float, C.c_int32, _VP, _VP, _VP, _VP, _VP]),
    "llie_scale_inplace": (C.c_int, [_VP, C.c_int64, _VP, _VP]),
    "llie_hybrid_combine": (C.c_int, [_VP, C.c_double, _VP, C.c_int32, C.c_double, _VP, C.c_int32, C.c_double, _VP,
                                      C.POINTER(C.c_float), _VP, _VP, _VP]),
    "llie_tile_merge": (C.c_int, [_VP, _VP, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _VP, _VP]),
    "llie_decode_u16": (C.c_int, [C.POINTER(DecodeDesc), _VP, _VP, C.c_float, _VP, _VP, _VP]),
    "llie_linear_to_srgb": (C.c_int, [_VP, C.c_int64, _VP, _VP, _VP]),
    "llie_eval_accumulate": (C.c_int, [C.POINTER(EvalDesc), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP,
                                       _VP, _VP, _VP, _VP, _VP]),
    "llie_luma_bt601": (C.c_int, [_VP, C.c_int32, C.c_int64, _VP, _VP]),
    "llie_selftest_div_const": (C.c_int, [C.c_int32, C.c_uint32, C.c_uint64, _VP, _VP]),
    "llie_selftest_srgb_to_linear": (C.c_int, [_VP, _VP, C.c_int64, _VP]),
    "llie_sobel_magnitude": (C.c_int, [_VP, _VP, C.c_int64, C.c_int32, C.c_int32, _VP]),
    "llie_select_workspace_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "llie_select_quantiles": (C.c_int, [C.POINTER(SelectDesc), _VP, _VP, _VP, C.POINTER(C.c_double), _VP, _VP, _VP,
                                        _VP, _VP]),
}


def declared_symbols() -> list[str]:
    return sorted(_SIGNATURES)


def lib():
    """Load (once) the sm_100a kernel library; raise if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise NativeLibraryError(
                    f"{LIB_PATH} is missing: the CUDA kernel library has not been built. "
                    "Run `python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc). "
                    "There is no CPU fallback.")
            handle = C.CDLL(LIB_PATH)
            for name, (res, args) in _SIGNATURES.items():
                fn = getattr(handle, name)  # AttributeError if the ABI drifted
                fn.restype = res
                fn.argtypes = args
            _lib = handle
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().llie_last_error().decode("utf-8", "replace")
        raise NativeLibraryError(f"{what} failed (code {rc}): {msg}")


def cuda_device_for(*tensors) -> torch.device:
    """Device the kernels will run on: the tensors' CUDA device, or the current one for host
    tensors.  Raises when CUDA is unavailable (no fallback)."""
    for t in tensors:
        if torch.is_tensor(t) and t.is_cuda:
            return t.device
    if not torch.cuda.is_available():
        raise NativeLibraryError(
            "lowlight_image_enhancement_b200 needs a CUDA (sm_100a) device; none is available and "
            "there is no CPU fallback.")
    return torch.device("cuda", torch.cuda.current_device())


def stage(t: torch.Tensor, dev: torch.device, dtype=None) -> torch.Tensor:
    """Contiguous tensor on `dev` (host tensors are uploaded asynchronously on the current stream)."""
    if not t.is_cuda or t.device != dev:
        t = t.to(dev, non_blocking=True)
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


_ws_cache: dict = {}
_checked_devices: set = set()


def ensure_device(dev: torch.device) -> None:
    """Raise unless `dev` is a compute-capability 10.x device (checked once per device)."""
    if dev.index not in _checked_devices:
        with torch.cuda.device(dev):
            check(lib().llie_check_device(), "llie_check_device")
        _checked_devices.add(dev.index)


def workspace(dev: torch.device) -> torch.Tensor:
    """Zero-initialised, self-resetting scratch buffer per (device, stream)."""
    stream = torch.cuda.current_stream(dev)
    key = (dev.index, stream.cuda_stream)
    ws = _ws_cache.get(key)
    if ws is None:
        ensure_device(dev)
        ws = torch.zeros(lib().llie_workspace_bytes(), dtype=torch.uint8, device=dev)
        _ws_cache[key] = ws
    return ws


def ptr(t) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def stream_ptr(dev: torch.device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


# launch counter: bench.py reports how many of OUR kernels ran in the timed region
launches = 0


def count_launch(k: int = 1) -> None:
    global launches
    launches += k
