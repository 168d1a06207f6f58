"""Device-side decode / normalise of Sony-SID 16-bit PNG pairs — the arithmetic of the reference dataset's
`__getitem__` (NAFNet_base/basicsr/data/sony_sid_lmdb_dataset.py:196-233) without the host round trip.

The reference does, per pair and on the CPU: ``astype(float32) / 65535`` on both images, ``np.clip(short * ratio, 0,
1)``, an optional (random or centre) crop of all four arrays, HWC -> CHW and four ``torch.from_numpy(...).float()``
tensors, which the trainer then uploads.  Here the decoded uint16 arrays are uploaded as they are (2 bytes per sample
instead of 4 + 4) and `llie_decode_u16` writes the planar float32 tensors in one pass each — bit-exact with the host
arithmetic (tests/test_gpu_ingest.py).  PNG decompression, LMDB and the crop policy stay where they are (host, out
of scope): `crop` here is the window the dataset chose.
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence, Union

import numpy as np
import torch
from torch import Tensor

from . import _functional as Fn
from . import _native as nat

MAX_16BIT_VALUE = 65535.0
__all__ = ["decode_pair", "decode_batch", "MAX_16BIT_VALUE"]


def _as_u16(img, dev: torch.device) -> Tensor:
    if isinstance(img, np.ndarray):
        if img.dtype == np.uint8:  # placeholder images, promoted like sony_sid_lmdb_dataset.py:48-49
            img = img.astype(np.uint16) * 257
        if img.dtype != np.uint16:
            raise TypeError(f"Expected uint16 image, got dtype={img.dtype}.")
        img = torch.from_numpy(np.ascontiguousarray(img))
    if not torch.is_tensor(img) or img.dtype != torch.uint16:
        raise TypeError(f"Expected a uint16 array / tensor, got {getattr(img, 'dtype', type(img))}.")
    if img.dim() == 2:
        img = img.unsqueeze(-1)
    if img.shape[-1] != 3:
        raise ValueError(f"Expected 3-channel image, got shape {tuple(img.shape)}.")
    return img.to(dev, non_blocking=True)


def decode_batch(short_u16, long_u16, expo_ratio: Union[float, Sequence[float], Tensor], *,
                 crop: Optional[Sequence[int]] = None, bgr: bool = False,
                 device: Optional[torch.device] = None) -> Dict[str, Tensor]:
    """[N, H, W, 3] uint16 short / long exposures -> the tensors of the reference sample dict, batched:
    ``lq`` = ``short`` = ``short_obs`` = clip(short / 65535 * ratio, 0, 1), ``gt`` = ``long`` = ``long_raw`` =
    long / 65535, ``short_raw`` = short / 65535, ``expo_ratio`` [N, 1, 1, 1]; all float32 [N, 3, h, w] on the device.
    ``crop`` = (top, left, h, w); ``bgr`` = the arrays are still in cv2's BGR order."""
    dev = torch.device(device) if device is not None else nat.cuda_device_for(short_u16, long_u16)
    s = _as_u16(short_u16, dev)
    g = _as_u16(long_u16, dev)
    if s.dim() == 3:
        s, g = s.unsqueeze(0), g.unsqueeze(0)
    if s.shape != g.shape:
        raise ValueError(f"short / long shapes differ: {tuple(s.shape)} vs {tuple(g.shape)}")
    n = s.shape[0]
    if torch.is_tensor(expo_ratio):
        rho = expo_ratio.reshape(-1).to(dev, torch.float32)
    else:
        rho = torch.as_tensor(np.asarray(expo_ratio, dtype=np.float64).reshape(-1), dtype=torch.float32).to(dev)
    if rho.numel() == 1 and n > 1:
        rho = rho.expand(n).contiguous()
    if crop is not None:
        top, left, ch, cw = (int(v) for v in crop)
        if ch > s.shape[1] or cw > s.shape[2]:
            raise ValueError(f"Patch size {(ch, cw)} exceeds source dimensions {(s.shape[1], s.shape[2])}.")
    short_raw, aligned = Fn.decode_u16(s, rho, crop=crop, swap_rb=bgr)
    long_raw, _ = Fn.decode_u16(g, 1.0, crop=crop, swap_rb=bgr, want_aligned=False)
    return {"lq": aligned, "gt": long_raw, "short": aligned, "long": long_raw, "short_raw": short_raw,
            "long_raw": long_raw, "short_obs": aligned, "expo_ratio": rho.view(n, 1, 1, 1)}


def decode_pair(short_u16, long_u16, expo_ratio: float, *, crop: Optional[Sequence[int]] = None, bgr: bool = False,
                device: Optional[torch.device] = None) -> Dict[str, Tensor]:
    """One pair, un-batched like the reference sample: [3, h, w] tensors and ``expo_ratio`` of shape [1, 1, 1]."""
    out = decode_batch(short_u16, long_u16, float(expo_ratio), crop=crop, bgr=bgr, device=device)
    return {k: (v[0] if k != "expo_ratio" else v.view(1, 1, 1)) for k, v in out.items()}
