"""TEST INFRASTRUCTURE ONLY — CPU restatement of the decode / normalise step in front of the path (SURVEY.md §8f
rank 3), in numpy / CPU torch exactly as the reference issues it.

* `decode_pair`     `NAFNet_base/basicsr/data/sony_sid_lmdb_dataset.py:196-233` (`__getitem__`: /65535, clip(short *
                    ratio, 0, 1), crop window `:186-187`) + `basicsr/utils/img_util.py:28-34` (HWC -> CHW float tensor)
* `linear_to_srgb`  `metrics/evaluation.py:27-38`;  `norm_to_lpips_range`  `:50-52`;
  `channelwise_sse`  `:55-62`;  `sobel_magnitude`  `:65-86` (manual branch)

Pinned against the unmodified reference functions by `tests/golden/make_golden_ingest.py` ->
`tests/golden/ingest_vectors.npz`.  Only tests/, smoke() and the benchmark's CPU legs may import this module.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

MAX_16BIT_VALUE = 65535.0


def decode_pair(short_u16: np.ndarray, long_u16: np.ndarray, expo_ratio: float, crop=None):
    """[H, W, 3] uint16 x2 -> dict of [3, h, w] float32 tensors with the reference sample's keys."""
    short_raw = short_u16.astype(np.float32) / MAX_16BIT_VALUE
    long_raw = long_u16.astype(np.float32) / MAX_16BIT_VALUE
    aligned = np.clip(short_raw * float(expo_ratio), 0.0, 1.0)
    if crop is not None:
        top, left, h, w = crop
        short_raw, long_raw, aligned = (a[top:top + h, left:left + w, :] for a in (short_raw, long_raw, aligned))

    def chw(a):
        return torch.from_numpy(np.ascontiguousarray(a.transpose(2, 0, 1))).float()

    s, g, sr = chw(aligned), chw(long_raw), chw(short_raw)
    return {"lq": s, "gt": g, "short": s, "long": g, "short_raw": sr, "long_raw": g, "short_obs": s,
            "expo_ratio": torch.full((1, 1, 1), float(expo_ratio), dtype=torch.float32)}


def linear_to_srgb(x: torch.Tensor) -> torch.Tensor:
    x = x.clamp(0.0, 1.0)
    a = 0.055
    threshold = 0.0031308
    low = 12.92 * x
    high = (1 + a) * torch.pow(x.clamp(min=threshold), 1.0 / 2.4) - a
    return torch.where(x <= threshold, low, high)


def norm_to_lpips_range(x: torch.Tensor) -> torch.Tensor:
    return x.mul(2.0).sub(1.0)


def channelwise_sse(pred: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    return ((pred - target) ** 2).sum(dim=(0, 2, 3))


def sobel_magnitude(gray: torch.Tensor) -> torch.Tensor:
    kx = torch.tensor([[1, 0, -1], [2, 0, -2], [1, 0, -1]], dtype=gray.dtype).view(1, 1, 3, 3)
    ky = torch.tensor([[1, 2, 1], [0, 0, 0], [-1, -2, -1]], dtype=gray.dtype).view(1, 1, 3, 3)
    gx = F.conv2d(gray, kx, padding=1)
    gy = F.conv2d(gray, ky, padding=1)
    return torch.sqrt(gx * gx + gy * gy + 1e-12)
