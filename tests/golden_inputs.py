"""Seeded input builders shared by tests/golden/make_golden.py (reference side, build container)
and the parity tests (oracle side and CUDA side).  CPU torch RNG only, so both sides see the
same bits.  No reference import here."""
from __future__ import annotations

import json
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def _gen(seed):
    return torch.Generator().manual_seed(seed)


# ------------------------------------------------------------------ PSFs
def dense_from_spec(spec) -> torch.Tensor:
    """Dense [3,3,kh,kw] PSFs — the only form the reference phys_cons_* accepts."""
    kind = spec[0]
    if kind == "delta":
        k = spec[1]
        p = torch.zeros(3, 3, k, k)
        for c in range(3):
            p[c, c, k // 2, k // 2] = 1.0
        return p
    if kind == "blur":
        k = spec[1]
        return torch.ones(3, 3, k, k) / (3 * k * k)
    if kind == "rand":  # random cross-channel, includes negative lobes
        kh, kw, seed = spec[1], spec[2], spec[3]
        p = torch.randn(3, 3, kh, kw, generator=_gen(seed)) * 0.1
        for c in range(3):
            p[c, c, kh // 2, kw // 2] += 1.0
        return p
    if kind == "diag":  # block-diagonal embedding of depthwise kernels [3,1,kh,kw]
        dw = depthwise_from_spec(spec[1])
        kh, kw = dw.shape[-2:]
        p = torch.zeros(3, 3, kh, kw)
        for c in range(3):
            p[c, c] = dw[c, 0]
        return p
    raise ValueError(spec)


def depthwise_from_spec(name) -> torch.Tensor:
    p2 = torch.tensor([[0.01, 0.02, 0.01], [0.02, 0.88, 0.02], [0.01, 0.02, 0.01]])
    if name == "P2":
        return p2.view(1, 1, 3, 3).expand(3, 1, 3, 3).contiguous()
    if name == "B2":
        r = torch.tensor([[0.0117, 0.0233, 0.0117], [0.0233, 0.86, 0.0233], [0.0117, 0.0233, 0.0117]])
        b = torch.tensor([[0.0083, 0.0167, 0.0083], [0.0167, 0.90, 0.0167], [0.0083, 0.0167, 0.0083]])
        return torch.stack([r, p2, b]).view(3, 1, 3, 3)
    if name == "R5x3":  # non-square, asymmetric
        return torch.rand(3, 1, 5, 3, generator=_gen(991))
    raise ValueError(name)


# ------------------------------------------------------------------ phys_cons cases
def _case(**kw):
    base = dict(srgb=False, clamp01=False, reduction="mean", padding="reflect", crop="valid",
                robust="none", eps=1e-12, rho="scalar", psf=("diag", "P2"), shape=(2, 3, 24, 20), seed=11)
    base.update(kw)
    return base


PHYS_CASES = [
    _case(),
    _case(srgb=True, clamp01=True, rho="scalar1"),
    _case(srgb=True, clamp01=True, robust="charbonnier", eps=1e-3, rho="scalar1", seed=12),
    _case(padding="replicate", crop="same", rho="batch", psf=("diag", "B2"), seed=13),
    _case(padding="zeros", crop="same", rho="batch", psf=("diag", "B2"), reduction="none", seed=14),
    _case(padding="zeros", crop="valid", rho="spatial", psf=("blur", 5), reduction="sum", seed=15),
    _case(padding="reflect", crop="same", rho="full", psf=("rand", 3, 5, 5), seed=16),
    _case(padding="replicate", crop="valid", rho="batch", psf=("rand", 5, 3, 6), robust="charbonnier",
          eps=1e-6, seed=17, shape=(3, 3, 17, 33)),
    _case(srgb=True, clamp01=True, padding="reflect", crop="same", rho="batch", psf=("diag", "R5x3"),
          seed=18, shape=(2, 3, 16, 40)),
    _case(padding="reflect", crop="valid", rho="scalar", psf=("blur", 9), seed=19, shape=(1, 3, 48, 48)),
    _case(padding="reflect", crop="same", rho="scalar", psf=("delta", 3), seed=20, normalize_psf=False),
]


def phys_inputs(case):
    n, c, h, w = case["shape"]
    g = _gen(case["seed"])
    pred = torch.rand(n, c, h, w, generator=g)
    obs = torch.rand(n, c, h, w, generator=g)
    kind = case["rho"]
    if kind == "scalar":
        rho = 0.7
    elif kind == "scalar1":
        rho = 1.0
    elif kind == "batch":
        rho = torch.tensor([0.8, 1.2, 1.7][:n])
    elif kind == "spatial":
        rho = 0.5 + torch.rand(n, 1, h, w, generator=g)
    elif kind == "full":
        rho = 0.5 + torch.rand(n, c, h, w, generator=g)
    else:
        raise ValueError(kind)
    return pred, obs, rho


def loss_inputs(mode):
    g = _gen(31 if mode == "mono" else 32)
    bhat = torch.rand(2, 3, 31, 29, generator=g)
    a = torch.rand(2, 3, 31, 29, generator=g) * 0.01
    rho = torch.tensor([100.0, 250.0]).view(2, 1, 1, 1)
    return bhat, a, rho


# ------------------------------------------------------------------ PSNR / SSIM inputs
SSIM_CASES = [
    dict(),
    dict(padding="replicate"),
    dict(padding="circular"),
    dict(padding="constant"),
    dict(kernel_size=7, sigma=1.0),
    dict(kernel_size=3, sigma=0.8, padding="replicate"),
    dict(gaussian=False, kernel_size=5),
    dict(data_range=2.0, k1=0.02, k2=0.05),
]


def quality_inputs():
    g = _gen(41)
    tgt = torch.rand(2, 3, 40, 36, generator=g)
    pred = (tgt + 0.05 * torch.randn(2, 3, 40, 36, generator=g)).clamp(0, 1)
    return pred, tgt


# ------------------------------------------------------------------ Lab pairs
# flat indices of the exact-opposite-hue pairs (|h1-h2| == pi): the CIEDE2000 formula is discontinuous
# there, fp32 vs fp64 runs of the REFERENCE differ by up to 42 units -> excluded from parity checks.
LAB_UNSTABLE = slice(32, 36)


def lab_inputs(json_path=None):
    """18 Sharma pairs (reference standard_tests/data/ciede2000_pairs.json, copied as data into
    tests/golden/ciede2000_pairs.json) followed by 110 random Lab pairs incl. greys and exact
    duplicates; layout [1,3,8,16]."""
    if json_path is None:
        json_path = os.path.join(HERE, "golden", "ciede2000_pairs.json")
    with open(json_path) as fh:
        pairs = json.load(fh)
    g = _gen(51)
    n_extra = 128 - len(pairs)
    l1 = torch.rand(n_extra, generator=g) * 100
    a1 = (torch.rand(n_extra, generator=g) - 0.5) * 200
    b1 = (torch.rand(n_extra, generator=g) - 0.5) * 200
    l2 = (l1 + torch.randn(n_extra, generator=g) * 10).clamp(0, 100)
    a2 = a1 + torch.randn(n_extra, generator=g) * 20
    b2 = b1 + torch.randn(n_extra, generator=g) * 20
    a1[:6] = 0.0
    b1[:6] = 0.0          # achromatic first colour
    a2[3:9] = 0.0
    b2[3:9] = 0.0         # achromatic second colour (overlaps -> both grey)
    l2[10:14], a2[10:14], b2[10:14] = l1[10:14], a1[10:14], b1[10:14]  # identical pairs
    a2[14:18], b2[14:18] = -a1[14:18], -b1[14:18]                      # opposite hues (|dh| = pi)
    first = torch.tensor([[p["L1"], p["a1"], p["b1"]] for p in pairs], dtype=torch.float32)
    second = torch.tensor([[p["L2"], p["a2"], p["b2"]] for p in pairs], dtype=torch.float32)
    lab1 = torch.cat([first, torch.stack([l1, a1, b1], 1)]).t().reshape(1, 3, 8, 16).contiguous()
    lab2 = torch.cat([second, torch.stack([l2, a2, b2], 1)]).t().reshape(1, 3, 8, 16).contiguous()
    return lab1, lab2


def l_plane_inputs():
    """L planes whose Sobel sums are exact in fp32 (multiples of 1/8 in [0,100]) -> the edge count
    is order-of-summation independent and must match bit-exactly."""
    g = _gen(61)
    lp = torch.randint(0, 801, (3, 1, 37, 53), generator=g).to(torch.float32) / 8.0
    lp[2] = 42.0  # constant image: every pixel equals the threshold -> all selected except borders differ
    return lp


# ------------------------------------------------------------------ BasicSR pixel-loss inputs
PIXEL_LOSS_CASES = [
    dict(name="l1_mean", loss="l1", reduction="mean", weight=None),
    dict(name="l1_sum", loss="l1", reduction="sum", weight=None),
    dict(name="l1_w1_mean", loss="l1", reduction="mean", weight=1),
    dict(name="l1_wc_mean", loss="l1", reduction="mean", weight=3),
    dict(name="mse_mean", loss="mse", reduction="mean", weight=None),
    dict(name="mse_wc_sum", loss="mse", reduction="sum", weight=3),
    dict(name="mse_w1_mean", loss="mse", reduction="mean", weight=1),
    dict(name="charb_mean", loss="charbonnier", reduction="mean", weight=None),
    dict(name="charb_eps_mean", loss="charbonnier", reduction="mean", weight=None, eps=1e-6),
    dict(name="charb_w1_sum", loss="charbonnier", reduction="sum", weight=1),
    dict(name="psnr", loss="psnr", toY=False),
    dict(name="psnr_toY", loss="psnr", toY=True),
]


def pixel_loss_inputs(shape=(3, 3, 24, 20)):
    g = _gen(71)
    tgt = torch.rand(*shape, generator=g)
    pred = (tgt + 0.1 * torch.randn(*shape, generator=g)).clamp(0, 1)
    w1 = torch.rand(shape[0], 1, *shape[2:], generator=g)
    wc = torch.rand(*shape, generator=g)
    return pred, tgt, {None: None, 1: w1, 3: wc}


# ------------------------------------------------------------------ decode / normalise step
INGEST_CASES = [
    dict(name="full", h=24, w=40, ratio=100.0, crop=None, seed=301),
    dict(name="crop_aligned", h=40, w=64, ratio=250.0, crop=(8, 12, 32, 32), seed=302),  # vector loads
    dict(name="crop_shifted", h=37, w=50, ratio=100.0, crop=(5, 6, 16, 16), seed=305),   # 2-byte loads, vector stores
    dict(name="crop_odd", h=37, w=50, ratio=300.0, crop=(3, 7, 21, 21), seed=303),       # width % 4 != 0: generic kernel
    dict(name="dark", h=16, w=16, ratio=1.0, crop=None, seed=304),
]


def ingest_inputs(case):
    """Two [H, W, 3] uint16 arrays (numpy): a short exposure concentrated near black (so that ratio x value straddles
    the clip at 1) with the extremes 0 and 65535 present, and a long exposure over the full range."""
    import numpy as np
    g = _gen(case["seed"])
    h, w = case["h"], case["w"]
    short = (torch.rand(h, w, 3, generator=g) ** 3 * (65535.0 * 4.0 / case["ratio"])).clamp(0, 65535).to(torch.int32)
    long = torch.randint(0, 65536, (h, w, 3), generator=g, dtype=torch.int32)
    short[0, 0, 0], short[0, 1, 1], long[0, 0, 2], long[1, 0, 0] = 0, 65535, 65535, 0
    return short.numpy().astype(np.uint16), long.numpy().astype(np.uint16)


def srgb_inputs():
    """[2, 3, 33, 47] linear values: uniform in [-0.1, 1.1] with the OETF knee, 0, 1 and tiny values planted."""
    g = _gen(311)
    x = torch.rand(2, 3, 33, 47, generator=g) * 1.2 - 0.1
    x[0, 0, 0, :8] = torch.tensor([0.0, 1.0, 0.0031308, 0.0031307, 0.0031309, 1e-8, 0.5, 0.18])
    x[1, 1, 1, :8] = torch.logspace(-6, 0, 8)
    return x


# ---- validation tiling (image_restoration_model.py:167-241): crop grids over odd sizes, ratio-defined crops, a single
# tile, a super-resolution scale
TILE_CASES = [
    dict(h=150, w=250, scale=1, val=dict(crop_size_h=64, crop_size_w=48)),
    dict(h=64, w=64, scale=1, val=dict(crop_size_h=64, crop_size_w=64)),
    dict(h=111, w=87, scale=1, val=dict(crop_size_h_ratio=0.5, crop_size_w_ratio=0.4)),
    dict(h=32, w=320, scale=1, val=dict(crop_size_h=32, crop_size_w=100)),
    dict(h=128, w=192, scale=2, val=dict(crop_size_h=51, crop_size_w=65)),
    dict(h=97, w=131, scale=1, val=dict(crop_size_h=48, crop_size_w=48)),
]


def tile_inputs(k):
    """(lq [1, 3, h / scale, w / scale], outs_fn(T, tiles) -> [T, 3, crop_h, crop_w] seeded network outputs)."""
    case = TILE_CASES[k]
    g = torch.Generator().manual_seed(7000 + k)
    s = case["scale"]
    lq = torch.rand(1, 3, case["h"] // s, case["w"] // s, generator=g)

    def outs_fn(t, tiles):
        return torch.rand(t, 3, tiles.shape[-2] * s, tiles.shape[-1] * s, generator=g)

    return lq, outs_fn
