"""Decode / normalise step in front of the path.

CPU: the oracle restatement vs the golden vectors produced by the reference's own `__getitem__` / evaluation helpers.
GPU: `llie_decode_u16` (bit-exact: integer samples, one IEEE division, one multiply, a clip) and
`llie_linear_to_srgb` (1e-5 relative, BASELINE's fp32 tolerance; measured ~3e-7) through the Python mirrors, vs the
golden vectors, vs the oracle on larger seeded inputs, and through size-independent properties at full SID size."""
from __future__ import annotations

import os

import numpy as np
import pytest
import torch

from oracle import ingest as O
from tests.golden_inputs import INGEST_CASES, ingest_inputs, srgb_inputs

GOLD = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "ingest_vectors.npz")))
KEYS = ("lq", "gt", "short_raw", "long_raw", "short_obs", "expo_ratio")
RTOL = 1e-5


# ------------------------------------------------------------------ CPU: oracle vs reference goldens
@pytest.mark.parametrize("case", INGEST_CASES, ids=[c["name"] for c in INGEST_CASES])
def test_oracle_decode_matches_reference_bits(case):
    short, long = ingest_inputs(case)
    out = O.decode_pair(short, long, case["ratio"], case["crop"])
    for k in KEYS:
        assert np.array_equal(out[k].numpy(), GOLD[f"{case['name']}_{k}"]), k
    assert out["lq"].max() == 1.0 or case["ratio"] == 1.0  # the clip at 1 is exercised


def test_oracle_srgb_helpers_match_reference():
    x = srgb_inputs()
    s = O.linear_to_srgb(x)
    np.testing.assert_allclose(s.numpy(), GOLD["srgb_out"], rtol=1e-6, atol=0)
    np.testing.assert_allclose(O.norm_to_lpips_range(s).numpy(), GOLD["lpips_out"], rtol=0, atol=2e-7)
    y = (x + 0.05 * torch.randn(x.shape, generator=torch.Generator().manual_seed(312))).clamp(0, 1)
    np.testing.assert_allclose(O.channelwise_sse(x.clamp(0, 1), y).numpy(), GOLD["sse_out"], rtol=1e-6)
    np.testing.assert_allclose(O.sobel_magnitude(s.mean(dim=1, keepdim=True)).numpy(), GOLD["sobel_out"], rtol=1e-5,
                               atol=1e-7)


def test_oracle_srgb_known_values():
    x = torch.tensor([0.0, 1.0, 0.0031308, 0.5, -3.0, 7.0])
    s = O.linear_to_srgb(x).double()
    want = torch.tensor([0.0, 1.0, 12.92 * 0.0031308, 1.055 * 0.5 ** (1 / 2.4) - 0.055, 0.0, 1.0], dtype=torch.float64)
    assert torch.allclose(s, want, rtol=1e-6, atol=1e-7)


# ------------------------------------------------------------------ GPU: kernels vs goldens / oracle
gpu = pytest.mark.gpu


@gpu
@pytest.mark.parametrize("case", INGEST_CASES, ids=[c["name"] for c in INGEST_CASES])
def test_decode_pair_bit_exact_vs_reference(case):
    from lowlight_image_enhancement_b200.sony_sid_decode import decode_pair
    short, long = ingest_inputs(case)
    out = decode_pair(short, long, case["ratio"], crop=case["crop"])
    for k in KEYS:
        got = out[k].cpu().numpy()
        assert got.dtype == np.float32 and got.shape == GOLD[f"{case['name']}_{k}"].shape, k
        assert np.array_equal(got, GOLD[f"{case['name']}_{k}"]), k


@gpu
def test_decode_batch_bgr_and_per_image_ratio():
    from lowlight_image_enhancement_b200.sony_sid_decode import decode_batch
    g = torch.Generator().manual_seed(320)
    n, h, w = 3, 20, 28
    short = torch.randint(0, 2000, (n, h, w, 3), generator=g, dtype=torch.int32).numpy().astype(np.uint16)
    long = torch.randint(0, 65536, (n, h, w, 3), generator=g, dtype=torch.int32).numpy().astype(np.uint16)
    ratios = [100.0, 250.0, 300.0]
    out = decode_batch(short[..., ::-1].copy(), long[..., ::-1].copy(), ratios, bgr=True)  # cv2 order in, RGB out
    for i in range(n):
        ref = O.decode_pair(short[i], long[i], ratios[i])
        for k in ("lq", "gt", "short_raw"):
            assert np.array_equal(out[k][i].cpu().numpy(), ref[k].numpy()), (i, k)
    assert out["expo_ratio"].shape == (n, 1, 1, 1)


@gpu
def test_decode_generic_channel_counts():
    from lowlight_image_enhancement_b200 import _functional as Fn
    g = torch.Generator().manual_seed(321)
    for c in (1, 2, 4):
        u = torch.randint(0, 65536, (2, 9, 13, c), generator=g, dtype=torch.int32).numpy().astype(np.uint16)
        raw, ali = Fn.decode_u16(torch.from_numpy(u).cuda(), 3.0, crop=(1, 2, 7, 9))
        want = (u.astype(np.float32) / 65535.0)[:, 1:8, 2:11, :].transpose(0, 3, 1, 2)
        assert np.array_equal(raw.cpu().numpy(), want)
        assert np.array_equal(ali.cpu().numpy(), np.clip(want * np.float32(3.0), 0.0, 1.0))


@gpu
def test_decode_errors():
    from lowlight_image_enhancement_b200 import _functional as Fn
    from lowlight_image_enhancement_b200.sony_sid_decode import decode_pair
    u = torch.zeros(8, 8, 3, dtype=torch.uint16, device="cuda")
    with pytest.raises(ValueError):
        Fn.decode_u16(u, 1.0, crop=(4, 4, 8, 8))
    with pytest.raises(TypeError):
        Fn.decode_u16(u.to(torch.int32), 1.0)
    with pytest.raises(TypeError):
        decode_pair(np.zeros((8, 8, 3), np.float32), np.zeros((8, 8, 3), np.uint16), 1.0)
    with pytest.raises(ValueError):
        decode_pair(np.zeros((8, 8, 4), np.uint16), np.zeros((8, 8, 4), np.uint16), 1.0)


@gpu
def test_decode_full_size_round_trip_and_clip():
    """One SID-sized frame: every sample must come back from raw * 65535 (rounded), planes must be the de-interleaved
    source, aligned must equal clip(raw * rho) elementwise (checked on the device in plain torch)."""
    from lowlight_image_enhancement_b200 import _functional as Fn
    h, w = 2848, 4256
    g = torch.Generator(device="cuda").manual_seed(322)
    u = torch.randint(0, 65536, (1, h, w, 3), generator=g, dtype=torch.int32, device="cuda").to(torch.uint16)
    raw, ali = Fn.decode_u16(u, 100.0)
    back = torch.round(raw * 65535.0).to(torch.int32).permute(0, 2, 3, 1)
    assert torch.equal(back, u.to(torch.int32))
    # tensor / tensor: ATen's CUDA "divide by a scalar" multiplies by the reciprocal, which is not numpy's division
    planes = u.to(torch.float32).permute(0, 3, 1, 2)
    assert torch.equal(raw, planes / torch.full_like(planes, 65535.0))
    assert torch.equal(ali, (raw * 100.0).clamp(0.0, 1.0))


@gpu
def test_linear_to_srgb_vs_reference_and_oracle():
    from lowlight_image_enhancement_b200.metrics import evaluation as E
    x = srgb_inputs()
    s = E._linear_to_srgb(x.cuda())
    np.testing.assert_allclose(s.cpu().numpy(), GOLD["srgb_out"], rtol=RTOL, atol=0)
    s2, lp = E.srgb_and_lpips_inputs(x.cuda())
    assert torch.equal(s, s2)
    assert torch.equal(lp, E._norm_to_lpips_range(s))
    np.testing.assert_allclose(lp.cpu().numpy(), GOLD["lpips_out"], rtol=0, atol=2e-6)
    # larger seeded input incl. non-finite values and an unaligned view
    g = torch.Generator().manual_seed(323)
    big = torch.rand(3 * 517 * 301 + 5, generator=g) * 1.3 - 0.15
    big[:4] = torch.tensor([float("nan"), float("inf"), -float("inf"), 0.0031308])
    for view in (big, big[1:]):
        got = E._linear_to_srgb(view.cuda()).cpu()
        ref = O.linear_to_srgb(view)
        assert torch.equal(torch.isnan(got), torch.isnan(ref))
        ok = ~torch.isnan(ref)
        rel = ((got[ok].double() - ref[ok].double()).abs() / ref[ok].double().abs().clamp_min(1e-30)).max()
        assert float(rel) <= RTOL, float(rel)
        # against float64 evaluation of the same formula: the kernel must not be further from it than eager fp32 + 4 ulp
        x64 = view[ok].double().clamp(0, 1)
        ref64 = torch.where(x64 <= 0.0031308, 12.92 * x64, 1.055 * x64.clamp(min=0.0031308) ** (1 / 2.4) - 0.055)
        assert float((got[ok].double() - ref64).abs().max()) <= 5e-7


@gpu
def test_evaluation_helpers_sse_sobel_to3ch():
    from lowlight_image_enhancement_b200.metrics import evaluation as E
    x = srgb_inputs()
    y = (x + 0.05 * torch.randn(x.shape, generator=torch.Generator().manual_seed(312))).clamp(0, 1)
    sse = E._compute_channelwise_psnr_sse(x.clamp(0, 1).cuda(), y.cuda())
    np.testing.assert_allclose(sse.cpu().numpy(), GOLD["sse_out"], rtol=RTOL)
    s = torch.from_numpy(GOLD["srgb_out"])
    mag = E._sobel_magnitude(s.mean(dim=1, keepdim=True).cuda())
    np.testing.assert_allclose(mag.cpu().numpy(), GOLD["sobel_out"], rtol=RTOL, atol=1e-7)
    one = torch.rand(2, 1, 4, 4, device="cuda")
    assert E._to_3ch(one).shape == (2, 3, 4, 4) and torch.equal(E._to_3ch(one)[:, 2], one[:, 0])
    assert E._to_3ch(torch.rand(1, 5, 4, 4)).shape == (1, 3, 4, 4)
