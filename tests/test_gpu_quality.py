"""GPU parity: PSNR family (and, further down, SSIM policies) vs golden vectors of the reference and the
CPU oracle.  BASELINE tolerance: 1e-5 relative for PSNR / SSIM."""
import math

import numpy as np
import pytest
import torch

from oracle import quality
from tests.golden_inputs import SSIM_CASES, quality_inputs

pytestmark = pytest.mark.gpu
RTOL = 1e-5


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _close(a, b, rtol=RTOL):
    a = a.detach().double().cpu().numpy() if torch.is_tensor(a) else np.asarray(a, dtype=np.float64)
    np.testing.assert_allclose(a, np.asarray(b, dtype=np.float64), rtol=rtol, atol=0)


def test_psnr_family_vs_golden(golden, golden_scalars, dev):
    from lowlight_image_enhancement_b200.metrics.channelwise import cpsnr_rgb, rgb_psnr
    from lowlight_image_enhancement_b200.metrics.linear import psnr_linear
    from lowlight_image_enhancement_b200.metrics.psnr import calculate_psnr
    pred, tgt = quality_inputs()
    p, t = pred.to(dev), tgt.to(dev)
    for dr in (1.0, 255.0):
        for red in ("none", "mean", "sum"):
            out = psnr_linear(p * dr, t * dr, data_range=dr, reduction=red)
            assert out.dtype == torch.float64
            _close(out, golden[f"psnr_linear_dr{int(dr)}_{red}"], 1e-9)
    _close(psnr_linear(p * 1.5 - 0.2, t, clamp=True, reduction="none"), golden["psnr_linear_clamp"], 1e-9)
    r = rgb_psnr(p, t, reduction="none")
    _close(torch.stack([r["R"], r["G"], r["B"], r["mean"]]), golden["rgb_psnr_none"], 1e-9)
    _close(cpsnr_rgb(p, t, reduction="none"), golden["cpsnr_none"], 1e-9)
    _close(cpsnr_rgb(p * 1.5 - 0.2, t, clamp=True, reduction="none"), golden["cpsnr_clamp"], 1e-9)
    assert abs(calculate_psnr(t, p, 1.0) - golden_scalars["calculate_psnr"]) < 1e-9
    assert calculate_psnr(t, t, 1.0) == float("inf")
    assert math.isinf(float(psnr_linear(p, p)))
    m = rgb_psnr(p, t, meta=True, domain="srgb", data_range=2.0)
    assert m["meta"] == {"domain": "srgb", "data_range": 2.0}


def test_psnr_contracts(dev):
    """standard_tests/test_linear_metrics.py:121-133, test_channelwise.py:54-63,212-222."""
    from lowlight_image_enhancement_b200.metrics.channelwise import cpsnr_rgb, rgb_psnr
    from lowlight_image_enhancement_b200.metrics.linear import psnr_linear
    x = torch.rand(2, 3, 16, 16, device=dev)
    with pytest.raises(TypeError):
        psnr_linear(x.to(torch.int32), x.to(torch.int32))
    with pytest.raises(ValueError):
        psnr_linear(x, x[:, :, :8])
    with pytest.raises(ValueError):
        psnr_linear(x, x, data_range=-1.0)
    with pytest.raises(ValueError):
        rgb_psnr(x[:, :2], x[:, :2])
    bad = x.clone()
    bad[1, 2, 3, 3] = float("inf")
    with pytest.raises(ValueError):
        cpsnr_rgb(bad, x)
    with pytest.raises(ValueError):
        psnr_linear(x, bad)
    assert psnr_linear(x[0], x[0] * 0.5).ndim == 0            # 3-D input, mean reduction
    assert psnr_linear(x.double(), x.double() * 0.5, reduction="none").shape == (2,)
    # odd sizes / unaligned views take the scalar path
    y = torch.rand(1, 3, 17, 13, device=dev)
    z = (y + 0.01).clamp(0, 1)
    _close(psnr_linear(y, z), quality.psnr_linear(y.cpu(), z.cpu()), 1e-9)
    v = torch.rand(1, 3, 20, 21, device=dev)[..., 1:]
    _close(psnr_linear(v, v * 0.9), quality.psnr_linear(v.cpu(), v.cpu() * 0.9), 1e-9)


def test_psnr_full_size_properties(dev):
    """C4 image size (2848x4256): data-range equivalence and an analytic MSE."""
    from lowlight_image_enhancement_b200.metrics.linear import psnr_linear
    from lowlight_image_enhancement_b200.metrics.channelwise import cpsnr_rgb
    g = torch.Generator(device=dev).manual_seed(4000)
    gt = torch.rand(1, 3, 2848, 4256, device=dev, generator=g)
    pred = gt + 0.125
    assert abs(float(psnr_linear(pred, gt)) - 10 * math.log10(1 / 0.125 ** 2)) < 1e-5
    a = float(psnr_linear(pred, gt, data_range=1.0))
    b = float(psnr_linear(pred * 255, gt * 255, data_range=255.0))
    assert abs(a - b) < 1e-5 * abs(a)
    assert abs(float(cpsnr_rgb(pred, gt)) - a) < 1e-5


@pytest.mark.parametrize("names", [
    ["linear_psnr", "linear_ssim", "deltae2000_mean", "edge_deltae2000_mean"],
    ["deltae2000_p95", "linear_ssim"],
    ["edge_deltae2000_mean", "deltae2000_mean", "deltae2000_p95", "linear_psnr", "linear_ssim"],
    ["linear_ssim"],
])
@pytest.mark.parametrize("batch", [1, 2])
def test_one_launch_eval_epilogue_matches_torch_glue(names, batch):
    """`accumulate_image_metrics` (raw device sums -> adapter values -> running vector in ONE launch,
    llie_eval_accumulate) against `image_metrics` + `ShardedMeans.add` (the same arithmetic as torch glue): identical
    float64 sums, counts and non-finite flags; requests it does not cover report False and launch nothing."""
    from lowlight_image_enhancement_b200.eval import ShardedMeans, accumulate_image_metrics, evaluate_pairs, image_metrics
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(41 + batch)
    pairs = []
    for i in range(3):
        gt = torch.round(255 * torch.nn.functional.avg_pool2d(torch.rand(batch, 3, 72, 100, generator=g), 5, 1, 2)) / 255
        pr = (gt + 0.04 * torch.randn(gt.shape, generator=g)).clamp(0, 1)
        pairs.append((pr.to(dev), gt.to(dev)))
    pairs[1][0][0, 1, 5, 5] = float("nan")  # one pair with a non-finite input
    fast, slow = ShardedMeans(names, dev), ShardedMeans(names, dev)
    flags = torch.zeros(1, dtype=torch.int32, device=dev)
    for pr, gt in pairs:
        assert accumulate_image_metrics(pr, gt, names, fast)
        flags.zero_()
        slow.add(image_metrics(pr, gt, names, flags), flags)
    a, b = fast.vec.cpu(), slow.vec.cpu()
    k = len(names)
    assert a[k] == b[k] == 3 and a[k + 1] == b[k + 1] == 1
    ok = ~torch.isnan(b[:k])
    assert torch.equal(torch.isnan(a[:k]), torch.isnan(b[:k]))
    assert torch.allclose(a[:k][ok], b[:k][ok], rtol=1e-12, atol=0), (a, b)
    # the driver takes the fast path and reports the same means
    clean = [pairs[0], pairs[2]]
    res = evaluate_pairs(clean, names, device=dev, rank=0, world_size=1)
    ref = ShardedMeans(names, dev)
    for pr, gt in clean:
        flags.zero_()
        ref.add(image_metrics(pr, gt, names, flags), flags)
    want = ref.reduce()
    for nm in names:
        assert res[nm] == pytest.approx(want[nm], rel=1e-12), nm
    assert res["count"] == 2 and res["nonfinite_inputs"] == 0
    # not covered: PSNR without a colour pass, unknown names, a vector on another device
    assert not accumulate_image_metrics(pairs[0][0], pairs[0][1], ["linear_psnr"], ShardedMeans(["linear_psnr"], dev))
    assert not accumulate_image_metrics(pairs[0][0], pairs[0][1], names, ShardedMeans(names, "cpu"))
