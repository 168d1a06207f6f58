"""Pin the CPU oracle against vectors produced by the UNMODIFIED reference modules
(tests/golden/make_golden.py).  CPU only.  Same torch ops in the same order -> we demand
(near) bit equality: 1e-7 absolute unless noted."""
import warnings

import numpy as np
import pytest
import torch

from oracle import colour, psf_ops, quality
from tests.golden_inputs import (LAB_UNSTABLE, PHYS_CASES, SSIM_CASES, dense_from_spec, l_plane_inputs, lab_inputs,
                                 loss_inputs, phys_inputs, quality_inputs)


def close(a, b, atol=1e-7, rtol=1e-6):
    a = a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    np.testing.assert_allclose(a, b, atol=atol, rtol=rtol)


@pytest.mark.parametrize("i", range(len(PHYS_CASES)))
def test_phys_metric_matches_reference(golden, i):
    case = PHYS_CASES[i]
    pred, obs, rho = phys_inputs(case)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        val, mp = psf_ops.phys_metric(pred, obs, dense_from_spec(case["psf"]), rho,
                                      reduction=case["reduction"], padding=case["padding"], crop=case["crop"],
                                      robust=case["robust"], eps=case["eps"], clamp01=case["clamp01"],
                                      normalize_psf=case.get("normalize_psf", True), return_map=True)
    close(val, golden[f"phys{i}_val"], atol=0, rtol=0)
    close(mp, golden[f"phys{i}_map"], atol=0, rtol=0)


@pytest.mark.parametrize("mode", ["mono", "rgb"])
def test_crosstalk_and_losses_match_reference(golden, mode):
    k = psf_ops.canonical_kernels(mode)
    close(k, golden[f"kern_{mode}"], atol=0, rtol=0)
    kn = psf_ops.normalise_crosstalk(k)
    close(kn, golden[f"kern_{mode}_norm"], atol=0, rtol=0)
    bhat, a, rho = loss_inputs(mode)
    gy = torch.randn(bhat.shape, generator=torch.Generator().manual_seed(77))
    close(psf_ops.crosstalk_forward(bhat, kn), golden[f"xtalk_{mode}_y"], atol=0, rtol=0)
    kd = kn.expand(3, 1, 3, 3).contiguous()
    close(psf_ops.newbp_forward(bhat, kd, 1, 3), golden[f"newbp_{mode}_fwd"], atol=0, rtol=0)
    close(psf_ops.newbp_adjoint(gy, kd, 1, 3), golden[f"newbp_{mode}_gx"], atol=0, rtol=0)
    close(psf_ops.newbp_adjoint(gy, kd, 1, 3), golden[f"xtalk_{mode}_gx"])
    x = bhat.clone().requires_grad_(True)
    ls = psf_ops.phys_loss_srgb(x, a, rho, kn)
    ls.backward()
    close(ls, golden[f"physloss_srgb_{mode}_val"], atol=0, rtol=0)
    close(x.grad, golden[f"physloss_srgb_{mode}_grad"], atol=0, rtol=0)
    for tag, kk in (("k3", k), ("k5", torch.ones(k.shape[0], 1, 5, 5) / 25.0)):
        x = bhat.clone().requires_grad_(True)
        lr = psf_ops.phys_loss_raw(x, a, rho.view(-1), kk)
        lr.backward()
        close(lr, golden[f"physloss_raw_{mode}_{tag}_val"], atol=0, rtol=0)
        close(x.grad, golden[f"physloss_raw_{mode}_{tag}_grad"], atol=0, rtol=0)


def test_psnr_family_matches_reference(golden, golden_scalars):
    pred, tgt = quality_inputs()
    for dr in (1.0, 255.0):
        for red in ("none", "mean", "sum"):
            close(quality.psnr_linear(pred * dr, tgt * dr, data_range=dr, reduction=red),
                  golden[f"psnr_linear_dr{int(dr)}_{red}"], atol=0, rtol=0)
    close(quality.psnr_linear(pred * 1.5 - 0.2, tgt, clamp=True, reduction="none"), golden["psnr_linear_clamp"], 0, 0)
    r = quality.rgb_psnr(pred, tgt, reduction="none")
    close(torch.stack([r["R"], r["G"], r["B"], r["mean"]]), golden["rgb_psnr_none"], 0, 0)
    close(quality.cpsnr_rgb(pred, tgt, reduction="none"), golden["cpsnr_none"], 0, 0)
    close(quality.cpsnr_rgb(pred * 1.5 - 0.2, tgt, clamp=True, reduction="none"), golden["cpsnr_clamp"], 0, 0)
    assert quality.calculate_psnr(tgt, pred, 1.0) == golden_scalars["calculate_psnr"]
    assert quality.calculate_psnr(tgt, tgt, 1.0) == float("inf")


@pytest.mark.parametrize("j", range(len(SSIM_CASES)))
def test_ssim_linear_matches_reference(golden, j):
    pred, tgt = quality_inputs()
    close(quality.ssim_linear(pred, tgt, reduction="none", channel_aggregate="none", **SSIM_CASES[j]),
          golden[f"ssim_linear{j}"], atol=0, rtol=0)


def test_rgb_ssim_and_f64_match_reference(golden):
    pred, tgt = quality_inputs()
    r = quality.rgb_ssim(pred, tgt, reduction="none")
    close(torch.stack([r["R"], r["G"], r["B"], r["mean"]]), golden["rgb_ssim_none"], 0, 0)
    close(quality.ssim_linear(pred.double(), tgt.double(), reduction="none", channel_aggregate="none"),
          golden["ssim_linear_f64"], atol=0, rtol=0)


def test_ciede2000_variants_match_reference(golden):
    lab1, lab2 = lab_inputs()
    close(lab1, golden["lab1"], 0, 0)
    close(colour.de00_metric_lab(lab1, lab2), golden["de_metric"], atol=0, rtol=0)
    close(colour.de00_metric_lab(lab1, lab2, kL=2.0, kC=1.5, kH=0.7), golden["de_metric_k"], atol=0, rtol=0)
    x = lab1.clone().requires_grad_(True)
    de = colour.de00_loss_lab(x, lab2)
    close(de, golden["de_loss"], atol=0, rtol=0)
    de.sum().backward()
    close(x.grad, golden["de_loss_grad_lab1"], atol=1e-6, rtol=1e-5)
    # the fp64 run bounds how noisy fp32 eager itself is on these pairs
    # (pairs 32..35 are exact opposite hues, |dh| = pi: a branch discontinuity of the formula where
    #  fp32 and fp64 legitimately disagree by tens of units -> excluded everywhere, see LAB_UNSTABLE)
    keep = np.ones(128, bool)
    keep[LAB_UNSTABLE] = False
    assert np.abs(golden["de_metric"] - golden["de_metric_f64"]).reshape(-1)[keep].max() < 1e-4
    assert np.abs(golden["de_loss"] - golden["de_loss_f64"]).reshape(-1)[keep].max() < 1e-4


def test_sobel_quantile_count_match_reference(golden):
    lp = l_plane_inputs()
    close(lp, golden["sobel_L"], 0, 0)
    for q in (0.85, 0.5, 0.999):
        g, thr, cnt = colour.edge_from_l(lp, q)
        close(g, golden["sobel_g"], atol=0, rtol=0)
        close(thr, golden[f"sobel_thr_q{q}"], atol=0, rtol=0)
        assert cnt.tolist() == golden[f"sobel_cnt_q{q}"].tolist()


def test_ssim_policies_agree_away_from_borders():
    """kornia / torchmetrics restatements are unpinned; cross-check them with the pinned ssim_linear:
    torchmetrics == interior mean of the ssim_linear map up to the var clamp/eps (tiny)."""
    pred, tgt = quality_inputs()
    tm = quality.ssim_torchmetrics(pred, tgt)
    km = quality.kornia_ssim_map(pred, tgt)
    inner = km[..., 5:-5, 5:-5].reshape(2, -1).mean(-1)
    close(tm, inner.numpy(), atol=2e-6, rtol=0)
    lin = quality.ssim_linear(pred, tgt, reduction="none")
    assert abs(float(km.mean()) - float(lin.mean())) < 1e-5


def test_rgb_to_lab_sanity_vs_cv2():
    cv2 = pytest.importorskip("cv2")
    g = torch.Generator().manual_seed(5)
    rgb = torch.rand(1, 3, 16, 16, generator=g)
    lab = colour.rgb_to_lab(rgb)[0].permute(1, 2, 0).numpy()
    ref = cv2.cvtColor(rgb[0].permute(1, 2, 0).numpy().copy(), cv2.COLOR_RGB2Lab)
    assert np.abs(lab - ref).max() < 0.5  # different white point rounding; loose by design


def test_rgb_to_lab_published_anchor_values():
    """The restated kornia 0.6.12 `rgb_to_lab` (absent from the reference tree and from both images: 'parity unpinned')
    is at least anchored on published colorimetry: CIE L*a*b* (D65 / 2 deg) of the sRGB primaries and secondaries as
    tabulated by Lindbloom's calculator for the IEC 61966-2-1 matrix -- agreement to the table's rounding and the
    4-digit matrix kornia uses (2e-2), white and grey neutral to 5e-3."""
    from oracle import colour
    table = {
        (1.0, 0.0, 0.0): (53.2408, 80.0925, 67.2032),
        (0.0, 1.0, 0.0): (87.7347, -86.1827, 83.1793),
        (0.0, 0.0, 1.0): (32.2970, 79.1875, -107.8602),
        (1.0, 1.0, 0.0): (97.1393, -21.5537, 94.4780),
        (0.0, 1.0, 1.0): (91.1132, -48.0875, -14.1312),
        (1.0, 0.0, 1.0): (60.3242, 98.2343, -60.8249),
        (1.0, 1.0, 1.0): (100.0, 0.0, 0.0),
        (0.0, 0.0, 0.0): (0.0, 0.0, 0.0),
    }
    rgb = torch.tensor(list(table.keys()), dtype=torch.float64).T.reshape(1, 3, 1, len(table))
    lab = colour.rgb_to_lab(rgb).reshape(3, len(table)).T
    want = torch.tensor(list(table.values()), dtype=torch.float64)
    assert float((lab - want).abs().max()) < 2e-2, (lab - want).abs().max()
    grey = colour.rgb_to_lab(torch.full((1, 3, 1, 5), 0.5, dtype=torch.float64) * torch.linspace(0.1, 1, 5).view(1, 1, 1, 5) * 2)
    assert float(grey[:, 1:].abs().max()) < 5e-3
