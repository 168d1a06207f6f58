"""Generate golden vectors by running the UNMODIFIED reference modules (imported from
/root/reference, CPU torch) on seeded inputs.  Run in the build container only:

    python tests/golden/make_golden.py

Writes tests/golden/reference_vectors.npz (inputs + reference outputs, small) and
tests/golden/reference_scalars.json (seed-defined larger cases: only scalars are stored, the inputs
are regenerated from the seed by tests/golden_inputs.py).  /root/reference does not exist on the
GPU box, so nothing at test time imports it: tests read these files only.

What is and is not covered: phys_consistency, newbp_layer, the two physics losses,
linear, channelwise, psnr, _deltaE00_lab_map, DeltaE00Loss._ciede2000, _sobel_magnitude and
torch.quantile are the reference's own code.  kornia / torchmetrics are not installed, so
sRGB->Lab, the kornia SSIM loss and the torchmetrics SSIM have NO golden here (parity unpinned).
"""
from __future__ import annotations

import json
import os
import sys
import warnings

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
sys.path.insert(0, REF)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from metrics import phys_consistency as ref_phys  # noqa: E402
from metrics import linear as ref_linear  # noqa: E402
from metrics import channelwise as ref_cw  # noqa: E402
from metrics import psnr as ref_psnr  # noqa: E402
from metrics import color_error as ref_ce  # noqa: E402
from NewBP_model import newbp_layer as ref_bp  # noqa: E402
from NewBP_model import losses as ref_losses  # noqa: E402

from tests.golden_inputs import (  # noqa: E402
    PHYS_CASES, SSIM_CASES, phys_inputs, quality_inputs, lab_inputs, l_plane_inputs, loss_inputs,
    dense_from_spec)

torch.set_num_threads(4)
vec: dict[str, np.ndarray] = {}
scal: dict[str, object] = {}


def put(name, t):
    vec[name] = t.detach().cpu().numpy() if torch.is_tensor(t) else np.asarray(t)


# ------------------------------------------------------------------ phys_cons_raw / phys_cons_srgb
for i, case in enumerate(PHYS_CASES):
    pred, obs, rho = phys_inputs(case)
    psf = dense_from_spec(case["psf"])
    fn = ref_phys.phys_cons_srgb if case["srgb"] else ref_phys.phys_cons_raw
    kw = dict(reduction=case["reduction"], padding=case["padding"], crop=case["crop"],
              robust=case["robust"], eps=case["eps"], return_map=True,
              normalize_psf=case.get("normalize_psf", True))
    if case["srgb"]:
        kw["clamp01"] = case["clamp01"]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        val, mp = fn(pred, obs, psf, rho, **kw)
    put(f"phys{i}_val", val)
    put(f"phys{i}_map", mp)

# ------------------------------------------------------------------ CrosstalkPSF forward + NewBP adjoint, physics losses
for mode in ("mono", "rgb"):
    spec = "P2" if mode == "mono" else "B2"
    k = ref_bp.build_psf_kernels(mode, spec)
    put(f"kern_{mode}", k)
    psf_mod = ref_bp.CrosstalkPSF(mode, k)
    put(f"kern_{mode}_norm", psf_mod.kernel)
    bhat, a, rho = loss_inputs(mode)
    x = bhat.clone().requires_grad_(True)
    y = psf_mod(x)
    gy = torch.randn(y.shape, generator=torch.Generator().manual_seed(77))
    y.backward(gy)
    put(f"xtalk_{mode}_y", y)
    put(f"xtalk_{mode}_gx", x.grad)
    kd = psf_mod.kernel.expand(3, 1, 3, 3).contiguous()
    put(f"newbp_{mode}_fwd", ref_bp.NewBPFunction.apply(bhat, kd, 1, 3))
    x2 = bhat.clone().requires_grad_(True)
    ref_bp.NewBPFunction.apply(x2, kd, 1, 3).backward(gy)
    put(f"newbp_{mode}_gx", x2.grad)
    # sRGB physics loss (losses.py:206-220)
    x3 = bhat.clone().requires_grad_(True)
    ls = ref_losses.PhysicalConsistencyLossSRGB(psf_mod)(x3, a, rho)
    ls.backward()
    put(f"physloss_srgb_{mode}_val", ls)
    put(f"physloss_srgb_{mode}_grad", x3.grad)
    # RAW physics loss (losses.py:158-192), 3x3 and a 5x5 variant
    for tag, kk in (("k3", k), ("k5", torch.ones(k.shape[0], 1, 5, 5) / 25.0)):
        x4 = bhat.clone().requires_grad_(True)
        lr = ref_losses.PhysicsConsistencyLoss(kk, device="cpu")(x4, a, rho.view(-1))
        lr.backward()
        put(f"physloss_raw_{mode}_{tag}_val", lr)
        put(f"physloss_raw_{mode}_{tag}_grad", x4.grad)

# ------------------------------------------------------------------ PSNR family, ssim_linear, rgb_ssim
pred, tgt = quality_inputs()
for dr in (1.0, 255.0):
    for red in ("none", "mean", "sum"):
        put(f"psnr_linear_dr{int(dr)}_{red}", ref_linear.psnr_linear(pred * dr, tgt * dr, data_range=dr, reduction=red))
put("psnr_linear_clamp", ref_linear.psnr_linear(pred * 1.5 - 0.2, tgt, clamp=True, reduction="none"))
r = ref_cw.rgb_psnr(pred, tgt, reduction="none")
put("rgb_psnr_none", torch.stack([r["R"], r["G"], r["B"], r["mean"]]))
put("cpsnr_none", ref_cw.cpsnr_rgb(pred, tgt, reduction="none"))
put("cpsnr_clamp", ref_cw.cpsnr_rgb(pred * 1.5 - 0.2, tgt, clamp=True, reduction="none"))
scal["calculate_psnr"] = ref_psnr.calculate_psnr(tgt, pred, 1.0)
scal["calculate_psnr_identical"] = ref_psnr.calculate_psnr(tgt, tgt, 1.0)
for j, sc in enumerate(SSIM_CASES):
    put(f"ssim_linear{j}", ref_linear.ssim_linear(pred, tgt, reduction="none", channel_aggregate="none", **sc))
r = ref_cw.rgb_ssim(pred, tgt, reduction="none")
put("rgb_ssim_none", torch.stack([r["R"], r["G"], r["B"], r["mean"]]))
put("ssim_linear_f64", ref_linear.ssim_linear(pred.double(), tgt.double(), reduction="none", channel_aggregate="none"))

# ------------------------------------------------------------------ CIEDE2000 (both variants) on Lab inputs
lab1, lab2 = lab_inputs(os.path.join(REF, "standard_tests/data/ciede2000_pairs.json"))
put("lab1", lab1)
put("lab2", lab2)
put("de_metric", ref_ce._deltaE00_lab_map(lab1, lab2, kL=1.0, kC=1.0, kH=1.0, eps=1e-12))
put("de_metric_k", ref_ce._deltaE00_lab_map(lab1, lab2, kL=2.0, kC=1.5, kH=0.7, eps=1e-12))
l1g = lab1.clone().requires_grad_(True)
de_loss = ref_losses.DeltaE00Loss._ciede2000(l1g, lab2, 1e-6)
put("de_loss", de_loss)
de_loss.sum().backward()
put("de_loss_grad_lab1", l1g.grad)
# fp64 run of the same reference functions (separates "we differ" from "fp32 eager is noisy")
put("de_metric_f64", ref_ce._deltaE00_lab_map(lab1.double(), lab2.double(), kL=1.0, kC=1.0, kH=1.0, eps=1e-12))
put("de_loss_f64", ref_losses.DeltaE00Loss._ciede2000(lab1.double(), lab2.double(), 1e-6))

# ------------------------------------------------------------------ Sobel magnitude + quantile threshold + count
lp = l_plane_inputs()
g = ref_ce._sobel_magnitude(lp)
put("sobel_L", lp)
put("sobel_g", g)
for q in (0.85, 0.5, 0.999):
    thr = torch.quantile(g.view(g.shape[0], -1), q, dim=1, keepdim=True)
    put(f"sobel_thr_q{q}", thr.view(-1))
    put(f"sobel_cnt_q{q}", (g >= thr.view(-1, 1, 1, 1)).flatten(1).sum(dim=1))

np.savez_compressed(os.path.join(HERE, "reference_vectors.npz"), **vec)
with open(os.path.join(HERE, "reference_scalars.json"), "w") as fh:
    json.dump(scal, fh, indent=1)
print("wrote", len(vec), "arrays;", sum(v.nbytes for v in vec.values()) // 1024, "KiB raw")
