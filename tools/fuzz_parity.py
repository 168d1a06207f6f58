"""Randomised parity sweep of the CUDA kernels against the CPU oracle (not part of the test-suite: a long-running
confidence check, `PYTHONPATH=. python tools/fuzz_parity.py [cases] [seed]`).  Shapes are drawn to hit the dispatch
boundaries: W % 4 == 0 or not (TMA ring vs cp.async / tile kernels), partial 256-column strips, planes shorter than one
row segment, windows from 3 to 25 taps, every padding mode, batches above the per-round limits of the finalizes.
Prints one line per failing case and a summary; exit code = number of failures."""
import random
import sys

import torch

from lowlight_image_enhancement_b200 import _functional as Fn
from lowlight_image_enhancement_b200.metrics.color_error import deltaE2000_map, edge_deltaE2000_stats
from lowlight_image_enhancement_b200.metrics.linear import psnr_linear, ssim_linear
from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw
from lowlight_image_enhancement_b200.metrics.ssim import calculate_ssim_per_image
from lowlight_image_enhancement_b200.NewBP_model.losses import DeltaE00Loss, SSIMLoss
from oracle import colour, psf_ops, quality

dev = torch.device("cuda:0")
cases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
rnd = random.Random(seed)
fails = []


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def check(name, ok, info):
    if not ok:
        fails.append((name, info))
        print("FAIL", name, info, flush=True)


def shape(min_hw=12, max_h=300, max_w=600, c=None):
    n = rnd.choice([1, 1, 2, 3, 9])
    c = c or rnd.choice([1, 2, 3])
    h = rnd.randint(min_hw, max_h)
    w = rnd.choice([rnd.randint(min_hw, max_w), 4 * rnd.randint(min_hw // 4 + 1, max_w // 4), 256, 260, 512])
    return n, c, h, w


for it in range(cases):
    g = torch.Generator().manual_seed(seed * 100003 + it)
    # ---- SSIMLoss value + gradient (kornia policy)
    n, c, h, w = shape()
    ws = rnd.choice([3, 5, 7, 11, 11, 11, 13, 15, 17, 25])
    if min(h, w) > ws:
        tgt = torch.rand(n, c, h, w, generator=g)
        gen = (tgt + 0.1 * torch.randn(n, c, h, w, generator=g)).clamp(-0.05, 1.05)
        xr = gen.double().requires_grad_(True)
        lr = quality.ssim_loss_kornia(xr, tgt.double(), ws, 1.0)
        lr.backward()
        x = gen.to(dev).requires_grad_(True)
        lo = SSIMLoss(window_size=ws)(x, tgt.to(dev))
        lo.backward()
        e_grad, e_f32 = rel(x.grad, xr.grad), 0.0
        if e_grad >= 5e-5:
            # small windows: sigma^2 = E[x^2] - mu^2 over 9 or 25 pixels cancels, and the float32 evaluation of the oracle
            # itself is that far from float64 (tools/diag_ssim3_grad.py); judge against it
            x32 = gen.clone().requires_grad_(True)
            quality.ssim_loss_kornia(x32, tgt, ws, 1.0).backward()
            e_f32 = rel(x32.grad, xr.grad)
        check("ssim_loss", abs(float(lo.detach()) - float(lr.detach())) <= 2e-5 * abs(float(lr.detach())) and
              e_grad < max(5e-5, 3.0 * e_f32), ((n, c, h, w), ws, float(lo.detach()), float(lr.detach()), e_grad, e_f32))
    # ---- ssim_linear / torchmetrics policy
    n, c, h, w = shape(min_hw=26)
    k = rnd.choice([3, 7, 11, 11, 11, 15, 21])
    pad = rnd.choice(["reflect", "replicate", "circular", "constant"])
    tgt = torch.rand(n, c, h, w, generator=g)
    pred = (tgt + 0.08 * torch.randn(n, c, h, w, generator=g)).clamp(0, 1)
    kw = dict(kernel_size=k, sigma=k / 7.0, padding=pad, reduction="none", channel_aggregate="none")
    ref = quality.ssim_linear(pred.double(), tgt.double(), **kw)
    out = ssim_linear(pred.to(dev), tgt.to(dev), **kw)
    check("ssim_linear", rel(out, ref) < 2e-5, ((n, c, h, w), k, pad, rel(out, ref)))
    ref = quality.ssim_torchmetrics(pred.double(), tgt.double(), data_range=1.0)
    out = calculate_ssim_per_image(tgt.to(dev), pred.to(dev), 1.0)
    check("ssim_tm", rel(out, ref) < 2e-5, ((n, c, h, w), rel(out, ref)))
    check("psnr", rel(psnr_linear(pred.to(dev), tgt.to(dev), reduction="none"),
                      quality.psnr_linear(pred, tgt, reduction="none")) < 1e-6, (n, c, h, w))
    # ---- dE00 metric map + edge statistics on 8-bit images (bit-exact count)
    n, _, h, w = shape(min_hw=16, max_h=200, max_w=400, c=3)
    gt8 = torch.round(255 * torch.rand(n, 3, h, w, generator=g)) / 255
    gt8 = torch.round(255 * torch.nn.functional.avg_pool2d(gt8, 3, 1, 1)) / 255
    pr8 = torch.round(255 * (gt8 + 0.05 * torch.randn(n, 3, h, w, generator=g)).clamp(0, 1)) / 255
    m = deltaE2000_map(pr8.to(dev), gt8.to(dev)).cpu().double()
    r64 = colour.de00_metric_lab(colour.rgb_to_lab(pr8.double()), colour.rgb_to_lab(gt8.double()))
    lab = colour.rgb_to_lab(pr8.double())
    lab2 = colour.rgb_to_lab(gt8.double())
    chroma_ok = ((lab[:, 1] ** 2 + lab[:, 2] ** 2) > 1.0) & ((lab2[:, 1] ** 2 + lab2[:, 2] ** 2) > 1.0)
    d = (m.view_as(r64) - r64).abs()
    frac = float((d[chroma_ok] > 1e-4).double().mean()) if chroma_ok.any() else 0.0
    # CIEDE2000 is discontinuous where the hue difference crosses 180 degrees (the mean hue jumps by pi): a pixel pair
    # on that branch changes by O(1) with the last bit of a', in float32 eager as here (seen once in ~900 cases: one pixel
    # off by 2.9).  Such pixels are counted and reported, the mean is taken over the others.
    jumps = int((d > 0.05).sum())
    if jumps:
        print("dE00 map: pixels on a formula discontinuity:", jumps, "of", d.numel(), "max", float(d.max()), flush=True)
    check("de00_map", frac < 2e-3 and float(d[d <= 0.05].mean()) < 3e-5 and jumps <= 2 + d.numel() // 100000,
          ((n, h, w), frac, float(d.mean()), float(d.max()), jumps))
    es, er = edge_deltaE2000_stats(pr8.to(dev), gt8.to(dev), q=0.85), colour.edge_stats(pr8, gt8, q=0.85)
    # the mask count is bit-exact GIVEN the L plane (stage test test_sobel_quantile_count_bit_exact); end to end it also
    # depends on the last bit of L = 116 f(Y) - 16, where this library's x^2.4 / cbrt (< 1 ulp) and the host libm of the
    # CPU oracle may round differently: a pixel whose gradient sits on the threshold can then change sides (seen once
    # in ~700 cases: 69 291 against 69 292 pixels).  One flip per image is tolerated and reported, more is a failure.
    if es["count"] != er["count"]:
        print("edge count differs by", es["count"] - er["count"], "of", er["count"], (n, h, w), flush=True)
    check("edge", abs(es["count"] - er["count"]) <= n and abs(es["mean"] - er["mean"]) < 2e-4,
          ((n, h, w), es["count"], er["count"], es["mean"], er["mean"]))
    # ---- DeltaE00Loss value + gradient
    xr = pr8.double().requires_grad_(True)
    lr = colour.de00_loss(xr, gt8.double()) if hasattr(colour, "de00_loss") else None
    if lr is not None:
        lr.backward()
        x = pr8.to(dev).requires_grad_(True)
        lo = DeltaE00Loss()(x, gt8.to(dev))
        lo.backward()
        gscale = float(xr.grad.abs().max())
        bad = float(((x.grad.cpu().double() - xr.grad).abs() > 1e-3 * gscale).double().mean())
        check("de00_loss", abs(float(lo) - float(lr)) < 1e-4 and bad < 2e-3, ((n, h, w), float(lo), float(lr), bad))
    # ---- phys_cons_raw value + gradient
    n, _, h, w = shape(min_hw=8, max_h=200, max_w=600, c=3)
    pad = rnd.choice(["reflect", "replicate", "zeros"])
    crop = rnd.choice(["same", "valid"])
    robust = rnd.choice(["none", "charbonnier"])
    spec = rnd.choice(["mono", "rgb"])
    rho = torch.tensor([rnd.uniform(0.5, 3.0) for _ in range(n)])
    pred = torch.rand(n, 3, h, w, generator=g)
    kern = psf_ops.canonical_kernels(spec)
    obs = (pred * rho.view(-1, 1, 1, 1) * 0.9 + 0.05 * torch.randn(n, 3, h, w, generator=g)).clamp(0, 1.2)
    xr = pred.double().requires_grad_(True)
    lr = psf_ops.phys_metric(xr, obs.double(), psf_ops.embed_dense(kern).double(), rho.double(), padding=pad, crop=crop,
                             robust=robust, keep_graph=True)
    lr.backward()
    x = pred.to(dev).requires_grad_(True)
    lo = phys_cons_raw(x, obs.to(dev), kern.to(dev), rho.to(dev), padding=pad, crop=crop, robust=robust,
                       differentiable=True)
    lo.backward()
    gs = float(xr.grad.abs().max())
    bad = float(((x.grad.cpu().double() - xr.grad).abs() > 2e-5 * gs).double().mean())
    # (both penalties are sign-like around r = 0 at the default eps 1e-12: a residual at the fp32 rounding level of the
    # stencil sum may take the other sign in the float32 oracle, a few pixels in 1e5)
    check("phys", abs(float(lo) - float(lr)) <= 2e-6 * abs(float(lr)) and bad < 2e-3,
          ((n, h, w), pad, crop, robust, spec, float(lo), float(lr), bad))
    # ---- exact quantile select
    per = rnd.choice([rnd.randint(5, 5000), 4 * rnd.randint(2, 50000), 1 << 20])
    imgs = rnd.choice([1, 2, 5])
    vals = torch.randn(imgs, per, generator=g).abs().pow(rnd.choice([0.5, 1, 3])).to(dev)
    q = rnd.choice([0.5, 0.85, 0.95, 0.0, 1.0])
    joint = rnd.choice([True, False])
    got, counts, _ = Fn.select_quantiles(vals, [q], joint=joint)
    want = torch.quantile(vals.flatten(), q).view(1) if joint else torch.quantile(vals, q, dim=1)
    check("select", torch.equal(got.flatten().cpu(), want.flatten().cpu()), (imgs, per, q, joint))
    if (it + 1) % 10 == 0:
        print(f"{it + 1} / {cases} cases, {len(fails)} failures", flush=True)
print("failures:", len(fails))
sys.exit(min(len(fails), 100))
