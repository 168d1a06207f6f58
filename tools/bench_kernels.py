"""Per-function device timings for the eval (C4-shaped) and loss (C3-shaped) paths: CUDA events on the
current stream, inputs rotated over buffers larger than L2.  Diagnostic companion of bench.py."""
import argparse
import json

import torch

from lowlight_image_enhancement_b200 import _functional as Fn
from lowlight_image_enhancement_b200 import _native as nat


def timed(fn, iters, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--h", type=int, default=2848)
    ap.add_argument("--w", type=int, default=4256)
    ap.add_argument("--pool", type=int, default=3)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    H, W = args.h, args.w
    g = torch.Generator(device=dev).manual_seed(4000)
    pool = []
    for _ in range(args.pool):
        gt = torch.round(255 * torch.rand(1, 3, H, W, device=dev, generator=g)) / 255
        gt = torch.round(255 * torch.nn.functional.avg_pool2d(gt, 5, 1, 2)) / 255
        pred = (gt + 0.03 * torch.randn(gt.shape, device=dev, generator=g)).clamp(0, 1)
        pool.append((pred.contiguous(), gt.contiguous()))
    px = H * W
    out = {}

    def rec(name, ms, bytes_per_px):
        out[name] = {"ms": round(ms, 4), "Mpix/s": round(px / ms / 1e3, 1), "alg_GB/s": round(bytes_per_px * px / ms / 1e6, 1)}
        print(name, out[name], flush=True)

    from lowlight_image_enhancement_b200.eval import METRIC_FUNCTIONS, image_metrics
    from lowlight_image_enhancement_b200.metrics.ssim import torchmetrics_ssim_scores
    from lowlight_image_enhancement_b200.metrics.color_error import _edge_core
    sel = set(args.only.split(",")) if args.only else None

    def want(n):
        return sel is None or n in sel

    if want("psnr"):
        rec("psnr", timed(lambda i: Fn.psnr_stats(pool[i % args.pool][1], pool[i % args.pool][0], 1, 1, 3 * px, cast_first=True,
                                                  clamp_hi=None, data_range=1.0, eps=1e-12), args.iters), 24)
    if want("ssim_tm"):
        rec("ssim_tm", timed(lambda i: torchmetrics_ssim_scores(pool[i % args.pool][0], pool[i % args.pool][1], data_range=1.0,
                                                                sigma=1.5, k1=0.01, k2=0.03), args.iters), 24)
    if want("ssim_sweep"):
        f = lambda i: torchmetrics_ssim_scores(pool[i % args.pool][0], pool[i % args.pool][1], data_range=1.0,
                                               sigma=1.5, k1=0.01, k2=0.03)
        nat.lib().llie_tune(3, 1)
        rec("ssim_tm_tile", timed(f, args.iters), 24)
        nat.lib().llie_tune(3, 0)
        for rows in (0, 48, 64, 96, 128, 192, 256, 384):
            nat.lib().llie_tune(4, rows)
            rec(f"ssim_tm_stream_rows{rows}", timed(f, args.iters), 24)
        nat.lib().llie_tune(4, 0)
    if want("de_sum"):
        rec("de_sum", timed(lambda i: Fn.deltae(pool[i % args.pool][0], pool[i % args.pool][1], variant=0, clamp_in=True), args.iters), 24)
    if want("de_map_l"):
        rec("de_map_l", timed(lambda i: Fn.deltae(pool[i % args.pool][0], pool[i % args.pool][1], variant=0, clamp_in=True,
                                                  want_map=True, want_l=True), args.iters), 32)
    if want("sobel"):
        lpl = torch.rand(1, H, W, device=dev) * 100
        rec("sobel", timed(lambda i: Fn.sobel_magnitude(lpl), args.iters), 8)
    if want("select"):
        v = torch.rand(1, px, device=dev)
        rec("select_q1", timed(lambda i: Fn.select_quantiles(v, [0.85], joint=False), args.iters), 4)
        thr = torch.tensor([0.85], device=dev)
        rec("select_gated", timed(lambda i: Fn.select_quantiles(v, [0.95], joint=True, gate=v, gate_thr=thr, want_sum=True),
                                  args.iters), 8)
    if want("edge"):
        rec("edge_total", timed(lambda i: _edge_core(pool[i % args.pool][0], pool[i % args.pool][1], 0.85, dict(clamp01=True)),
                                args.iters), 36)
    if want("all5"):
        names = sorted(METRIC_FUNCTIONS)
        rec("all5_metrics", timed(lambda i: image_metrics(pool[i % args.pool][0], pool[i % args.pool][1], names), args.iters), 36)
    if want("eval4"):
        names = ["linear_psnr", "linear_ssim", "deltae2000_mean", "edge_deltae2000_mean"]
        rec("eval4_metrics", timed(lambda i: image_metrics(pool[i % args.pool][0], pool[i % args.pool][1], names), args.iters), 36)

    # ---- loss path, C3 shape
    if want("loss"):
        from lowlight_image_enhancement_b200.NewBP_model.losses import DeltaE00Loss, HybridLossPlus, SSIMLoss
        n = 16
        B = torch.rand(n, 3, 512, 512, device=dev)
        sets = []
        for _ in range(4):
            bh = (B + 0.05 * torch.randn_like(B)).clamp(0, 1)
            sets.append(bh)
        A = (B * 0.9).contiguous()
        rho = torch.ones(n, 1, 1, 1, device=dev)
        pxl = n * 512 * 512
        kern = torch.tensor([[.01, .02, .01], [.02, .88, .02], [.01, .02, .01]], device=dev).view(1, 1, 3, 3)

        def recl(name, ms, bpp):
            out[name] = {"ms": round(ms, 4), "Mpix/s": round(pxl / ms / 1e3, 1), "alg_GB/s": round(bpp * pxl / ms / 1e6, 1)}
            print(name, out[name], flush=True)

        de, ss = DeltaE00Loss(), SSIMLoss()
        hyb = HybridLossPlus(device="cuda", w_phys=0.05, physics_kernel=kern, check_finite=False, perceptual_module=False)

        def run(mod):
            def f(i):
                x = sets[i % 4].detach().requires_grad_(True)
                mod(x, B).backward()
            return f

        def run_h(i):
            x = sets[i % 4].detach().requires_grad_(True)
            tot, _ = hyb(Bhat_raw=x, B_raw=B, A_raw=A, expo_ratio=rho, Bhat_srgb01=x, B_srgb01=B)
            tot.backward()

        with torch.no_grad():
            recl("de_loss_fwd_only", timed(lambda i: de(sets[i % 4], B), args.iters), 24)
            recl("ssim_loss_fwd_only", timed(lambda i: ss(sets[i % 4], B), args.iters), 24)
        recl("de_loss_fwd_bwd", timed(run(de), args.iters), 24 + 36)
        recl("ssim_loss_fwd_bwd", timed(run(ss), args.iters), 24 + 36)
        recl("hybrid_fwd_bwd", timed(run_h, args.iters), 36 + 48)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
