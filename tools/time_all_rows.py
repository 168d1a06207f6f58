"""Every public entry point of the mirror at bench shapes, one line each: us per call (incl. the Python path) and the
algorithmic bytes it moves per second -- a sweep for rows that still run on a slow generic kernel."""
import torch, bench
from lowlight_image_enhancement_b200 import _functional as Fn
from lowlight_image_enhancement_b200 import basicsr_losses as bl
from lowlight_image_enhancement_b200.eval import TileGrid
from lowlight_image_enhancement_b200.metrics import channelwise, color_error, linear, phys_consistency, psnr, ssim
from lowlight_image_enhancement_b200.NewBP_model import losses as L
from lowlight_image_enhancement_b200.NewBP_model.newbp_layer import create_crosstalk_psf
from oracle import psf_ops

dev = torch.device("cuda:0")
peak, _ = bench.peaks()
B = [torch.rand(16, 3, 512, 512, device=dev) for _ in range(4)]
T = [(b + 0.05 * torch.randn_like(b)).clamp(0, 1) for b in B]
big = [bench.synth_eval_pair(4000 + i, dev) for i in range(2)]
kern = psf_ops.canonical_kernels("rgb").to(dev)
rho = torch.rand(16, device=dev) + 0.5


def row(name, fn, nbytes, iters=12, grad=False):
    t = bench.timed_warm(fn, iters, dev) / iters
    print(f"{name:58s} {t*1e3:9.1f} us   {nbytes/t/1e6:8.0f} GB/s  ({nbytes/t/1e6/peak:.2f} of HBM)", flush=True)


e3 = B[0].numel()
e4 = big[0][0].numel()


def lossfb(crit, *extra):
    def f(i):
        x = T[i % 4].detach().requires_grad_(True)
        out = crit(x, B[i % 4], *extra)
        out.backward()
    return f


with torch.no_grad():
    row("phys_cons_raw fwd (C3 shape)", lambda i: phys_consistency.phys_cons_raw(T[i % 4], B[i % 4], kern, rho), 8 * e3)
    row("phys_cons_srgb fwd", lambda i: phys_consistency.phys_cons_srgb(T[i % 4], B[i % 4], kern, rho), 8 * e3)
    row("phys_cons_raw fwd, dense 3x3x3x3 PSF (generic)", lambda i: phys_consistency.phys_cons_raw(
        T[i % 4], B[i % 4], psf_ops.embed_dense(kern) + 0.01, rho), 8 * e3)
    row("phys_cons_raw fwd, 5x5 PSF (generic)", lambda i: phys_consistency.phys_cons_raw(
        T[i % 4], B[i % 4], torch.ones(3, 1, 5, 5, device=dev), rho), 8 * e3)
    row("psnr_linear", lambda i: linear.psnr_linear(T[i % 4], B[i % 4]), 8 * e3)
    row("rgb_psnr", lambda i: channelwise.rgb_psnr(T[i % 4], B[i % 4]), 8 * e3)
    row("cpsnr_rgb", lambda i: channelwise.cpsnr_rgb(T[i % 4], B[i % 4]), 8 * e3)
    row("calculate_psnr (C4 pair)", lambda i: psnr.calculate_psnr(big[i % 2][1], big[i % 2][0], 1.0), 8 * e4)
    row("ssim_linear (C3 shape)", lambda i: linear.ssim_linear(T[i % 4], B[i % 4]), 8 * e3)
    row("ssim_linear replicate pad, k=7", lambda i: linear.ssim_linear(T[i % 4], B[i % 4], kernel_size=7, sigma=1.0,
                                                                       padding="replicate"), 8 * e3)
    row("rgb_ssim", lambda i: channelwise.rgb_ssim(T[i % 4], B[i % 4]), 8 * e3)
    row("calculate_ssim (C4 pair, torchmetrics policy)", lambda i: ssim.calculate_ssim(big[i % 2][1], big[i % 2][0], 1.0), 8 * e4)
    ev_y = ssim.SSIMEvaluator(color_space="y", data_range=1.0)
    row("SSIMEvaluator color_space='y' (C4 pair, luma)", lambda i: ev_y(big[i % 2][1], big[i % 2][0]), 8 * e4, iters=6)
    row("deltaE2000_map (C4 pair)", lambda i: color_error.deltaE2000_map(big[i % 2][0], big[i % 2][1]), 28 * e4 // 3)
    row("deltaE2000_summary p50/p95 (C4 pair)", lambda i: color_error.deltaE2000_summary(big[i % 2][0], big[i % 2][1]), 28 * e4 // 3, iters=6)
    row("edge_deltaE2000_stats (C4 pair)", lambda i: color_error.edge_deltaE2000_stats(big[i % 2][0], big[i % 2][1]), 28 * e4 // 3, iters=6)
    grid = TileGrid(2848, 4256, 512, 512)
    outs = torch.rand(len(grid.indices), 3, 512, 512, device=dev)
    row(f"TileGrid.merge ({len(grid.indices)} tiles of 512^2 -> 2848x4256)", lambda i: grid.merge(outs), 4 * outs.numel() + 4 * e4 // 1)
psf_mod = create_crosstalk_psf("rgb", "B2").to(dev)
row("DeltaE00Loss fwd+bwd (C3)", lossfb(L.DeltaE00Loss()), 36 * e3 // 3)
row("SSIMLoss fwd+bwd (C3)", lossfb(L.SSIMLoss()), 44 * e3)
row("PhysicalConsistencyLossSRGB fwd+bwd (C3)", lossfb(L.PhysicalConsistencyLossSRGB(psf_mod), 1.0), 12 * e3)
row("basicsr L1Loss fwd+bwd", lossfb(bl.L1Loss()), 12 * e3)
row("basicsr MSELoss fwd+bwd", lossfb(bl.MSELoss()), 12 * e3)
row("basicsr CharbonnierLoss fwd+bwd", lossfb(bl.CharbonnierLoss()), 12 * e3)
row("basicsr PSNRLoss fwd+bwd", lossfb(bl.PSNRLoss()), 12 * e3)
