"""Timing of the generic physics-consistency kernels (csrc/phys.cu): the shapes the TMA row ring does not take -- odd
widths, 5x5 and dense cross-channel PSFs -- at 16x3x512xW, forward metric and forward + backward, us per call."""
import sys
import torch, bench
from lowlight_image_enhancement_b200.metrics import phys_consistency
from oracle import psf_ops

dev = torch.device("cuda:0")
peak, _ = bench.peaks()
k3 = psf_ops.canonical_kernels("rgb").to(dev)
rho = torch.rand(16, device=dev) + 0.5
cases = [("dw 3x3, W=512 (ring)", 512, k3), ("dw 3x3, W=510", 510, k3), ("dw 3x3, W=511", 511, k3),
         ("dense 3x3x3x3, W=512", 512, psf_ops.embed_dense(k3.cpu()).to(dev) + 0.01),
         ("dw 5x5, W=512", 512, torch.ones(3, 1, 5, 5, device=dev)), ("dw 7x7, W=512", 512, torch.ones(3, 1, 7, 7, device=dev))]
only = sys.argv[1] if len(sys.argv) > 1 else ""
for name, w, k in cases:
    if only not in name:
        continue
    B = [torch.rand(16, 3, 512, w, device=dev) for _ in range(4)]
    T = [(b + 0.05 * torch.randn_like(b)).clamp(0, 1) for b in B]
    e = B[0].numel()

    def fwd(i):
        with torch.no_grad():
            phys_consistency.phys_cons_raw(T[i % 4], B[i % 4], k, rho)

    def fb(i):
        x = T[i % 4].detach().requires_grad_(True)
        phys_consistency.phys_cons_raw(x, B[i % 4], k, rho, differentiable=True).backward()

    tf = bench.timed_warm(fwd, 12, dev) / 12
    tb = bench.timed_warm(fb, 12, dev) / 12
    print(f"{name:28s} fwd {tf*1e3:8.1f} us ({8*e/tf/1e6/peak:.2f} of HBM)   fwd+bwd {tb*1e3:8.1f} us", flush=True)
