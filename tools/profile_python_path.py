"""Where the host time of one mirror call goes (cProfile over 300 calls, kernels are asynchronous): validation, tensor
allocation, ctypes marshalling, result assembly."""
import cProfile, pstats, sys
import torch
from lowlight_image_enhancement_b200.metrics import linear, phys_consistency
from oracle import psf_ops

dev = torch.device("cuda:0")
B = torch.rand(16, 3, 512, 512, device=dev)
T = (B + 0.05 * torch.randn_like(B)).clamp(0, 1)
kern = psf_ops.canonical_kernels("rgb").to(dev)
rho = torch.rand(16, device=dev) + 0.5
which = sys.argv[1] if len(sys.argv) > 1 else "phys"
fn = {"phys": lambda: phys_consistency.phys_cons_raw(T, B, kern, rho),
      "psnr": lambda: linear.psnr_linear(T, B),
      "ssim": lambda: linear.ssim_linear(T, B)}[which]
with torch.no_grad():
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(300):
        fn()
    pr.disable()
    torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(22)
