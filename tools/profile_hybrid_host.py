"""Host-side view of one HybridLossPlus step at C3: wall / CUDA-event time per step and a cProfile of 50 steps."""
import cProfile
import pstats
import time

import torch

from lowlight_image_enhancement_b200.NewBP_model.losses import HybridLossPlus

dev = torch.device("cuda:0")
n = 16
B = torch.rand(n, 3, 512, 512, device=dev)
xs = [(B + 0.05 * torch.randn_like(B)).clamp(0, 1) for _ in range(4)]
A = (B * 0.9).contiguous()
rho = torch.ones(n, 1, 1, 1, device=dev)
kern = torch.tensor([[.01, .02, .01], [.02, .88, .02], [.01, .02, .01]], device=dev).view(1, 1, 3, 3).repeat(3, 1, 1, 1)
hyb = HybridLossPlus(device="cuda", w_phys=0.05, physics_kernel=kern, check_finite=False, perceptual_module=False)


def step(i):
    x = xs[i % 4].detach().requires_grad_(True)
    tot, _ = hyb(Bhat_raw=x, B_raw=B, A_raw=A, expo_ratio=rho, Bhat_srgb01=x, B_srgb01=B)
    tot.backward()


for i in range(10):
    step(i)
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
s.record()
for i in range(50):
    step(i)
t_host = time.perf_counter() - t0   # host time to ENQUEUE 50 steps
e.record()
torch.cuda.synchronize()
print(f"host enqueue {t_host / 50 * 1e6:.1f} us/step, device (events) {s.elapsed_time(e) / 50 * 1e3:.1f} us/step")
pr = cProfile.Profile()
pr.enable()
for i in range(50):
    step(i)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
