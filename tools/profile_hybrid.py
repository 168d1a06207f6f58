"""torch.profiler view of one HybridLossPlus step at C3 (kernel list with device times, total vs busy)."""
import torch
from torch.profiler import ProfilerActivity, profile

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


for i in range(5):
    step(i)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for i in range(5):
        step(i)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
tot = {}
for e in ev:
    tot.setdefault(e.name[:70], [0, 0.0])
    tot[e.name[:70]][0] += 1
    tot[e.name[:70]][1] += e.device_time
busy = sum(v[1] for v in tot.values())
t0 = min(e.time_range.start for e in ev)
t1 = max(e.time_range.end for e in ev)
print(f"span {(t1 - t0) / 5:.1f} us/step, busy {busy / 5:.1f} us/step")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[1] / 5:9.1f} us  x{v[0] / 5:.0f}  {k}")
