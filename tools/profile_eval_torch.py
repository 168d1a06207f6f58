"""torch.profiler view of image_metrics (4 eval metrics, or the 7 of configs[3] with --c4) on one full-resolution pair."""
import torch
from torch.profiler import ProfilerActivity, profile

from lowlight_image_enhancement_b200.eval import image_metrics

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(4000)
gt = torch.round(255 * torch.rand(1, 3, 2848, 4256, device=dev, generator=g)) / 255
gt = torch.round(255 * torch.nn.functional.avg_pool2d(gt, 5, 1, 2)) / 255
pred = (gt + 0.03 * torch.randn(gt.shape, device=dev, generator=g)).clamp(0, 1)
import sys
from lowlight_image_enhancement_b200.eval import C4_METRICS
names = list(C4_METRICS) if "--c4" in sys.argv else ["linear_psnr", "linear_ssim", "deltae2000_mean", "edge_deltae2000_mean"]
print(names)
for _ in range(3):
    image_metrics(pred, gt, names)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(5):
        image_metrics(pred, gt, names)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
tot = {}
for e in ev:
    tot.setdefault(e.name[:80], [0, 0.0])
    tot[e.name[:80]][0] += 1
    tot[e.name[:80]][1] += e.device_time
busy = sum(v[1] for v in tot.values())
t0 = min(e.time_range.start for e in ev)
t1 = max(e.time_range.end for e in ev)
print(f"span {(t1 - t0) / 5:.1f} us/image, busy {busy / 5:.1f} us/image")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[1] / 5:9.1f} us  x{v[0] / 5:.0f}  {k}")
