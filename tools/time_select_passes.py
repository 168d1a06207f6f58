import torch
from torch.profiler import ProfilerActivity, profile
from lowlight_image_enhancement_b200.eval import image_metrics
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(4000)
gt = torch.round(255 * torch.rand(1, 3, 2848, 4256, device=dev, generator=g)) / 255
gt = torch.round(255 * torch.nn.functional.avg_pool2d(gt, 5, 1, 2)) / 255
pred = (gt + 0.03 * torch.randn(gt.shape, device=dev, generator=g)).clamp(0, 1)
names = ["deltae2000_mean", "edge_deltae2000_mean", "deltae2000_p95"]
for _ in range(3):
    image_metrics(pred, gt, names)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    image_metrics(pred, gt, names)
    torch.cuda.synchronize()
ev = sorted([e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA], key=lambda e: e.time_range.start)
for e in ev:
    if "llie" in e.name:
        print(f"{e.device_time:8.1f} us  {e.name[:70]}")
