"""Host-side view of the sharded eval driver at C4: host time to ENQUEUE an image vs device time per image, and a
cProfile of the enqueue loop (is the driver GPU-bound?)."""
import cProfile
import pstats
import time

import torch

from lowlight_image_enhancement_b200.eval import ShardedMeans, accumulate_image_metrics

dev = torch.device("cuda:0")
H, W = 2848, 4256
g = torch.Generator(device=dev).manual_seed(4000)
pool = []
for _ in range(3):
    gt = torch.round(255 * torch.rand(1, 3, H, W, device=dev, generator=g)) / 255
    pool.append(((gt + 0.03 * torch.randn(gt.shape, device=dev, generator=g)).clamp(0, 1), gt))
names = ["linear_psnr", "linear_ssim", "deltae2000_mean", "edge_deltae2000_mean"]
acc = ShardedMeans(names, dev)
for i in range(6):
    accumulate_image_metrics(*pool[i % 3], names, acc)
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
s.record()
for i in range(24):
    accumulate_image_metrics(*pool[i % 3], names, acc)
t_host = time.perf_counter() - t0
e.record()
torch.cuda.synchronize()
print(f"host enqueue {t_host / 24 * 1e6:.1f} us/image, device (events) {s.elapsed_time(e) / 24 * 1e3:.1f} us/image")
pr = cProfile.Profile()
pr.enable()
for i in range(24):
    accumulate_image_metrics(*pool[i % 3], names, acc)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
