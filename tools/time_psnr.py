"""Time psnr_stats on a full-size pair with the loaded library build (A/B variants)."""
import torch, bench
from lowlight_image_enhancement_b200 import _functional as Fn
dev = torch.device("cuda:0")
shape = (1, 3, 2848, 4256)
pool = [(torch.rand(shape, device=dev), torch.rand(shape, device=dev)) for _ in range(3)]
n, c, h, w = shape
t = bench.timed_warm(lambda i: Fn.psnr_stats(pool[i % 3][1], pool[i % 3][0], n, c, h * w, cast_first=True, clamp_hi=None,
                                             data_range=1.0, eps=1e-12), 60, dev) / 60
print(f"psnr_stats {t*1e3:.1f} us  {2*4*n*c*h*w/t/1e6/6553:.3f} of peak")
