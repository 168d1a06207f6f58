"""Time the decode / sRGB kernels of the loaded library build on a 2848x4256 frame (A/B variants)."""
import torch, bench
from lowlight_image_enhancement_b200 import _functional as Fn
from lowlight_image_enhancement_b200.metrics.evaluation import srgb_and_lpips_inputs
dev = torch.device("cuda:0")
frames = [torch.randint(0, 65536, (1, 2848, 4256, 3), device=dev, dtype=torch.int32).to(torch.uint16) for _ in range(3)]
lin = [torch.rand(1, 3, 2848, 4256, device=dev) for _ in range(3)]
t1 = bench.timed_warm(lambda i: Fn.decode_u16(frames[i % 3], 100.0), 60, dev) / 60
t2 = bench.timed_warm(lambda i: Fn.decode_u16(frames[i % 3], 1.0, want_aligned=False), 60, dev) / 60
t3 = bench.timed_warm(lambda i: srgb_and_lpips_inputs(lin[i % 3]), 60, dev) / 60
px = 2848 * 4256 * 3
print(f"decode raw+aligned {t1*1e3:.1f} us ({px*10/t1/1e6/6553:.3f})  raw {t2*1e3:.1f} us ({px*6/t2/1e6/6553:.3f})  "
      f"srgb+lpips {t3*1e3:.1f} us ({px*12/t3/1e6/6553:.3f})")
