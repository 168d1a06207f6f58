"""Time the Sobel magnitude kernel of the loaded library build on a 2848x4256 plane (A/B variants)."""
import torch, bench
from lowlight_image_enhancement_b200 import _functional as Fn
dev = torch.device("cuda:0")
L = [100 * torch.rand(1, 1, 2848, 4256, device=dev) for _ in range(4)]
t = bench.timed_warm(lambda i: Fn.sobel_magnitude(L[i % 4]), 80, dev) / 80
print(f"sobel {t*1e3:.1f} us")
