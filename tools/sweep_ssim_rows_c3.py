"""Sweep LLIE_TUNE_SSIM_SEG_ROWS for the DSSIM loss forward + backward at 16x3x512x512 and 8x3x512x512 (diagnostic)."""
import torch, bench
from lowlight_image_enhancement_b200 import _native as nat
from lowlight_image_enhancement_b200.NewBP_model.losses import SSIMLoss
dev = torch.device("cuda:0")
nat.ensure_device(dev)
loss = SSIMLoss()
for shape in ((16, 3, 512, 512), (4, 3, 1024, 1024)):
    xs = [torch.rand(shape, device=dev, requires_grad=True) for _ in range(3)]
    ys = [torch.rand(shape, device=dev) for _ in range(3)]
    def fwd(i):
        return loss(xs[i % 3], ys[i % 3])
    def both(i):
        xs[i % 3].grad = None
        loss(xs[i % 3], ys[i % 3]).backward()
    for rows in (0, 128, 103, 86, 64, 52, 43, 128, 0):
        nat.lib().llie_tune(4, rows)
        tf = bench.timed_warm(fwd, 40, dev) / 40
        tb = bench.timed_warm(both, 40, dev) / 40
        print(f"{shape} seg_rows {rows:4d}: fwd {tf*1e3:.1f} us  fwd+bwd {tb*1e3:.1f} us")
nat.lib().llie_tune(4, 0)
