"""DSSIM loss forward + backward on one 2848x4256 pair: heuristic segment rows vs explicit values (diagnostic)."""
import torch, bench
from lowlight_image_enhancement_b200 import _native as nat
from lowlight_image_enhancement_b200.NewBP_model.losses import SSIMLoss
dev = torch.device("cuda:0")
nat.ensure_device(dev)
loss = SSIMLoss()
shape = (1, 3, 2848, 4256)
xs = [torch.rand(shape, device=dev, requires_grad=True) for _ in range(3)]
ys = [torch.rand(shape, device=dev) for _ in range(3)]
def both(i):
    xs[i % 3].grad = None
    loss(xs[i % 3], ys[i % 3]).backward()
for rows in (0, 168, 84, 112, 0, 168, 84):
    nat.lib().llie_tune(4, rows)
    tf = bench.timed_warm(lambda i: loss(xs[i % 3], ys[i % 3]), 30, dev) / 30
    tb = bench.timed_warm(both, 30, dev) / 30
    print(f"seg_rows {rows:4d}: fwd {tf*1e3:.1f} us  fwd+bwd {tb*1e3:.1f} us")
nat.lib().llie_tune(4, 0)
