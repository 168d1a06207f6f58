"""Time the L1 value + gradient kernel and the BasicSR pixel losses of the loaded library build (A/B variants)."""
import torch, bench
from lowlight_image_enhancement_b200 import _functional as Fn
from lowlight_image_enhancement_b200 import basicsr_losses as B
dev = torch.device("cuda:0")
for shape in ((16, 3, 512, 512), (1, 3, 2848, 4256)):
    xs = [torch.rand(shape, device=dev, requires_grad=True) for _ in range(3)]
    ys = [torch.rand(shape, device=dev) for _ in range(3)]
    n = xs[0].numel()
    out = []
    for name, fn in (("L1MeanFn", lambda a, b: Fn.L1MeanFn.apply(a, b)), ("L1Loss", B.L1Loss()), ("MSELoss", B.MSELoss()),
                     ("Charbonnier", B.CharbonnierLoss())):
        t = bench.timed_warm(lambda i: fn(xs[i % 3], ys[i % 3]), 60, dev) / 60
        out.append(f"{name} {t*1e3:.1f} us ({n*12/t/1e6/6553:.2f})")
    print(shape, "  ".join(out))
