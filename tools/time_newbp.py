"""NewBP operator (llie_newbp_forward / llie_newbp_adjoint, csrc/conv.cu) at training shapes: us per launch and the
fraction of the measured HBM peak (8 B/element: one read, one write)."""
import torch, bench
from lowlight_image_enhancement_b200.NewBP_model.newbp_layer import NewBPFunction, create_crosstalk_psf

dev = torch.device("cuda:0")
peak, _ = bench.peaks()
for shape, dt in (((16, 3, 512, 512), torch.float32), ((32, 3, 512, 512), torch.float32), ((16, 3, 512, 512), torch.bfloat16)):
    xs = [torch.rand(shape, device=dev, dtype=dt).requires_grad_(True) for _ in range(6)]
    k = create_crosstalk_psf("rgb", "B2").to(dev).kernel.to(dt)

    def fwd(i):
        with torch.no_grad():
            return NewBPFunction.apply(xs[i % 6], k, 1, 3)

    def fb(i):
        x = xs[i % 6]
        y = NewBPFunction.apply(x, k, 1, 3)
        y.backward(y.detach())
        x.grad = None

    t_f = bench.timed_warm(fwd, 30, dev) / 30
    t_fb = bench.timed_warm(fb, 30, dev) / 30
    el = 1
    for v in shape:
        el *= v
    b = 2 * el * (4 if dt == torch.float32 else 2)
    print(f"{shape} {dt}: forward {t_f*1e3:.1f} us ({b/t_f/1e6/peak:.2f} of HBM), forward + adjoint {t_fb*1e3:.1f} us")
