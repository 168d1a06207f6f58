"""Reference points for the HBM roofline on this box: torch copy / add / sum and our PSNR streaming kernel,
timed with CUDA graphs of 16 launches over rotating buffers (diagnostic)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from lowlight_image_enhancement_b200 import _functional as Fn, _native as nat
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
shape = tuple(int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "8x3x512x512").split("x"))
n, c, h, w = shape
elems = n * c * h * w
POOL = min(16, max(2, int(700e6 // (12 * elems)) + 1))
xs = [torch.rand(shape, device=dev) for _ in range(POOL)]
ys = [torch.rand(shape, device=dev) for _ in range(POOL)]
zs = [torch.empty(shape, device=dev) for _ in range(POOL)]
L = 16
def graph_time(fn):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(3): fn(i)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for i in range(L): fn(i)
        g.replay(); torch.cuda.synchronize()
        best = 1e9
        for rep in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(s); g.replay(); b.record(s); torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b) / L)
    return best * 1e3
tests = {
  "torch copy (8 B/el)": (lambda i: zs[i % POOL].copy_(xs[i % POOL]), 8),
  "torch add  (12 B/el)": (lambda i: torch.add(xs[i % POOL], ys[i % POOL], out=zs[i % POOL]), 12),
  "torch sum x (4 B/el)": (lambda i: xs[i % POOL].sum(), 4),
  "torch (x-y).abs().sum (eager 3 kernels)": (lambda i: (xs[i % POOL] - ys[i % POOL]).abs().sum(), 8),
  "llie psnr_stats (8 B/el)": (lambda i: Fn.psnr_stats(xs[i % POOL], ys[i % POOL], n, c, h * w, cast_first=False, clamp_hi=None, data_range=1.0, eps=1e-12), 8),
}
print("shape", shape, "pool", POOL)
for name, (fn, bpe) in tests.items():
    t = graph_time(fn)
    print(f"{name:42s} {t:8.2f} us  {bpe*elems/t/1e3:8.1f} GB/s", flush=True)
