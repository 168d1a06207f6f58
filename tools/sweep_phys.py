"""A/B sweep of the phys-consistency kernels (diagnostic; not part of the product or the bench contract).
Times fwd / bwd with CUDA graphs of 16 launches over rotating buffer sets (> L2)."""
import sys, os, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from lowlight_image_enhancement_b200 import _functional as Fn, _native as nat

dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
shape = tuple(int(v) for v in (sys.argv[1].split("x") if len(sys.argv) > 1 else "8x3x512x512".split("x")))
POOL = max(2, int(700e6 // (3 * 4 * shape[0] * shape[1] * shape[2] * shape[3])) + 1)
POOL = min(POOL, 16)
n, c, h, w = shape
sets = [(torch.rand(shape, device=dev), torch.rand(shape, device=dev), torch.full((n,), 1.5, device=dev)) for _ in range(POOL)]
grads = [torch.empty(shape, device=dev) for _ in range(POOL)]
kern = torch.tensor([[0.01, 0.02, 0.01], [0.02, 0.88, 0.02], [0.01, 0.02, 0.01]]).view(1, 1, 3, 3).repeat(c, 1, 1, 1).to(dev)
psf = Fn.prepare_psf(kern, c, c, dev, normalize=True, nonneg=False, eps=1e-12)
cfg = Fn.PhysCfg(nat.PAD["reflect"], True, False, 1e-12, 0, False, False)
gout = torch.ones(1, device=dev)
st = (1, 0, 0, 0)
L = 16

def fwd(i):
    p, o, r = sets[i % POOL]
    Fn.phys_forward(p, o, psf, r, 0.0, st, cfg, False)

def bwd(i):
    p, o, r = sets[i % POOL]
    Fn.phys_backward(p, o, psf, r, 0.0, st, cfg, gout, 1e-6, out=grads[i % POOL])

def graph_time(fn):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(3):
            fn(i)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for i in range(L):
                fn(i)
        g.replay(); torch.cuda.synchronize()
        best = 1e9
        for rep in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(s); g.replay(); b.record(s); torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b) / L)
    return best * 1e3  # us

elems = n * c * h * w
res = []
rows_list = [int(v) for v in os.environ.get("SWEEP_ROWS", "0,8,12,16,24,32,64").split(",")]
configs = ([("generic", 2, 0)] if os.environ.get("SWEEP_GENERIC") else []) + [("ring", 0, r) for r in rows_list]
print("lib:", nat.LIB_PATH, "shape:", shape)
for name, path, rows in configs:
    nat.lib().llie_tune(1, path); nat.lib().llie_tune(2, rows)
    try:
        tf, tb = graph_time(fwd), graph_time(bwd)
    except Exception as e:
        print(name, rows, "failed", e); continue
    print(f"{name:5s} rows={rows:3d} fwd {tf:8.2f} us {8*elems/tf/1e3:8.1f} GB/s | bwd {tb:8.2f} us {12*elems/tb/1e3:8.1f} GB/s", flush=True)
nat.lib().llie_tune(1, 0); nat.lib().llie_tune(2, 0)
