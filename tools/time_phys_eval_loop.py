"""Dataset-level phys-consistency loop (eval.evaluate_phys_consistency) with and without the per-call stream-overlap
promise, through the public API: ms per batch incl. Python launch overhead, for C2-sized and 4 x C2-sized batches."""
import time
import torch
from lowlight_image_enhancement_b200.eval import evaluate_phys_consistency
from oracle import psf_ops

dev = torch.device("cuda:0")
psf = psf_ops.canonical_kernels("rgb").to(dev)
for n in (8, 32, 128):
    pool = []
    for i in range(8 if n <= 32 else 3):
        pred = torch.rand(n, 3, 512, 512, device=dev)
        rho = torch.rand(n, device=dev) + 0.5
        pool.append((pred, (pred * 0.9).contiguous(), rho))
    batches = [pool[i % len(pool)] for i in range(96)]
    for overlap in (False, True):
        evaluate_phys_consistency(batches, psf, overlap=overlap)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            r = evaluate_phys_consistency(batches, psf, overlap=overlap)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / 3 / len(batches)
        px = n * 512 * 512
        print(f"batch {n:3d}x3x512x512 overlap={overlap!s:5s}: {dt*1e6:7.1f} us per batch = {px/dt/1e9:7.1f} Gpix/s "
              f"({8*3*px/dt/1e9:6.0f} GB/s algorithmic), mean {r['phys_cons_mean']:.6f}")
    del pool, batches
    torch.cuda.empty_cache()
