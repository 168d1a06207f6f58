#!/usr/bin/env python
"""bench.py — throughput of the B200-native loss/metric hot path (contract: see DESIGN.md §Measurement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload phys_c2]

A *step* is one pass of the hot path over one synthetic batch of BASELINE.json configs[1]:
`phys_cons_raw` (rgb B2 depthwise PSF, per-image rho in {100,250,300}, reflect pad, valid crop, L1)
forward + the adjoint backward on 8x3x512x512 fp32.  value = Mpix/s (pixels = N*H*W) with inputs
resident in HBM; e2e = the same through the public Python API from pinned HOST buffers (H2D of
pred/obs and D2H of the loss inside the timed region).  One JSON line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

RHO = [100.0, 250.0, 300.0, 100.0, 250.0, 300.0, 100.0, 250.0]
SHAPE = (8, 3, 512, 512)
# dram bytes per launch from the committed `ncu --set full` capture of this workload (profiles/); None = not captured
# dram__bytes_read.sum + dram__bytes_write.sum per launch, `ncu --set full` of this command
# (profiles/r1_bench_phys_c2_ncu_full.txt, r1_bench_phys_c2_fused_ncu_full.txt).  Reads equal the algorithmic 50.33 MB (no halo re-reads reach
# DRAM); the 25.2 MB gradient written by the backward kernel is still held by the 126 MB L2 when the kernel
# ends, so only ~1.3 MB of its write-back is attributed to the launch.
TRAFFIC = {"phys_fwd_ring": 50.40e6, "phys_bwd_ring": 51.78e6, "phys_fused_ring": 51.96e6}
POOL = 8  # rotating buffer sets: 8 x (pred+obs+grad = 75.5 MB) = 604 MB >> 126 MB L2


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def synth_batch(seed, device):
    """pred = U[0,1)/rho_n, obs = clamp(rho_n*(B2 (*) pred) + 0.01 N, 0, 1)."""
    from oracle import psf_ops  # input synthesis only (not timed, not the product path)
    g = torch.Generator().manual_seed(seed)
    rho = torch.tensor(RHO)
    pred = torch.rand(SHAPE, generator=g) / rho.view(-1, 1, 1, 1)
    k = psf_ops.normalise_crosstalk(psf_ops.canonical_kernels("rgb"))
    obs = (rho.view(-1, 1, 1, 1) * psf_ops.crosstalk_forward(pred, k) + 0.01 * torch.randn(SHAPE, generator=g)).clamp(0, 1)
    return pred.to(device), obs.to(device), rho.to(device)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop = index, [], threading.Event()
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(timeout=6)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower() == "active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's algorithm (oracle restatement of metrics/phys_consistency.py, same torch ops)
# ------------------------------------------------------------------------------------------------
def cpu_step_fn():
    from oracle import psf_ops
    dense = psf_ops.embed_dense(psf_ops.canonical_kernels("rgb"))

    def step(pred, obs, rho):
        x = pred.detach().clone().requires_grad_(True)
        loss = psf_ops.phys_metric(x, obs, dense, rho, keep_graph=True)
        loss.backward()
        return float(loss.detach())
    return step


def time_cpu(steps, warmup):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sets = [synth_batch(102 + i, "cpu") for i in range(2)]
    step = cpu_step_fn()
    for i in range(warmup):
        step(*sets[i % 2])
    t0 = time.perf_counter()
    for i in range(steps):
        step(*sets[i % 2])
    dt = time.perf_counter() - t0
    mpix = SHAPE[0] * SHAPE[2] * SHAPE[3] / 1e6
    return mpix * steps / dt, dt / steps * 1e3, cores


def run_reference(args, rank):
    if rank != 0:
        return None
    steps = max(1, args.steps)
    val, ms, cores = time_cpu(steps, max(1, min(args.warmup, 3)))
    sample = f"{steps} full C2 batches (8x3x512x512) fwd+bwd, oracle port of phys_cons_raw, torch CPU {cores} threads"
    return {
        "impl": "reference", "metric": "Mpix/s phys-cons fwd+bwd", "value": round(val, 3), "unit": "Mpix/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": round(ms, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "phys_cons_raw rgb B2 depthwise PSF, per-image rho, 8x3x512x512 fp32, fwd+bwd (BASELINE configs[1])"},
        "cpu_baseline": {"value": round(val, 3), "unit": "Mpix/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(val, 3), "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def timed(fn, iters, dev):
    """CUDA-event time (ms) of `iters` calls of fn(i) on the current stream, synchronised both sides."""
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    s.record()
    for i in range(iters):
        fn(i)
    e.record()
    torch.cuda.synchronize(dev)
    return s.elapsed_time(e)


def timed_warm(fn, iters, dev, warm=3):
    """`timed` after `warm` untimed calls (first calls create the per-stream workspace and grow the caching allocator)."""
    for i in range(warm):
        fn(i)
    return timed(fn, iters, dev)


class GraphLoop:
    """`steps` calls of fn(i) captured into CUDA graphs of <= 64 calls each (the inner loop is launch-bound
    from Python: one ctypes call costs more than the 17 us kernel) and replayed on a side stream."""

    def __init__(self, fn, steps, dev):
        self.dev, self.stream = dev, torch.cuda.Stream(dev)
        g = max(d for d in range(1, 65) if steps % d == 0)
        self.replays = steps // g
        with torch.cuda.stream(self.stream):
            for i in range(3):
                fn(i)
            torch.cuda.synchronize(dev)
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph, stream=self.stream):
                for i in range(g):
                    fn(i)
        torch.cuda.synchronize(dev)

    def run(self):
        """-> elapsed ms of all `steps` calls (CUDA events on the launching stream, synchronised both sides)."""
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(self.dev)
        with torch.cuda.stream(self.stream):
            s.record(self.stream)
            for _ in range(self.replays):
                self.graph.replay()
            e.record(self.stream)
        torch.cuda.synchronize(self.dev)
        return s.elapsed_time(e)


def run_ours(args, rank, world, dev):
    import torch.distributed as dist
    from lowlight_image_enhancement_b200 import _functional as Fn
    from lowlight_image_enhancement_b200 import _native as nat
    from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw
    from oracle import psf_ops  # kernel constants for input synthesis only

    if args.no_pdl:
        nat.check(nat.lib().llie_tune(5, 0), "llie_tune")  # LLIE_TUNE_PHYS_PDL
    if args.ring_rows:
        nat.check(nat.lib().llie_tune(2, args.ring_rows), "llie_tune")  # LLIE_TUNE_PHYS_ROWS (A/B)
    sets = [synth_batch(1000 * rank + 102 + i, dev) for i in range(POOL)]
    grads = [torch.empty(SHAPE, device=dev) for _ in range(POOL)]
    kern = psf_ops.canonical_kernels("rgb").to(dev)
    psf = Fn.prepare_psf(kern, 3, 3, dev, normalize=True, nonneg=False, eps=1e-12)
    cfg = Fn.PhysCfg(nat.PAD["reflect"], True, False, 1e-12, 0, False, False)
    n, c, h, w = SHAPE
    count = c * (h - 2) * (w - 2)
    gout = torch.ones(1, device=dev)
    gscale = 1.0 / (n * count)
    st = (1, 0, 0, 0)

    def fwd(i):
        p, o, r = sets[i % POOL]
        return Fn.phys_forward(p, o, psf, r, 0.0, st, cfg, False)

    def bwd(i):
        p, o, r = sets[i % POOL]
        return Fn.phys_backward(p, o, psf, r, 0.0, st, cfg, gout, gscale, out=grads[i % POOL])

    def fused(i):
        p, o, r = sets[i % POOL]
        return Fn.phys_value_and_grad(p, o, psf, r, 0.0, st, cfg, gout, gscale, out=grads[i % POOL])

    def step2(i):
        fwd(i)
        bwd(i)

    # a step = loss value (per-sample means, batch mean, flags) AND d loss / d pred.  One fused launch
    # (llie_phys_value_and_grad) yields both from a single pass over pred / obs; --two-pass times the separate
    # forward + backward kernels instead (same outputs, 20 instead of 12 bytes per element).
    step = step2 if args.two_pass else fused
    K, W = args.steps, max(3, args.warmup)
    # serialised reference timing first (every launch waits for the previous grid before its first global access);
    # the `early` flag is a kernel argument, so it is frozen into a graph at capture time
    g_su = GraphLoop(fused, K, dev)
    g_su.run()
    t_u_serial = min(g_su.run() for _ in range(3)) / K
    del g_su
    overlap = not (args.no_overlap or args.no_pdl)
    if overlap:
        # the K steps read / write 8 rotating, disjoint buffer sets: exactly the stream-overlap contract of
        # LLIE_TUNE_PHYS_OVERLAP (include/llie_b200.h) -- consecutive launches only share the workspace
        Fn.phys_stream_overlap(True)
    g_step, g_f, g_b = GraphLoop(step, K, dev), GraphLoop(fwd, K, dev), GraphLoop(bwd, K, dev)
    g_2 = GraphLoop(step2, K, dev)
    g_u = GraphLoop(fused, K, dev)
    for _ in range(max(1, W * g_step.replays // K)):
        g_step.run()
    # per-kernel timings (live, CUDA events on the launching stream): K launches of each kernel alone
    t_f = min(g_f.run() for _ in range(3)) / K
    t_b = min(g_b.run() for _ in range(3)) / K
    t_u = min(g_u.run() for _ in range(3)) / K
    t_2 = min(g_2.run() for _ in range(3)) / K
    if world > 1:
        dist.barrier()
    l0 = nat.launches
    with ClockSampler(dev.index) as clk:
        ms = g_step.run()
        time.sleep(0.25)  # let the sampler see the clocks right after the (sub-ms .. ms) timed region
    launches = (2 if args.two_pass else 1) * K  # kernels per step, replayed from the captured graph
    Fn.phys_stream_overlap(False)  # the e2e leg below runs under autograd: default (serialised) launches
    del l0
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        ms = float(t)
    mpix = n * h * w / 1e6
    value = world * mpix * K / (ms / 1e3)

    # ---- e2e: public API, pinned host inputs, H2D + D2H inside the timed region
    hp = [(s[0].cpu().pin_memory(), s[1].cpu().pin_memory()) for s in sets[:4]]
    rho_d = sets[0][2]

    def e2e_step(i):
        ph_, oh_ = hp[i % 4]
        x = ph_.to(dev, non_blocking=True).requires_grad_(True)
        o = oh_.to(dev, non_blocking=True)
        loss = phys_cons_raw(x, o, kern, rho_d, differentiable=True)
        loss.backward()
        return loss.item()

    Ke = max(3, min(K, 50))
    if args.skip_e2e:
        Ke = 1
    for i in range(1 if args.skip_e2e else 3):
        e2e_step(i)
    ms_e = timed(e2e_step, Ke, dev)
    if world > 1:
        t = torch.tensor([ms_e], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e = float(t)
    e2e_val = world * mpix * Ke / (ms_e / 1e3)

    if rank != 0:
        return None
    peak, peak_src = peaks()
    elems = n * c * h * w
    bytes_f, bytes_b = 8 * elems, 12 * elems
    ach_b = bytes_b / (t_b * 1e-3) / 1e9
    ach_f = bytes_f / (t_f * 1e-3) / 1e9
    ach_u = bytes_b / (t_u * 1e-3) / 1e9   # the fused pass moves the backward's 12 B/element
    roof = ({"kernel": "phys_bwd_ring", "achieved": ach_b, "us": t_b} if args.two_pass else
            {"kernel": "phys_bwd_ring<FUSED> (value + gradient)", "achieved": ach_u, "us": t_u})
    line = {
        "metric": "Mpix/s phys-cons fwd+bwd", "value": round(value, 1), "unit": "Mpix/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": round(ms / K, 5), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "phys_cons_raw rgb B2 depthwise PSF, per-image rho, 8x3x512x512 fp32, fwd+bwd (BASELINE configs[1])",
                   "step": ("forward kernel + backward kernel" if args.two_pass else
                            "loss value + gradient from one fused launch (llie_phys_value_and_grad); "
                            f"two-pass forward + backward = {round(t_2 * 1e3, 2)} us per step"),
                   "l2": f"inputs rotate over {POOL} resident buffer sets (604 MB per GPU) > 126 MB L2",
                   "launch": f"steps captured in CUDA graphs of {K // g_step.replays} steps, replayed {g_step.replays}x; "
                             + ("programmatic dependent launch with the stream-overlap contract (LLIE_TUNE_PHYS_OVERLAP: "
                                "consecutive steps touch disjoint buffer sets, so a launch waits for the previous grid "
                                "only before it touches the shared workspace); "
                                f"fully serialised launches: {round(t_u_serial * 1e3, 2)} us per fused launch"
                                if overlap else "serialised launches"),
                   "per_gpu_batch": list(SHAPE)},
        "roofline": {"bound": "hbm", "kernel": roof["kernel"], "achieved": round(roof["achieved"], 1), "peak": peak,
                     "unit": "GB/s", "frac": round(roof["achieved"] / peak, 4),
                     "traffic": TRAFFIC.get("phys_bwd_ring" if args.two_pass else "phys_fused_ring"),
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": bytes_b,
                     "us_per_launch": round(roof["us"] * 1e3, 2),
                     "us_per_launch_serialised": round(t_u_serial * 1e3, 2),
                     "frac_serialised": round(bytes_b / (t_u_serial * 1e-3) / 1e9 / peak, 4)},
        "kernels": [
            {"kernel": "phys_fwd_ring", "us": round(t_f * 1e3, 2), "algorithmic_bytes": bytes_f,
             "achieved_gbs": round(ach_f, 1), "frac": round(ach_f / peak, 4), "traffic": TRAFFIC.get("phys_fwd_ring")},
            {"kernel": "phys_bwd_ring", "us": round(t_b * 1e3, 2), "algorithmic_bytes": bytes_b,
             "achieved_gbs": round(ach_b, 1), "frac": round(ach_b / peak, 4), "traffic": TRAFFIC.get("phys_bwd_ring")},
            {"kernel": "phys_bwd_ring<FUSED>", "us": round(t_u * 1e3, 2), "algorithmic_bytes": bytes_b,
             "achieved_gbs": round(ach_u, 1), "frac": round(ach_u / peak, 4), "traffic": TRAFFIC.get("phys_fused_ring")}],
        "e2e": {"value": round(e2e_val, 1), "unit": "Mpix/s", "h2d_bytes_per_step": 2 * 4 * elems,
                "d2h_bytes_per_step": 4, "steps": Ke, "api": "metrics.phys_consistency.phys_cons_raw(differentiable=True)+backward"},
        "gpu_launches": launches,
        "clocks": clk.summary(),
    }
    if world == 1 and not args.skip_cpu:
        v, ms_c, cores = time_cpu(3, 1)
        line["cpu_baseline"] = {"value": round(v, 3), "unit": "Mpix/s", "cores": cores, "kind": "port",
                                "sample": "3 full C2 batches fwd+bwd after 1 warm-up, oracle port, torch CPU all threads"}
    return line


# ------------------------------------------------------------------------------------------------
# secondary workloads (reported inside the same JSON line): BASELINE configs[3] eval and configs[2] hybrid loss
# ------------------------------------------------------------------------------------------------
EVAL_SHAPE = (1, 3, 2848, 4256)
EVAL_METRICS = ["linear_psnr", "linear_ssim", "deltae2000_mean", "edge_deltae2000_mean"]


def synth_eval_pair(seed, dev, shape=EVAL_SHAPE):
    """gt = 8-bit noise low-passed by a 5x5 box and re-quantised, pred = clamp(gt + 0.03 N)."""
    g = torch.Generator(device=dev).manual_seed(seed)
    gt = torch.round(255 * torch.rand(shape, device=dev, generator=g)) / 255
    gt = torch.round(255 * torch.nn.functional.avg_pool2d(gt, 5, 1, 2)) / 255
    pred = (gt + 0.03 * torch.randn(shape, device=dev, generator=g)).clamp(0, 1)
    return pred.contiguous(), gt.contiguous()


def stressed_kernels(stress, dev):
    """SURVEY.md §8d C3 (bench-defined; the reference names the family but gives no numbers):
    K_c = delta + s * alpha_c * P2_offcentre, alpha = (1.3, 1.0, 0.7), renormalised to sum 1."""
    p2 = torch.tensor([[.01, .02, .01], [.02, 0.0, .02], [.01, .02, .01]])
    ks = []
    for alpha in (1.3, 1.0, 0.7):
        k = stress * alpha * p2
        k[1, 1] = 1.0
        ks.append(k / k.sum())
    return torch.stack(ks).view(3, 1, 3, 3).to(dev)


def run_extra(args, rank, world, dev):
    """-> dict with 'eval_c4' (sharded eval, one all-reduce) and 'hybrid_c3' (HybridLossPlus fwd+bwd)."""
    import torch.distributed as dist
    from lowlight_image_enhancement_b200 import _functional as Fn
    from lowlight_image_enhancement_b200.eval import (ShardedMeans, accumulate_image_metrics, evaluate_pairs,
                                                     image_metrics)
    from lowlight_image_enhancement_b200.metrics.ssim import torchmetrics_ssim_scores
    from lowlight_image_enhancement_b200.NewBP_model.losses import HybridLossPlus
    out = {}
    peak, _ = peaks()

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t)
        return ms

    # ---- eval: images sharded by index over the ranks, per-image metrics accumulated on the device,
    # ONE all-reduce + ONE host sync at the end (lowlight_image_enhancement_b200/eval.py)
    pool = [synth_eval_pair(4000 + 17 * rank + i, dev) for i in range(3)]   # 3 x 291 MB > L2
    per_rank = max(6, min(48, args.steps))
    n_img = per_rank * world
    px = EVAL_SHAPE[2] * EVAL_SHAPE[3]
    getter = lambda i: pool[(i // world) % 3]
    evaluate_pairs(getter, EVAL_METRICS, n_items=3 * world, device=dev)    # warm-up (also NCCL)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    s_ev, e_ev = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s_ev.record()
    res = evaluate_pairs(getter, EVAL_METRICS, n_items=n_img, device=dev)
    e_ev.record()
    torch.cuda.synchronize(dev)
    wall = (time.perf_counter() - t0) * 1e3
    ms = max_over_ranks(s_ev.elapsed_time(e_ev))
    pr, gt = pool[0]
    t_ssim = timed_warm(lambda i: torchmetrics_ssim_scores(pool[i % 3][0], pool[i % 3][1], data_range=1.0, sigma=1.5,
                                                      k1=0.01, k2=0.03), 6, dev) / 6
    t_de = timed_warm(lambda i: Fn.deltae(pool[i % 3][0], pool[i % 3][1], variant=0, clamp_in=True), 6, dev) / 6
    t_ps = timed_warm(lambda i: Fn.psnr_stats(pool[i % 3][1], pool[i % 3][0], 1, 1, 3 * px, cast_first=True, clamp_hi=None,
                                         data_range=1.0, eps=1e-12), 24, dev) / 24
    t_all = timed_warm(lambda i: image_metrics(pool[i % 3][0], pool[i % 3][1], EVAL_METRICS), 6, dev) / 6
    acc_k = ShardedMeans(EVAL_METRICS, dev)
    t_fast = timed_warm(lambda i: accumulate_image_metrics(pool[i % 3][0], pool[i % 3][1], EVAL_METRICS, acc_k), 6,
                        dev) / 6

    def kern(name, t, bpp, bound):
        return {"kernel": name, "ms": round(t, 4), "Mpix/s": round(px / t / 1e3, 1),
                "algorithmic_bytes": bpp * px, "achieved_gbs": round(bpp * px / t / 1e6, 1),
                "frac_hbm": round(bpp * px / t / 1e6 / peak, 4), "bound": bound}
    out["eval_c4"] = {
        "workload": "PSNR + SSIM(torchmetrics policy) + dE00 mean + edge dE00 on 1x3x2848x4256 sRGB pairs (BASELINE configs[3]), "
                    f"{per_rank} images per GPU sharded by index, device-side accumulation, one all-reduce, one host sync",
        "value": round(n_img * px / 1e6 / (ms / 1e3), 1), "unit": "Mpix/s", "images": n_img, "n_gpus": world,
        "ms_per_image_per_gpu": round(ms / per_rank, 4), "wall_ms": round(wall, 2),
        "means": {k: (round(v, 6) if isinstance(v, float) else v) for k, v in res.items()},
        "kernels": [kern("psnr_stats_kernel", t_ps, 24, "hbm"),
                    kern("ssim_stream2_kernel<11, 4 maps>", t_ssim, 24, "instruction issue (88 FMA + ~85 other instructions per element)"),
                    kern("deltae_kernel", t_de, 24, "instruction issue (fp32 + sfu)"),
                    kern("all 4 metrics of one image, image_metrics (generic adapters, torch glue between kernels)", t_all, 36,
                         "mixed"),
                    kern("all 4 metrics of one image, accumulate_image_metrics (SSIM on a side stream, one-launch epilogue)",
                         t_fast, 36, "mixed")],
    }
    if world == 1 and not args.skip_cpu:
        # reference CPU path of the same four adapters on ONE full-resolution pair (bounded sample: ~15-30 s)
        from oracle import colour, quality
        torch.set_num_threads(os.cpu_count() or 1)
        cp, cg = pool[0][0].cpu(), pool[0][1].cpu()
        t0 = time.perf_counter()
        quality.calculate_psnr(cg, cp, 1.0)
        float(quality.ssim_torchmetrics(cp, cg).mean())
        colour.de00_summary(cp, cg, percentiles=(95.0,))          # deltae2000_mean pays the p95 sort in the reference
        colour.edge_stats(cp, cg, q=0.85)
        dt = time.perf_counter() - t0
        out["eval_c4"]["cpu_baseline"] = {"value": round(px / 1e6 / dt, 3), "unit": "Mpix/s", "cores": os.cpu_count(),
                                          "kind": "port", "sample": f"1 full-resolution pair, the same 4 adapters, {dt:.1f} s"}
        del cp, cg
    del pool, pr, gt
    torch.cuda.empty_cache()

    # ---- decode / normalise step in front of the path: SID-sized 16-bit frames -> planar fp32,
    # and the linear -> sRGB (+ LPIPS range) map of the evaluation helpers.  HBM-bound streaming kernels.
    from lowlight_image_enhancement_b200.metrics.evaluation import srgb_and_lpips_inputs
    h_, w_ = EVAL_SHAPE[2], EVAL_SHAPE[3]
    gi = torch.Generator(device=dev).manual_seed(5000 + rank)
    frames = [(torch.randint(0, 65536, (1, h_, w_, 3), generator=gi, dtype=torch.int32, device=dev) // 64).to(torch.uint16)
              for _ in range(3)]                                                        # 3 x 72.7 MB in, 291 MB out each
    lin = [torch.rand(EVAL_SHAPE, device=dev, generator=gi) for _ in range(3)]
    el = 3 * px
    for _ in range(2):  # warm-up (allocator + first launch)
        Fn.decode_u16(frames[0], 100.0), Fn.decode_u16(frames[0], 1.0, want_aligned=False), srgb_and_lpips_inputs(lin[0])
    t_dec = timed_warm(lambda i: Fn.decode_u16(frames[i % 3], 100.0), 27, dev) / 27
    t_dec1 = timed_warm(lambda i: Fn.decode_u16(frames[i % 3], 1.0, want_aligned=False), 27, dev) / 27
    t_srgb = timed_warm(lambda i: srgb_and_lpips_inputs(lin[i % 3]), 27, dev) / 27

    def kern_b(name, t, bpe):
        return {"kernel": name, "ms": round(t, 4), "Mpix/s": round(px / t / 1e3, 1), "algorithmic_bytes": bpe * el,
                "achieved_gbs": round(bpe * el / t / 1e6, 1), "frac_hbm": round(bpe * el / t / 1e6 / peak, 4),
                "bound": "hbm"}
    out["ingest_c4"] = {
        "workload": "decode of one 2848x4256x3 uint16 frame to planar fp32 (raw + exposure-aligned / raw only) and "
                    "linear -> sRGB + LPIPS-range map of one 1x3x2848x4256 image; outputs allocated per call",
        "kernels": [kern_b("decode_u16_c3_kernel (raw + aligned)", t_dec, 10),
                    kern_b("decode_u16_c3_kernel (raw)", t_dec1, 6),
                    kern_b("linear_to_srgb_kernel (srgb + lpips)", t_srgb, 12)],
    }
    if world == 1 and not args.skip_cpu:
        from oracle import ingest as oing
        torch.set_num_threads(os.cpu_count() or 1)
        f_np = frames[0][0].cpu().numpy()
        x_cpu = lin[0].cpu()
        t0 = time.perf_counter()
        oing.decode_pair(f_np, f_np, 100.0)
        dt_d = time.perf_counter() - t0
        t0 = time.perf_counter()
        oing.norm_to_lpips_range(oing.linear_to_srgb(x_cpu))
        dt_s = time.perf_counter() - t0
        out["ingest_c4"]["cpu_baseline"] = {
            "decode_pair_Mpix/s": round(2 * px / 1e6 / dt_d, 2), "linear_to_srgb_Mpix/s": round(px / 1e6 / dt_s, 2),
            "unit": "Mpix/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"one short + long frame decode ({dt_d:.2f} s, numpy is single-threaded) and one sRGB map ({dt_s:.2f} s)"}
        del f_np, x_cpu
    del frames, lin
    torch.cuda.empty_cache()

    # ---- hybrid loss, C3: L1 + 0.05 DSSIM + 0.02 dE00 + 0.05 phys (stressed rgb kernel, RAW convention)
    n = 16
    g = torch.Generator(device=dev).manual_seed(103 + rank)
    B = torch.rand(n, 3, 512, 512, device=dev, generator=g)
    kern3 = stressed_kernels(2.0, dev)
    A = (torch.nn.functional.conv2d(torch.nn.functional.pad(B, (1, 1, 1, 1), mode="replicate"), kern3, groups=3)
         + 0.02 * torch.randn(B.shape, device=dev, generator=g)).clamp(0, 1)
    xs = [(B + 0.05 * torch.randn(B.shape, device=dev, generator=g)).clamp(0, 1) for _ in range(4)]   # 4 x 50 MB + grads
    rho = torch.ones(n, 1, 1, 1, device=dev)
    hyb = HybridLossPlus(device=dev, w_l1_raw=1.0, w_perc=0.0, w_deltaE=0.02, w_ssim=0.05, w_phys=0.05,
                         physics_kernel=kern3, check_finite=False)

    def hstep(i):
        x = xs[i % 4].detach().requires_grad_(True)
        tot, _ = hyb(Bhat_raw=x, B_raw=B, A_raw=A, expo_ratio=rho, Bhat_srgb01=x, B_srgb01=B)
        tot.backward()
        return tot

    for i in range(3):
        hstep(i)
    kh = max(6, min(50, args.steps))
    ms_h = max_over_ranks(timed(hstep, kh, dev)) / kh
    pxl = n * 512 * 512
    out["hybrid_c3"] = {
        "workload": "HybridLossPlus L1 + 0.05 DSSIM + 0.02 dE00 + 0.05 phys (stressed rgb PSF s=2) fwd+bwd, 16x3x512x512 fp32 "
                    "(BASELINE configs[2]), per GPU",
        "value": round(world * pxl / 1e6 / (ms_h / 1e3), 1), "unit": "Mpix/s", "ms_per_step": round(ms_h, 4),
        "algorithmic_bytes_per_step": (36 + 48) * pxl,
        "achieved_gbs": round((36 + 48) * pxl / ms_h / 1e6, 1), "n_gpus": world,
    }
    if world == 1 and not args.skip_cpu:
        from oracle import hybrid as ohyb
        cb, ca, ck = B.cpu(), A.cpu(), kern3.cpu()
        cx = xs[0].cpu().requires_grad_(True)
        t0 = time.perf_counter()
        tot, _ = ohyb.hybrid_loss(Bhat_raw=cx, B_raw=cb, A_raw=ca, expo_ratio=rho.cpu(), Bhat_srgb01=cx, B_srgb01=cb,
                                  w_l1=1.0, w_de=0.02, w_ssim=0.05, w_phys=0.05, physics_kernel=ck)
        tot.backward()
        dt = time.perf_counter() - t0
        out["hybrid_c3"]["cpu_baseline"] = {"value": round(pxl / 1e6 / dt, 3), "unit": "Mpix/s", "cores": os.cpu_count(),
                                            "kind": "port", "sample": f"1 step (no warm-up) of the oracle hybrid loss fwd+bwd, {dt:.1f} s"}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="phys_c2")
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs: 1 untimed-quality e2e step only")
    ap.add_argument("--skip-cpu", action="store_true", help="profiling runs: no cpu_baseline leg")
    ap.add_argument("--skip-extra", action="store_true", help="profiling runs: no eval_c4 / hybrid_c3 sections")
    ap.add_argument("--two-pass", action="store_true", help="time separate forward + backward kernels per step")
    ap.add_argument("--ring-rows", type=int, default=0, help="A/B: rows per work item of the phys ring kernels")
    ap.add_argument("--no-overlap", action="store_true", help="A/B: every launch waits for the previous grid at its top")
    ap.add_argument("--no-pdl", action="store_true", help="A/B: plain launches instead of programmatic dependent launch")
    args = ap.parse_args()
    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if args.impl == "reference":
        line = run_reference(args, rank)
        if line is not None:
            print(json.dumps(line), flush=True)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    line = run_ours(args, rank, world, dev)
    if not args.skip_extra:
        extra = run_extra(args, rank, world, dev)
        if line is not None:
            line.update(extra)
    if line is not None:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
