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
# BASELINE.json's metric, verbatim: `value` is its first half (phys-cons fwd+bwd, Mpix/s), `eval_c4.value` (last key of
# the line, both arms) its second half (SSIM / dE00 eval, Mpix/s), `roofline` the HBM fraction
METRIC = "Mpix/s phys-cons fwd+bwd & SSIM/\u0394E00 eval; HBM GB/s vs B200 peak, 1\u20138 GPU"
WORKLOAD = "phys_cons_raw rgb B2 depthwise PSF, per-image rho, 8x3x512x512 fp32, fwd+bwd (BASELINE configs[1])"


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, read at run time from the committed summary
    of the `ncu --set full` capture (profiles/ncu_traffic.json, written by tools/ncu_traffic.py from the .ncu-rep);
    (None, reason) when no capture is recorded -- never a constant in this file."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(path) as fh:
            rec = json.load(fh)[kernel]
        return float(rec["dram_bytes_per_launch"]), rec.get("source", "profiles/ncu_traffic.json")
    except Exception as exc:  # missing file / kernel
        return None, f"no ncu capture recorded ({type(exc).__name__})"
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
    """SM clocks / throttle reasons sampled DURING the run (B200_PROFILING.md), in-process through NVML (nvidia_ml_py):
    no process is forked anywhere near a timed region (the `nvidia-smi` fork sat inside the one graph replay
    it timed).  The thread runs from before the warm-up to the end; samples are kept only while `load` is set, i.e.
    while this rank keeps the GPU busy with timed work (not during the CPU-baseline legs).  `nvidia-smi` polling is only
    the fallback when NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, dev):
        self.index = dev.index
        self.rows, self.stop, self.load = [], threading.Event(), threading.Event()
        self.nv = self.handle = None
        try:
            import pynvml
            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(dev).uuid)
            self.handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            self.nv = pynvml
        except Exception:
            self.nv = None
        self.t = threading.Thread(target=self._run, daemon=True)

    def _sample_nvml(self):
        nv, h = self.nv, self.handle
        sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        r = int(get(h))
        bit = lambda *names: next((getattr(nv, n) for n in names if hasattr(nv, n)), 0)
        flags = [bit("nvmlClocksEventReasonHwSlowdown", "nvmlClocksThrottleReasonHwSlowdown"),
                 bit("nvmlClocksEventReasonHwThermalSlowdown", "nvmlClocksThrottleReasonHwThermalSlowdown"),
                 bit("nvmlClocksEventReasonSwThermalSlowdown", "nvmlClocksThrottleReasonSwThermalSlowdown"),
                 bit("nvmlClocksEventReasonSwPowerCap", "nvmlClocksThrottleReasonSwPowerCap")]
        return [str(sm), str(mx)] + ["Active" if (r & f) else "Not Active" for f in flags]

    def _run(self):
        while not self.stop.is_set():
            try:
                if not self.load.is_set():
                    pass
                elif self.nv is not None:
                    self.rows.append(self._sample_nvml())
                else:
                    out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                    self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.02 if self.nv is not None else 0.1)

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
                "reasons": reasons, "samples": len(sm), "source": "nvml" if self.nv is not None else "nvidia-smi"}


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


def reference_forward_check():
    """The UNMODIFIED reference module (oracle/_ref/metrics/phys_consistency.py, copied by oracle/build_ref.py) on the same
    batch: its `phys_cons_raw` is a no-grad metric, so it times the FORWARD half only and pins the
    port's forward value; the backward half of the step exists in the reference only as autograd of the same torch ops,
    which is what the port runs."""
    try:
        from oracle import build_ref, psf_ops
        if not build_ref.available():
            return {"reference_module": "oracle/_ref absent (built only where /root/reference exists)"}
        ref = build_ref.load("metrics/phys_consistency.py")
        pred, obs, rho = synth_batch(102, "cpu")
        dense = psf_ops.embed_dense(psf_ops.canonical_kernels("rgb"))
        ref.phys_cons_raw(pred, obs, dense, rho)
        t0 = time.perf_counter()
        for _ in range(5):
            v_ref = float(ref.phys_cons_raw(pred, obs, dense, rho))
        dt = (time.perf_counter() - t0) / 5
        v_port = float(psf_ops.phys_metric(pred, obs, dense, rho))
        mpix = SHAPE[0] * SHAPE[2] * SHAPE[3] / 1e6
        return {"reference_forward": {"kind": "reference", "value": round(mpix / dt, 3), "unit": "Mpix/s (forward only)",
                                      "sample": "5 calls of the unmodified reference phys_cons_raw (no-grad metric) after 1 warm-up",
                                      "port_minus_reference": v_port - v_ref}}
    except Exception as exc:  # the check must never take the bench line down
        return {"reference_module": f"unavailable: {type(exc).__name__}: {exc}"}


def cpu_eval_c4(pairs, n_timed):
    """The reference CPU path of the seven configs[3] metrics (oracle port) on `pairs` (host tensors): one warm-up on
    a quarter-size crop, then `n_timed` full-resolution pairs.  -> (Mpix/s, seconds)"""
    from oracle import colour, quality
    torch.set_num_threads(os.cpu_count() or 1)

    def cpu_metrics(cp, cg):
        quality.calculate_psnr(cg, cp, 1.0)
        float(quality.ssim_torchmetrics(cp, cg).mean())
        quality.cpsnr_rgb(cp, cg)
        quality.rgb_psnr(cp, cg)
        quality.rgb_ssim(cp, cg)
        colour.de00_summary(cp, cg, percentiles=(95.0,))      # deltae2000_mean pays the p95 sort in the reference
        colour.edge_stats(cp, cg, q=0.85)

    hh, ww = pairs[0][0].shape[-2] // 2, pairs[0][0].shape[-1] // 2
    cpu_metrics(pairs[0][0][..., :hh, :ww].contiguous(), pairs[0][1][..., :hh, :ww].contiguous())
    t0 = time.perf_counter()
    for i in range(n_timed):
        cpu_metrics(*pairs[i % len(pairs)])
    dt = time.perf_counter() - t0
    px = pairs[0][0].shape[-2] * pairs[0][0].shape[-1]
    return n_timed * px / 1e6 / dt, dt


def run_reference(args, rank):
    if rank != 0:
        return None
    steps = max(1, args.steps)
    val, ms, cores = time_cpu(steps, max(1, min(args.warmup, 3)))
    sample = (f"{steps} full C2 batches (8x3x512x512) fwd+bwd: the reference's torch ops for phys_cons_raw with autograd "
              f"enabled (oracle port; the reference function itself is a no-grad metric), torch CPU {cores} threads")
    cpu = {"value": round(val, 3), "unit": "Mpix/s", "cores": cores, "kind": "port", "sample": sample}
    cpu.update(reference_forward_check())
    return {
        "impl": "reference", "metric": METRIC, "value": round(val, 3), "unit": "Mpix/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": round(ms, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD},
        "cpu_baseline": cpu,
        "e2e": {"value": round(val, 3), "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        # second half of the metric (LAST key, as in the GPU arm's line): the same seven configs[3] metrics on the host
        "eval_c4": cpu_eval_line(args),
    }


def cpu_eval_line(args):
    if args.skip_extra:
        return None
    pairs = [synth_eval_pair(4000 + i, torch.device("cpu")) for i in range(2)]
    v, dt = cpu_eval_c4(pairs, 2)
    return {"value": round(v, 3), "unit": "Mpix/s", "images": 2, "n_gpus": args.gpus, "metrics": 7, "kind": "port",
            "cores": os.cpu_count(),
            "sample": f"2 full-resolution 2848x4256 pairs after a quarter-size warm-up, oracle port, {dt:.1f} s"}


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

    def run(self, rounds=1):
        """`rounds` back-to-back passes over all `steps` calls, each bracketed by its own pair of CUDA events on the
        launching stream (no host synchronisation in between, nothing else enqueued) -> list of elapsed ms."""
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(rounds + 1)]
        torch.cuda.synchronize(self.dev)
        with torch.cuda.stream(self.stream):
            ev[0].record(self.stream)
            for r in range(rounds):
                for _ in range(self.replays):
                    self.graph.replay()
                ev[r + 1].record(self.stream)
        torch.cuda.synchronize(self.dev)
        return [ev[r].elapsed_time(ev[r + 1]) for r in range(rounds)]

    def per_step(self, steps, min_ms=100.0, min_rounds=5):
        """Median time per step (ms) over enough rounds to fill >= `min_ms` of device time (>= `min_rounds`)."""
        first = self.run(1)[0]
        rounds = max(min_rounds, int(min_ms / max(first, 1e-3)) + 1)
        t = sorted(self.run(min(rounds, 4000)))
        return t[len(t) // 2] / steps, len(t), sum(t)


def run_ours(args, rank, world, dev):
    import torch.distributed as dist
    from lowlight_image_enhancement_b200 import _functional as Fn
    from lowlight_image_enhancement_b200 import _native as nat
    from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw
    from oracle import psf_ops  # kernel constants for input synthesis only

    if args.prefetch_rows >= 0:
        nat.check(nat.lib().llie_tune(7, args.prefetch_rows), "llie_tune")  # LLIE_TUNE_PHYS_PREFETCH
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
    clk = ClockSampler(dev)
    clk.__enter__()
    clk.load.set()
    # Headline = the DEFAULT launch mode of the product (what phys_cons_raw, the loss classes and autograd run): every
    # launch is ordered after the previous grid.  The stream-overlap contract (llie_phys_desc.stream_overlap, loops over
    # independent batches only) is measured next to it and reported as `stream_overlap`, not as `value`.
    # (the flag is a kernel argument: frozen into a graph at capture time)
    g_step = GraphLoop(step, K, dev)
    g_f, g_b, g_u, g_2 = GraphLoop(fwd, K, dev), GraphLoop(bwd, K, dev), GraphLoop(fused, K, dev), GraphLoop(step2, K, dev)
    for _ in range(max(1, W // max(1, K // g_step.replays) + 1)):
        g_step.run()
    t_f = g_f.per_step(K, 30.0, 3)[0]
    t_b = g_b.per_step(K, 30.0, 3)[0]
    t_u = g_u.per_step(K, 30.0, 3)[0]
    t_2 = g_2.per_step(K, 30.0, 3)[0]
    if world > 1:
        dist.barrier()
    # timed region: R rounds of EXACTLY K steps, >= 100 ms of device time in total, median round (one replay of a
    # 20-step graph is 0.3 ms: a single sample of it measured the launch jitter of eight ranks, not the kernel)
    ms_step, rounds, ms_total = g_step.per_step(K, 130.0, 5)
    launches = (2 if args.two_pass else 1) * K * rounds
    overlap = None
    if not (args.no_overlap or args.no_pdl):
        with Fn.phys_stream_overlap():  # per-call promise, scoped: everything outside this block is default mode
            g_o = GraphLoop(fused, K, dev)
            overlap = g_o.per_step(K, 50.0, 5)[0]
            del g_o
    if world > 1:
        t = torch.tensor([ms_step], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        ms_step = float(t)
    ms = ms_step * K
    mpix = n * h * w / 1e6
    value = world * mpix / (ms_step / 1e3)

    # ---- e2e: public API (phys_cons_raw(differentiable=True) + backward) from pinned HOST buffers; the H2D copies of
    # every step's pred / obs and the D2H read of its loss are inside the timed region.  Double-buffered: the upload of
    # step i + 1 runs on a copy stream while step i computes and its loss is read back (the copy engine never idles;
    # the step is PCIe-bound: 50.3 MB per step).
    hp = [(s[0].cpu().pin_memory(), s[1].cpu().pin_memory()) for s in sets[:4]]
    rho_d = sets[0][2]
    copy_st = torch.cuda.Stream(dev)
    dbuf = [(torch.empty(SHAPE, device=dev), torch.empty(SHAPE, device=dev)) for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    freed = [torch.cuda.Event() for _ in range(2)]

    def upload(i):
        with torch.cuda.stream(copy_st):
            copy_st.wait_event(freed[i % 2])       # the step that last used this device buffer pair has finished
            dbuf[i % 2][0].copy_(hp[i % 4][0], non_blocking=True)
            dbuf[i % 2][1].copy_(hp[i % 4][1], non_blocking=True)
            ready[i % 2].record(copy_st)

    def e2e_run(steps):
        main = torch.cuda.current_stream(dev)
        for e in freed:
            e.record(main)
        upload(0)
        out = 0.0
        for i in range(steps):
            if i + 1 < steps:
                upload(i + 1)
            main.wait_event(ready[i % 2])
            x = dbuf[i % 2][0].detach().requires_grad_(True)
            loss = phys_cons_raw(x, dbuf[i % 2][1], kern, rho_d, differentiable=True)
            loss.backward()
            freed[i % 2].record(main)
            out += loss.item()                     # D2H read of the step's result (host sync)
        return out

    Ke = 1 if args.skip_e2e else max(5, min(K, 50))
    e2e_run(1 if args.skip_e2e else 3)
    s_e, e_e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    s_e.record()
    e2e_run(Ke)
    e_e.record()
    torch.cuda.synchronize(dev)
    ms_e = s_e.elapsed_time(e_e)
    if world > 1:
        t = torch.tensor([ms_e], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e = float(t)
    e2e_val = world * mpix * Ke / (ms_e / 1e3)

    # secondary end-to-end leg, the shape of a real evaluation: B_hat is the network's output and already lives on the
    # device; what crosses PCIe per step is the OBSERVED frame in the dataset's own 16-bit form (2 B/sample, decoded
    # and exposure-aligned on the GPU by llie_decode_u16, sony_sid_lmdb_dataset.py:208-233) and the 4-byte result
    from lowlight_image_enhancement_b200 import _functional as Fn2
    hu = [torch.randint(0, 65536, (n, h, w, c), dtype=torch.int32).to(torch.uint16).pin_memory() for _ in range(4)]
    du = [torch.empty((n, h, w, c), dtype=torch.uint16, device=dev) for _ in range(2)]
    ready_u = [torch.cuda.Event() for _ in range(2)]
    freed_u = [torch.cuda.Event() for _ in range(2)]
    one = torch.ones(n, device=dev)

    def upload_u(i):
        with torch.cuda.stream(copy_st):
            copy_st.wait_event(freed_u[i % 2])
            du[i % 2].copy_(hu[i % 4], non_blocking=True)
            ready_u[i % 2].record(copy_st)

    def e2e_u16(steps):
        main = torch.cuda.current_stream(dev)
        for e in freed_u:
            e.record(main)
        upload_u(0)
        out = 0.0
        for i in range(steps):
            if i + 1 < steps:
                upload_u(i + 1)
            main.wait_event(ready_u[i % 2])
            _, obs_i = Fn2.decode_u16(du[i % 2], one, want_raw=False)
            x = dbuf[i % 2][0].detach().requires_grad_(True)
            loss = phys_cons_raw(x, obs_i, kern, rho_d, differentiable=True)
            loss.backward()
            freed_u[i % 2].record(main)
            out += loss.item()
        return out

    e2e_u16(1 if args.skip_e2e else 3)
    s_u, e_u = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    s_u.record()
    e2e_u16(Ke)
    e_u.record()
    torch.cuda.synchronize(dev)
    ms_u = s_u.elapsed_time(e_u)
    if world > 1:
        t = torch.tensor([ms_u], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_u = float(t)
    e2e_u16_val = world * mpix * Ke / (ms_u / 1e3)
    del hu, du
    clk.load.clear()

    if rank != 0:
        return None, clk
    peak, peak_src = peaks()
    elems = n * c * h * w
    bytes_f, bytes_b = 8 * elems, 12 * elems
    ach_b = bytes_b / (t_b * 1e-3) / 1e9
    ach_f = bytes_f / (t_f * 1e-3) / 1e9
    ach_u = bytes_b / (t_u * 1e-3) / 1e9   # the fused pass moves the backward's 12 B/element
    roof = ({"kernel": "phys_bwd_ring", "achieved": ach_b, "us": t_b} if args.two_pass else
            {"kernel": "phys_bwd_ring<FUSED> (value + gradient)", "achieved": ach_u, "us": t_u})
    traffic, traffic_src = ncu_traffic("phys_bwd_ring" if args.two_pass else "phys_fused_ring")
    line = {
        "metric": METRIC, "value": round(value, 1), "unit": "Mpix/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": round(ms_step, 6), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD},
        "notes": {"step": ("forward kernel + backward kernel" if args.two_pass else
                           "loss value + gradient from one fused launch (llie_phys_value_and_grad); "
                           f"two-pass forward + backward = {round(t_2 * 1e3, 2)} us per step"),
                  "l2": f"inputs rotate over {POOL} resident buffer sets (604 MB per GPU) > 126 MB L2",
                  "launch": f"default (serialised) launches, steps captured in CUDA graphs of {K // g_step.replays} steps; "
                            f"timed region = {rounds} rounds of {K} steps = {round(ms_total, 1)} ms of device time, "
                            "median round reported",
                  "per_gpu_batch": list(SHAPE)},
        "roofline": {"bound": "hbm", "kernel": roof["kernel"], "achieved": round(roof["achieved"], 1), "peak": peak,
                     "unit": "GB/s", "frac": round(roof["achieved"] / peak, 4), "traffic": traffic,
                     "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": bytes_b,
                     "us_per_launch": round(roof["us"] * 1e3, 2), "launch_mode": "default (serialised)"},
        "stream_overlap": (None if overlap is None else {
            "contract": "llie_phys_desc.stream_overlap, per call: loops over independent batches only (never under autograd)",
            "us_per_launch": round(overlap * 1e3, 2), "value": round(world * mpix / (overlap / 1e3), 1), "unit": "Mpix/s",
            "frac": round(bytes_b / (overlap * 1e-3) / 1e9 / peak, 4)}),
        "kernels": [
            {"kernel": "phys_fwd_ring", "us": round(t_f * 1e3, 2), "algorithmic_bytes": bytes_f,
             "achieved_gbs": round(ach_f, 1), "frac": round(ach_f / peak, 4), "traffic": ncu_traffic("phys_fwd_ring")[0]},
            {"kernel": "phys_bwd_ring", "us": round(t_b * 1e3, 2), "algorithmic_bytes": bytes_b,
             "achieved_gbs": round(ach_b, 1), "frac": round(ach_b / peak, 4), "traffic": ncu_traffic("phys_bwd_ring")[0]},
            {"kernel": "phys_bwd_ring<FUSED>", "us": round(t_u * 1e3, 2), "algorithmic_bytes": bytes_b,
             "achieved_gbs": round(ach_u, 1), "frac": round(ach_u / peak, 4), "traffic": ncu_traffic("phys_fused_ring")[0]}],
        "e2e": {"value": round(e2e_val, 1), "unit": "Mpix/s", "h2d_bytes_per_step": 2 * 4 * elems,
                "d2h_bytes_per_step": 4, "steps": Ke,
                "h2d_gbs_per_gpu": round(2 * 4 * elems * Ke / (ms_e * 1e-3) / 1e9, 1),   # the leg is bound by this link
                "api": "metrics.phys_consistency.phys_cons_raw(differentiable=True)+backward, double-buffered uploads"},
        "e2e_u16_obs": {"value": round(e2e_u16_val, 1), "unit": "Mpix/s", "h2d_bytes_per_step": 2 * elems,
                        "d2h_bytes_per_step": 4, "steps": Ke,
                        "api": "B_hat resident (network output), observed frame uploaded as the dataset's uint16 "
                               "[n,h,w,3] and decoded / exposure-aligned on the GPU (llie_decode_u16), then the same "
                               "phys_cons_raw + backward"},
        "gpu_launches": launches,
    }
    if world == 1 and not args.skip_cpu:
        v, ms_c, cores = time_cpu(5, 2)
        line["cpu_baseline"] = {"value": round(v, 3), "unit": "Mpix/s", "cores": cores, "kind": "port",
                                "sample": "5 full C2 batches fwd+bwd after 2 warm-up steps, oracle port of phys_cons_raw "
                                          "(autograd backward), torch CPU all threads"}
        line["cpu_baseline"].update(reference_forward_check())
    return line, clk


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


def run_extra(args, rank, world, dev, clk=None):
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

    # ---- eval (BASELINE configs[3]): ALL seven metrics the config names, images sharded by index over the ranks,
    # per-image metrics accumulated on the device, ONE all-reduce + ONE host sync at the end (eval.py)
    from lowlight_image_enhancement_b200.eval import C4_METRICS
    pool = [synth_eval_pair(4000 + 17 * rank + i, dev) for i in range(3)]   # 3 x 291 MB > L2
    per_rank = args.eval_images
    n_img = per_rank * world
    px = EVAL_SHAPE[2] * EVAL_SHAPE[3]
    getter = lambda i: pool[(i // world) % 3]
    evaluate_pairs(getter, C4_METRICS, n_items=3 * world, device=dev)    # warm-up (also NCCL)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    s_ev, e_ev = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s_ev.record()
    res = evaluate_pairs(getter, C4_METRICS, n_items=n_img, device=dev)
    e_ev.record()
    torch.cuda.synchronize(dev)
    wall = (time.perf_counter() - t0) * 1e3
    ms = max_over_ranks(s_ev.elapsed_time(e_ev))
    # the four adapters of round 1 alone (comparison with profiles/r1_*), same driver
    evaluate_pairs(getter, EVAL_METRICS, n_items=3 * world, device=dev)
    torch.cuda.synchronize(dev)
    s_ev.record()
    evaluate_pairs(getter, EVAL_METRICS, n_items=max(6, per_rank // 4) * world, device=dev)
    e_ev.record()
    torch.cuda.synchronize(dev)
    ms4 = max_over_ranks(s_ev.elapsed_time(e_ev)) / max(6, per_rank // 4)
    win_l = Fn.window_1d("linear", 11, 1.5, dev)
    from lowlight_image_enhancement_b200 import _native as nat_
    t_ssim = timed_warm(lambda i: torchmetrics_ssim_scores(pool[i % 3][0], pool[i % 3][1], data_range=1.0, sigma=1.5,
                                                      k1=0.01, k2=0.03), 12, dev) / 12
    t_ssim5 = timed_warm(lambda i: Fn.ssim_stats(pool[i % 3][0], pool[i % 3][1], win_l, pad_mode=nat_.PAD["reflect"],
                                                 c1=1e-4, c2=9e-4, eps=1e-12, clamp_var=True), 12, dev) / 12
    win_t = Fn.window_1d("torchmetrics", 11, 1.5, dev)
    t_dual = timed_warm(lambda i: Fn.ssim_stats_dual(pool[i % 3][0], pool[i % 3][1], win_l, win_t, c1=1e-4, c2=9e-4,
                                                     eps=1e-12, c1_b=1e-4, c2_b=9e-4), 12, dev) / 12
    t_de = timed_warm(lambda i: Fn.deltae(pool[i % 3][0], pool[i % 3][1], variant=0, clamp_in=True), 12, dev) / 12
    t_ps = timed_warm(lambda i: Fn.psnr_stats(pool[i % 3][0], pool[i % 3][1], 1, 3, px, cast_first=True, clamp_hi=None,
                                         data_range=1.0, eps=1e-12), 24, dev) / 24
    t_all = timed_warm(lambda i: image_metrics(pool[i % 3][0], pool[i % 3][1], C4_METRICS), 6, dev) / 6
    acc_k = ShardedMeans(C4_METRICS, dev)
    t_fast = timed_warm(lambda i: accumulate_image_metrics(pool[i % 3][0], pool[i % 3][1], C4_METRICS, acc_k), 12,
                        dev) / 12

    # FP32 roofline of the filter kernels: 148 SMs x 128 FP32 lanes x SM clock multiply-adds per second (the packed
    # FFMA2 form occupies the pipe for two cycles, so it does not raise this peak: tools/micro/pipe_rates.cu)
    sm_hz = (clk.summary().get("sm_max_mhz") or 1965) * 1e6 if clk is not None else 1965e6
    peak_fma = 148 * 128 * sm_hz

    def kern(name, t, bpp, bound, fma_per_elem=None):
        rec = {"kernel": name, "ms": round(t, 4), "Mpix/s": round(px / t / 1e3, 1),
               "algorithmic_bytes": bpp * px, "achieved_gbs": round(bpp * px / t / 1e6, 1),
               "frac_hbm": round(bpp * px / t / 1e6 / peak, 4), "bound": bound}
        if fma_per_elem is not None:   # elements = 3 channels x pixels
            rate = fma_per_elem * 3 * px / (t * 1e-3)
            rec.update({"algorithmic_fma_per_element": fma_per_elem, "achieved_tfma": round(rate / 1e12, 2),
                        "peak_tfma": round(peak_fma / 1e12, 2), "frac_fp32": round(rate / peak_fma, 4)})
        return rec
    out["eval_c4_detail"] = {
        "workload": "PSNR + SSIM(torchmetrics policy) + CPSNR + RGB-PSNR + RGB-SSIM + dE00 mean + edge dE00 on 1x3x2848x4256 "
                    f"sRGB pairs (BASELINE configs[3]), {per_rank} images per GPU sharded by index, device-side "
                    "accumulation, one all-reduce, one host sync",
        "wall_ms": round(wall, 2),
        "means": {k: (round(v, 6) if isinstance(v, float) else v) for k, v in res.items()},
        "ms_per_image_four_adapters_only": round(ms4, 4),
        "kernels": [kern("psnr_stats_kernel (per-channel table: CPSNR, RGB-PSNR)", t_ps, 24, "hbm"),
                    kern("ssim_ring_kernel<4 maps> (SSIM, torchmetrics policy)", t_ssim, 24, "fp32 pipe / issue (88 FMA per element)", 88),
                    kern("ssim_ring_kernel<5 maps> (RGB-SSIM, ssim_linear policy)", t_ssim5, 24, "fp32 pipe / issue (110 FMA per element)", 110),
                    kern("ssim_ring_kernel<5 maps, DUAL> (both policies from one pass: what the driver launches)", t_dual, 24,
                         "fp32 pipe / issue (110 FMA per element + second formula)", 110),
                    kern("deltae_kernel", t_de, 24, "instruction issue (fp32 + sfu)"),
                    kern("all 7 metrics of one image, image_metrics (generic path)", t_all, 36, "mixed"),
                    kern("all 7 metrics of one image, accumulate_image_metrics (one-launch epilogue)", t_fast, 36, "mixed")],
    }
    # compact summary: main() appends it as the LAST keys of the JSON line (driver tails keep the end of stdout)
    out["eval_c4"] = {"value": round(n_img * px / 1e6 / (ms / 1e3), 1), "unit": "Mpix/s", "images": n_img,
                      "n_gpus": world, "ms_per_image_per_gpu": round(ms / per_rank, 4), "metrics": len(C4_METRICS)}
    if world == 1 and not args.skip_cpu:
        if clk is not None:
            clk.load.clear()              # CPU leg: the idle GPU must not enter the clock record
        # reference CPU path of the same seven metrics: 1 warm-up pair (quarter size) + 4 full-resolution pairs
        host = [(p_.cpu(), g_.cpu()) for p_, g_ in pool]
        v_cpu, dt = cpu_eval_c4(host, 4)
        out["eval_c4_detail"]["cpu_baseline"] = {
            "value": round(v_cpu, 3), "unit": "Mpix/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"4 full-resolution pairs after a quarter-size warm-up, the same 7 metrics, {dt:.1f} s"}
        out["eval_c4"]["cpu_Mpix/s"] = out["eval_c4_detail"]["cpu_baseline"]["value"]
        del host
    if clk is not None:
        clk.load.set()
    del pool
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
        if clk is not None:
            clk.load.clear()              # CPU leg: the idle GPU must not enter the clock record
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
    if clk is not None:
        clk.load.set()
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
                         physics_kernel=kern3, check_finite=False, perceptual_module=False)

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
        if clk is not None:
            clk.load.clear()              # CPU leg: the idle GPU must not enter the clock record
        from oracle import hybrid as ohyb
        cb, ca, ck, cr = B.cpu(), A.cpu(), kern3.cpu(), rho.cpu()
        cxs = [x_.cpu() for x_ in xs[:2]]

        def cstep(i):
            cx = cxs[i % 2].detach().requires_grad_(True)
            tot, _ = ohyb.hybrid_loss(Bhat_raw=cx, B_raw=cb, A_raw=ca, expo_ratio=cr, Bhat_srgb01=cx, B_srgb01=cb,
                                      w_l1=1.0, w_de=0.02, w_ssim=0.05, w_phys=0.05, physics_kernel=ck)
            tot.backward()

        cstep(0)
        t0 = time.perf_counter()
        for i in range(5):
            cstep(i + 1)
        dt = (time.perf_counter() - t0) / 5
        out["hybrid_c3"]["cpu_baseline"] = {"value": round(pxl / 1e6 / dt, 3), "unit": "Mpix/s", "cores": os.cpu_count(),
                                            "kind": "port", "sample": f"5 steps after 1 warm-up of the oracle hybrid loss fwd+bwd, {dt:.2f} s per step"}
    if clk is not None:
        clk.load.set()

    # ---- dataset-level phys-consistency loop through the public API (eval.evaluate_phys_consistency): the product
    # caller of the per-call stream-overlap promise.  128-image batches: one launch outlasts the Python launch path.
    from lowlight_image_enhancement_b200.eval import evaluate_phys_consistency
    nb = 128
    pool_b = []
    for i in range(3):
        pb = torch.rand(nb, 3, 512, 512, device=dev, generator=torch.Generator(device=dev).manual_seed(7100 + i))
        pool_b.append((pb, (pb * 0.9).contiguous(), torch.full((nb,), 1.1, device=dev)))
    seq = [pool_b[i % 3] for i in range(48)]
    from oracle import psf_ops as _pops   # PSF constants only (canonical kernels); the metric itself runs on the GPU kernels
    psf_b = _pops.canonical_kernels("rgb").to(dev)
    loop = {}
    for _ in range(6):   # the CPU baseline legs above left the GPU idle for seconds: bring the clocks back up first
        evaluate_phys_consistency(seq, psf_b, overlap=False, rank=0, world_size=1)
    for name, ov in (("serialised", False), ("overlapped", True)):
        evaluate_phys_consistency(seq, psf_b, overlap=ov, rank=0, world_size=1)
        torch.cuda.synchronize(dev)
        s_l, e_l = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s_l.record()
        for _ in range(4):
            evaluate_phys_consistency(seq, psf_b, overlap=ov, rank=0, world_size=1)
        e_l.record()
        torch.cuda.synchronize(dev)
        us = s_l.elapsed_time(e_l) * 1e3 / (4 * len(seq))
        gbs = 8 * 3 * nb * 512 * 512 / (us * 1e-6) / 1e9
        loop[name] = {"us_per_batch": round(us, 1), "Mpix/s": round(nb * 512 * 512 / us, 1),
                      "achieved_gbs": round(gbs, 1), "frac_hbm": round(gbs / peak, 4)}
    out["phys_eval_loop"] = {"api": "eval.evaluate_phys_consistency (forward metric, 48 resident batches of 128x3x512x512 "
                                    "fp32 per call, device-side table, one host sync); `overlapped` = launches 2..B carry "
                                    "llie_phys_desc.stream_overlap", "per_gpu": loop}
    del pool_b, seq
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------
# BASELINE configs[4]: the loss path inside a full NewBP-NAFNet training step (bf16 autocast, AdamW, DDP)
# ------------------------------------------------------------------------------------------------
def run_c5(args, rank, world, dev):
    """One training step as `ImageRestorationModel.optimize_parameters` runs it (image_restoration_model.py:248-322):
    autocast(bf16) { out = NAFNet(lq); HybridLossPlus(out, ...) }, backward, grad-clip 0.01, AdamW step.  The backbone is
    stock PyTorch (nafnet_stock.NAFNet, width 32, enc [2,2,4,8], middle 12, dec [2,2,2,2], configs/colab/
    sid_newbp_rgb.yml:48-58, random init); the loss path is either this repo's kernels or the eager torch ops of the
    reference (oracle restatement run on the GPU under the same autocast).  Global batch `--c5-batch` split over the
    ranks (strong scaling), DDP gradient all-reduce over NCCL.  Reports step time with either loss path and the loss
    path's own forward + backward time -> its share of the step, before / after."""
    import torch.distributed as dist
    from lowlight_image_enhancement_b200.NewBP_model.losses import HybridLossPlus
    from lowlight_image_enhancement_b200.NewBP_model.nafnet_stock import NAFNet
    from lowlight_image_enhancement_b200.NewBP_model.newbp_layer import CrosstalkPSF
    from oracle import hybrid as ohyb
    if args.c5_batch % world:
        return {"c5_train_step": {"skipped": f"global batch {args.c5_batch} not divisible by {world} ranks"}}
    b = args.c5_batch // world
    torch.manual_seed(105)
    net = NAFNet(img_channel=3, width=32, middle_blk_num=12, enc_blk_nums=[2, 2, 4, 8], dec_blk_nums=[2, 2, 2, 2]).to(dev)
    for prm in net.parameters():           # beta / gamma start at zero upstream (identity blocks): make every block live
        if prm.dim() == 4 and prm.shape[0] == 1 and float(prm.detach().abs().sum()) == 0.0:
            torch.nn.init.normal_(prm, std=0.1)
    model = net
    if world > 1:
        model = torch.nn.parallel.DistributedDataParallel(net, device_ids=[dev.index])
    opt = torch.optim.AdamW(net.parameters(), lr=1e-4, betas=(0.9, 0.9), weight_decay=0.0)
    g = torch.Generator(device=dev).manual_seed(105 + rank)
    gt = torch.rand(b, 3, 512, 512, device=dev, generator=g)
    lq = (0.6 * gt + 0.02 * torch.randn(gt.shape, device=dev, generator=g)).clamp(0, 1)
    rho = torch.ones(b, 1, 1, 1, device=dev)
    psf = CrosstalkPSF("rgb", stressed_kernels(2.0, dev)).to(dev)
    ours = HybridLossPlus(device=dev, w_l1_raw=1.0, w_deltaE=0.02, w_ssim=0.05, w_phys=0.10, physics_psf_module=psf,
                          perceptual_module=False, check_finite=False)
    kpsf = psf.kernel.detach().clone()

    def loss_ours(out):
        return ours(Bhat_raw=out, B_raw=gt, A_raw=lq, expo_ratio=rho, Bhat_srgb01=out.clamp(0.0, 1.0),
                    B_srgb01=gt.clamp(0.0, 1.0), A_srgb01=lq.clamp(0.0, 1.0))[0]

    def loss_eager(out):
        return ohyb.hybrid_loss(Bhat_raw=out, B_raw=gt, A_raw=lq, expo_ratio=rho, Bhat_srgb01=out.clamp(0.0, 1.0),
                                B_srgb01=gt.clamp(0.0, 1.0), A_srgb01=lq.clamp(0.0, 1.0), w_l1=1.0, w_de=0.02,
                                w_ssim=0.05, w_phys=0.10, psf_kernel=kpsf)[0]

    def train_step(loss_fn):
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = model(lq)
            total = loss_fn(out)
        total.backward()
        torch.nn.utils.clip_grad_norm_(net.parameters(), 0.01)
        opt.step()
        return total

    out_fixed = None

    def loss_only(loss_fn):
        x = out_fixed.detach().requires_grad_(True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            total = loss_fn(x)
        total.backward()
        return total

    def t_ms(fn, arg, iters):
        for _ in range(2):
            fn(arg)
        s_, e_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        s_.record()
        for _ in range(iters):
            fn(arg)
        e_.record()
        torch.cuda.synchronize(dev)
        ms = s_.elapsed_time(e_) / iters
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms

    try:
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            out_fixed = model(lq)                      # bf16 network output, as the loss path sees it in training
        res = {}
        for name, fn in (("b200_kernels", loss_ours), ("eager_reference_ops", loss_eager)):
            step_ms = t_ms(train_step, fn, 4)
            loss_ms = t_ms(loss_only, fn, 8)
            res[name] = {"ms_per_step": round(step_ms, 3), "loss_path_ms": round(loss_ms, 3),
                         "loss_path_share": round(loss_ms / step_ms, 4),
                         "images_per_s": round(args.c5_batch / (step_ms / 1e3), 1)}
        vo, ve = float(loss_ours(out_fixed.float())), float(loss_eager(out_fixed.float()))
        out = {"workload": "NewBP-NAFNet (width 32, enc [2,2,4,8], mid 12, dec [2,2,2,2]) training step, bf16 autocast, "
                           f"AdamW, grad clip, global batch {args.c5_batch}x3x512x512 over {world} GPU(s) (DDP), HybridLossPlus "
                           "L1 + 0.02 dE00 + 0.05 DSSIM + 0.10 phys_srgb (stressed rgb PSF), VGG term excluded in both arms "
                           "(BASELINE configs[4])",
               "n_gpus": world, "per_gpu_batch": b, **res,
               "loss_value_check": {"b200": round(vo, 6), "eager_fp32": round(ve, 6)},
               "step_speedup": round(res["eager_reference_ops"]["ms_per_step"] / res["b200_kernels"]["ms_per_step"], 3),
               "loss_path_speedup": round(res["eager_reference_ops"]["loss_path_ms"] / res["b200_kernels"]["loss_path_ms"], 2)}
    except torch.cuda.OutOfMemoryError as exc:   # never take the headline line down
        out = {"skipped": f"out of memory at per-GPU batch {b}: {str(exc)[:120]}"}
    del net, model, opt
    torch.cuda.empty_cache()
    return {"c5_train_step": out}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="phys_c2")
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs: 1 untimed-quality e2e step only")
    ap.add_argument("--skip-cpu", action="store_true", help="profiling runs: no cpu_baseline leg")
    ap.add_argument("--skip-extra", action="store_true", help="profiling runs: no eval_c4 / hybrid_c3 / c5 sections")
    ap.add_argument("--skip-c5", action="store_true", help="no NAFNet training-step section (BASELINE configs[4])")
    ap.add_argument("--eval-images", type=int, default=64, help="full-resolution pairs per GPU in the eval_c4 section")
    ap.add_argument("--c5-batch", type=int, default=32, help="GLOBAL batch of the C5 training step (split over the ranks)")
    ap.add_argument("--two-pass", action="store_true", help="time separate forward + backward kernels per step")
    ap.add_argument("--ring-rows", type=int, default=0, help="A/B: rows per work item of the phys ring kernels")
    ap.add_argument("--no-overlap", action="store_true", help="A/B: every launch waits for the previous grid at its top")
    ap.add_argument("--no-pdl", action="store_true", help="A/B: plain launches instead of programmatic dependent launch")
    ap.add_argument("--prefetch-rows", type=int, default=-1, help="A/B: rows of a ring CTA's first item prefetched into L2 before its grid wait (0 = off)")
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
    line, clk = run_ours(args, rank, world, dev)
    extra = {}
    if not args.skip_extra:
        clk.load.set()
        extra = run_extra(args, rank, world, dev, clk)
        if not args.skip_c5:
            extra.update(run_c5(args, rank, world, dev))
        clk.load.clear()
    clk.__exit__()
    if line is not None:
        tail = extra.pop("eval_c4", None)
        line.update(extra)
        line["clocks"] = clk.summary()
        if tail is not None:
            line["eval_c4"] = tail          # LAST key: the sharded-eval scaling evidence survives a truncated stdout tail
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
