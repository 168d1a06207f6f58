"""Sharded validation driver (lowlight_image_enhancement_b200/eval.py).

CPU part: world_size-2 gloo run of the host logic (image -> rank rule of irm:359-361, per-rank float64
sums, ONE all-reduce, means as irm:459-464) with an injected host metric function.
GPU part: the device metric scalars equal the adapters' floats and the driver's means equal the
serial loop over adapters."""
import math
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lowlight_image_enhancement_b200.eval import ShardedMeans, evaluate_pairs, shard_indices

NAMES = ["a", "b", "c"]


def _fake_metric(pred, target, names, flags, **_):
    # deterministic per-image values that depend on the pair only
    s = float(pred.sum()) - float(target.sum())
    if not torch.isfinite(pred).all():
        flags += 1
    return torch.tensor([s, s * s, 1.0 + pred.numel()], dtype=torch.float64)


def _pairs(n):
    g = torch.Generator().manual_seed(5)
    return [(torch.rand(3, 4 + i % 3, 5, generator=g), torch.rand(3, 4 + i % 3, 5, generator=g)) for i in range(n)]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, shard, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        touched = []
        pairs = _pairs(n)

        def getter(i):
            touched.append(i)
            return pairs[i]

        res = evaluate_pairs(getter, NAMES, n_items=n, device="cpu", shard=shard, metric_fn=_fake_metric)
        out.put((rank, res, touched))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,shard", [(7, "interleave"), (8, "block"), (1, "interleave")])
def test_two_rank_gloo_matches_serial(n, shard):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, shard, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    serial = evaluate_pairs(_pairs(n), NAMES, device="cpu", metric_fn=_fake_metric, rank=0, world_size=1)
    seen = []
    for rank, res, touched in got:
        assert touched == shard_indices(n, rank, world, shard)       # a rank reads only its own images
        seen += touched
        assert res["count"] == n and res["nonfinite_inputs"] == 0
        for k in NAMES:
            assert res[k] == pytest.approx(serial[k], rel=1e-12)
    assert sorted(seen) == list(range(n))                            # every image exactly once


def test_shard_rules_and_accumulator():
    assert shard_indices(10, 1, 4) == [1, 5, 9]                       # i % world == rank
    assert [shard_indices(10, r, 4, "block") for r in range(4)] == [[0, 1, 2], [3, 4, 5], [6, 7], [8, 9]]
    assert shard_indices(2, 3, 4) == []
    with pytest.raises(ValueError):
        shard_indices(4, 4, 4)
    acc = ShardedMeans(["x", "y"], "cpu")
    acc.add(torch.tensor([1.0, 10.0]))
    acc.add(torch.tensor([3.0, float("inf")]), flags=torch.tensor([2], dtype=torch.int32))
    out = acc.reduce()
    assert out["x"] == 2.0 and out["y"] == float("inf") and out["count"] == 2 and out["nonfinite_inputs"] == 1
    empty = ShardedMeans(["x"], "cpu").reduce()
    assert empty["count"] == 0 and empty["x"] != empty["x"]
    with pytest.raises(ValueError):
        acc.add(torch.tensor([1.0]))


@pytest.mark.gpu
def test_device_metrics_equal_adapters():
    from lowlight_image_enhancement_b200 import lowlight_metrics as lm
    from lowlight_image_enhancement_b200.eval import METRIC_FUNCTIONS, image_metrics
    dev = torch.device("cuda:0")
    names = sorted(k for k in METRIC_FUNCTIONS if hasattr(lm, k))      # the five adapter names
    g = torch.Generator().manual_seed(12)
    pairs = []
    for i in range(5):
        gt = torch.rand(1, 3, 40 + 8 * i, 72, generator=g)
        gt = torch.round(255 * torch.nn.functional.avg_pool2d(gt, 5, 1, 2)) / 255
        pred = (gt + 0.04 * torch.randn(gt.shape, generator=g))
        pairs.append((pred.to(dev), gt.to(dev)))
    serial = {k: 0.0 for k in names}
    for pred, gt in pairs:
        vals = image_metrics(pred, gt, names)
        assert vals.dtype == torch.float64 and vals.is_cuda
        for k, v in zip(names, vals.tolist()):
            ref = getattr(lm, k)(pred, gt)
            assert v == pytest.approx(ref, rel=1e-6, abs=1e-9), k
            serial[k] += ref / len(pairs)
    res = evaluate_pairs(pairs, names, device=dev)
    assert res["count"] == 5 and res["nonfinite_inputs"] == 0
    for k in names:
        assert res[k] == pytest.approx(serial[k], rel=1e-6)
    bad = pairs[0][0].clone()
    bad[0, 0, 0, 0] = float("nan")
    assert evaluate_pairs([(bad, pairs[0][1])], ["linear_psnr", "deltae2000_mean"], device=dev)["nonfinite_inputs"] == 1
    with pytest.raises(KeyError):
        image_metrics(pairs[0][0], pairs[0][1], ["nope"])


def _reference_grid(h, w, ch, cw, scale=1):
    """image_restoration_model.py:167-208 index walk restated for the test."""
    import math
    ch, cw = ch // scale * scale, cw // scale * scale
    nr, nc = (h - 1) // ch + 1, (w - 1) // cw + 1
    sj = cw if nc == 1 else math.ceil((w - cw) / (nc - 1) - 1e-8)
    si = ch if nr == 1 else math.ceil((h - ch) / (nr - 1) - 1e-8)
    si, sj = si // scale * scale, sj // scale * scale
    out, i, last_i = [], 0, False
    while i < h and not last_i:
        j = 0
        if i + ch >= h:
            i, last_i = h - ch, True
        last_j = False
        while j < w and not last_j:
            if j + cw >= w:
                j, last_j = w - cw, True
            out.append((i, j))
            j += sj
        i += si
    return out


@pytest.mark.parametrize("h,w,ch,cw", [(100, 130, 64, 64), (64, 64, 64, 64), (257, 300, 128, 96), (90, 50, 40, 50)])
def test_tile_grid_indices(h, w, ch, cw):
    from lowlight_image_enhancement_b200.eval import TileGrid
    g = TileGrid(h, w, ch, cw)
    assert g.indices == _reference_grid(h, w, ch, cw)
    lq = torch.arange(h * w, dtype=torch.float32).view(1, 1, h, w)
    tiles = g.split(lq)
    assert tiles.shape == (len(g.indices), 1, g.crop_h, g.crop_w)
    for t, (i, j) in enumerate(g.indices):
        assert torch.equal(tiles[t, 0], lq[0, 0, i:i + g.crop_h, j:j + g.crop_w])


def _tile_golden():
    import os
    import numpy as np
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "tile_vectors.npz"))


def test_tile_grid_matches_reference_grids():
    """Tile origins and the tile batch against the reference's OWN `grids` (image_restoration_model.py:167-213), run
    unmodified by tests/golden/make_golden_tiles.py: absolute and ratio-defined crops, one tile, scale 2."""
    from lowlight_image_enhancement_b200.eval import TileGrid
    from tests.golden_inputs import TILE_CASES, tile_inputs
    gold = _tile_golden()
    for k, case in enumerate(TILE_CASES):
        g = TileGrid.from_val_opt(case["h"], case["w"], case["val"], case["scale"])
        assert [list(ij) for ij in g.indices] == gold[f"idx{k}"].tolist(), k
        lq, _ = tile_inputs(k)
        tiles = g.split(lq)
        assert list(tiles.shape) == gold[f"tiles_shape{k}"].tolist(), k
        assert float(tiles.double().sum()) == float(gold[f"tiles_sum{k}"]), k


@pytest.mark.gpu
def test_tile_merge_matches_reference_grids_inverse():
    """`TileGrid.merge` (llie_tile_merge, one kernel on the device) against the reference's own `grids_inverse`
    (image_restoration_model.py:215-241: host accumulation, count map, division): bit for bit."""
    from lowlight_image_enhancement_b200.eval import TileGrid
    from tests.golden_inputs import TILE_CASES, tile_inputs
    gold = _tile_golden()
    dev = torch.device("cuda:0")
    for k, case in enumerate(TILE_CASES):
        g = TileGrid.from_val_opt(case["h"], case["w"], case["val"], case["scale"])
        lq, outs_fn = tile_inputs(k)
        outs = outs_fn(len(g.indices), g.split(lq))
        got = g.merge(outs.to(dev))
        assert got.is_cuda and torch.equal(got.cpu(), torch.from_numpy(gold[f"merged{k}"])), k


@pytest.mark.gpu
def test_tile_merge_equals_host_average():
    """grids_inverse (irm:215-241) on the device: identical fp32 result to the host accumulation."""
    from lowlight_image_enhancement_b200.eval import TileGrid
    dev = torch.device("cuda:0")
    h, w = 257, 300
    g = TileGrid(h, w, 128, 96)
    outs = torch.rand(len(g.indices), 3, g.crop_h, g.crop_w, generator=torch.Generator().manual_seed(3))
    preds, cnt = torch.zeros(1, 3, h, w), torch.zeros(1, 1, h, w)
    for t, (i, j) in enumerate(g.indices):
        preds[0, :, i:i + g.crop_h, j:j + g.crop_w] += outs[t]
        cnt[0, 0, i:i + g.crop_h, j:j + g.crop_w] += 1.0
    ref = preds / cnt
    got = g.merge(outs.to(dev))
    assert got.is_cuda and torch.equal(got.cpu(), ref)


def _oracle_metrics(pred, gt):
    """All fourteen driver metrics of one pair from the CPU oracle (float64 where the reference is float64)."""
    from oracle import colour, quality
    pc, gc = pred.clamp(0, 1), gt.clamp(0, 1)                          # the colour adapters clamp (lowlight_metrics.py:236)
    rp = quality.rgb_psnr(pred, gt)
    # float64 evaluation: on smooth 8-bit images the fp32 eager path itself carries ~1e-5 of cancellation noise in
    # E[x^2] - mu^2 (parity section)
    rs = {k: quality.ssim_linear(pred[:, c:c + 1].double(), gt[:, c:c + 1].double())
          for c, k in enumerate(("R", "G", "B"))}
    rs["mean"] = sum(rs.values()) / 3
    out = {"linear_psnr": float(quality.calculate_psnr(gt, pred, 1.0)),
           "linear_ssim": float(quality.ssim_torchmetrics(pred, gt, compute_dtype=torch.float64).mean()),
           "cpsnr_rgb": float(quality.cpsnr_rgb(pred, gt)),
           "deltae2000_mean": colour.de00_summary(pc, gc, percentiles=())["mean"],
           "deltae2000_p95": colour.de00_summary(pc, gc, percentiles=(95.0,))["p95"],
           "edge_deltae2000_mean": colour.edge_stats(pc, gc, q=0.85)["mean"]}
    for k in ("R", "G", "B", "mean"):
        out[f"rgb_psnr_{k}"] = float(rp[k])
        out[f"rgb_ssim_{k}"] = float(rs[k])
    return out


@pytest.mark.gpu
def test_evaluate_pairs_all_metrics_vs_oracle():
    """the sharded driver (one-launch epilogue AND the generic path) against the ORACLE, not against the
    repo's own adapters, for every metric BASELINE configs[3] names plus the per-channel values.  Tolerances of
    BASELINE.json: 1e-5 relative PSNR / SSIM, 1e-4 absolute dE00; the edge mean allows mask ties on random floats."""
    from lowlight_image_enhancement_b200.eval import C4_METRICS, METRIC_FUNCTIONS, ShardedMeans, image_metrics
    from lowlight_image_enhancement_b200.eval import accumulate_image_metrics
    dev = torch.device("cuda:0")
    names = list(METRIC_FUNCTIONS)
    assert set(C4_METRICS) <= set(names) and len(names) == 14
    g = torch.Generator().manual_seed(2024)
    pairs, want = [], {k: 0.0 for k in names}
    for i in range(4):
        gt = torch.rand(1, 3, 48 + 4 * i, 64 + 8 * i, generator=g)     # W % 4 == 0: the TMA-ring SSIM kernels
        gt = torch.round(255 * torch.nn.functional.avg_pool2d(gt, 5, 1, 2)) / 255
        pred = torch.round(255 * (gt + 0.04 * torch.randn(gt.shape, generator=g)).clamp(0, 1)) / 255
        pairs.append((pred.to(dev), gt.to(dev)))
        for k, v in _oracle_metrics(pred, gt).items():
            want[k] += v / 4

    def check(res):
        assert res["count"] == 4 and res["nonfinite_inputs"] == 0
        for k in names:
            if k.startswith("deltae"):
                assert abs(res[k] - want[k]) < 1e-4, (k, res[k], want[k])
            elif k.startswith("edge"):
                assert abs(res[k] - want[k]) < 5e-3, (k, res[k], want[k])
            else:
                assert res[k] == pytest.approx(want[k], rel=1e-5), (k, res[k], want[k])

    check(evaluate_pairs(pairs, names, device=dev))                     # fast path: llie_eval_accumulate
    acc = ShardedMeans(names, dev)
    assert accumulate_image_metrics(pairs[0][0], pairs[0][1], names, acc)   # ... and it really is the fast path
    check(evaluate_pairs(pairs, names, device=dev, metric_fn=image_metrics))  # generic path, same kernels
    check({**evaluate_pairs(pairs, C4_METRICS, device=dev), **{k: want[k] for k in names if k not in C4_METRICS}})


def test_validation_metrics_option_handling():
    """ValidationMetrics reads opt['val']['metrics'] like irm:346-351, 420-428: a 'type' per name, unknown types fail
    the way `getattr(metric_module, type)` does."""
    from lowlight_image_enhancement_b200.eval import ValidationMetrics
    with pytest.raises(ValueError):
        ValidationMetrics({}, device="cpu")
    with pytest.raises(KeyError):
        ValidationMetrics({"psnr": {"data_range": 1.0}}, device="cpu")
    with pytest.raises(AttributeError):
        ValidationMetrics({"psnr": {"type": "calculate_niqe"}}, device="cpu")
    vm = ValidationMetrics({"PSNR": {"type": "linear_psnr", "data_range": 1.0}, "SSIM": {"type": "linear_ssim"}},
                           device="cpu")
    assert vm.names == ["PSNR", "SSIM"] and vm.types == ["linear_psnr", "linear_ssim"] and vm._fast
    assert not ValidationMetrics({"a": {"type": "linear_psnr"}, "b": {"type": "linear_psnr", "data_range": 2.0}},
                                 device="cpu")._fast


@pytest.mark.gpu
def test_validation_metrics_vs_oracle():
    """The shim for the metric section of dist_validation against the oracle: configured NAMES differ from the metric
    types (fast path, one-launch epilogue) and one type appears twice with different kwargs (generic path)."""
    from lowlight_image_enhancement_b200.eval import ValidationMetrics
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(77)
    opt = {"PSNR": {"type": "linear_psnr", "data_range": 1.0}, "SSIM": {"type": "linear_ssim", "data_range": 1.0},
           "CPSNR": {"type": "cpsnr_rgb"}, "RGB-SSIM": {"type": "rgb_ssim_mean"}, "dE": {"type": "deltae2000_mean"},
           "edge dE": {"type": "edge_deltae2000_mean", "q": 0.85}}
    twice = {"psnr1": {"type": "linear_psnr", "data_range": 1.0}, "psnr2": {"type": "linear_psnr", "data_range": 2.0}}
    vm, vm2 = ValidationMetrics(opt, dev), ValidationMetrics(twice, dev)
    want = {k: 0.0 for k in opt}
    want2 = {"psnr1": 0.0, "psnr2": 0.0}
    for i in range(3):
        gt = torch.rand(1, 3, 52 + 4 * i, 72, generator=g)
        gt = torch.round(255 * torch.nn.functional.avg_pool2d(gt, 5, 1, 2)) / 255
        pred = torch.round(255 * (gt + 0.04 * torch.randn(gt.shape, generator=g)).clamp(0, 1)) / 255
        vm.update(pred.to(dev), gt.to(dev))
        vm2.update(pred[0].to(dev), gt[0].to(dev))                     # [C, H, W] as `visuals['result']` may be
        o = _oracle_metrics(pred, gt)
        for name, cfg in opt.items():
            want[name] += o[cfg["type"]] / 3
        want2["psnr1"] += o["linear_psnr"] / 3
        want2["psnr2"] += (o["linear_psnr"] + 20.0 * math.log10(2.0)) / 3
    res, res2 = vm.reduce(), vm2.reduce()
    assert res["count"] == 3 and res2["count"] == 3 and res["nonfinite_inputs"] == 0
    for name in opt:
        tol = 1e-4 if name == "dE" else (5e-3 if name == "edge dE" else 1e-5 * abs(want[name]))
        assert abs(res[name] - want[name]) <= tol, (name, res[name], want[name])
    for name in want2:
        assert res2[name] == pytest.approx(want2[name], rel=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("srgb", [False, True])
def test_evaluate_phys_consistency_matches_metric_loop(srgb):
    """The dataset-level phys-consistency driver (the one product loop that makes the stream-overlap promise itself)
    against a plain loop over phys_cons_raw / phys_cons_srgb: serialised launches bit for bit, overlapped launches to
    1e-6 relative; iterators, host tensors and non-finite inputs are rejected."""
    from lowlight_image_enhancement_b200.eval import evaluate_phys_consistency
    from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw, phys_cons_srgb
    from oracle import psf_ops
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(5)
    psf = psf_ops.canonical_kernels("rgb").to(dev)
    batches = []
    for n in (4, 2, 8, 1, 3, 8):
        pred = torch.rand(n, 3, 128, 256, generator=g)
        rho = torch.rand(n, generator=g) + 0.5
        obs = (pred * rho.view(-1, 1, 1, 1) * 0.9 + 0.05 * torch.randn(pred.shape, generator=g)).clamp(0, 1.2)
        batches.append((pred.to(dev), obs.to(dev), rho.to(dev)))
    fn = phys_cons_srgb if srgb else phys_cons_raw
    per = torch.cat([fn(p, o, psf, r, reduction="none", crop="same", robust="charbonnier", eps=1e-3) for p, o, r in batches])
    want = float(per.double().mean())
    kw = dict(srgb=srgb, crop="same", robust="charbonnier", eps=1e-3)
    ser = evaluate_phys_consistency(batches, psf, overlap=False, **kw)
    ovl = evaluate_phys_consistency(batches, psf, overlap=True, **kw)
    assert ser["count"] == ovl["count"] == 26 and ser["nonfinite_inputs"] == 0
    assert ser["phys_cons_mean"] == pytest.approx(want, rel=1e-14)      # same float32 per-image values, float64 sums
    assert ovl["phys_cons_mean"] == pytest.approx(want, rel=1e-6)
    half = evaluate_phys_consistency(batches, psf, rank=1, world_size=2, **kw)        # shard of rank 1: batches 1, 3, 5
    assert half["count"] == 2 + 1 + 8
    with pytest.raises(TypeError):
        evaluate_phys_consistency(iter(batches), psf, **kw)
    with pytest.raises(ValueError):
        evaluate_phys_consistency([(batches[0][0].cpu(), batches[0][1].cpu(), 1.0)], psf, **kw)
    bad = batches[2][0].clone()
    bad[1, 0, 5, 5] = float("nan")
    with pytest.raises(ValueError):
        evaluate_phys_consistency(batches[:2] + [(bad, batches[2][1], batches[2][2])], psf, **kw)


@pytest.mark.gpu
def test_tile_merge_many_overlapping_tiles():
    """More than 64 tiles over one 8 x 128 output rectangle: the kernel's culled list overflows and it falls back to the
    scan over all tiles; arbitrary (unordered, duplicated) origins, same fp32 result as the host accumulation."""
    from lowlight_image_enhancement_b200.eval import TileGrid
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(11)
    h, w, th, tw = 40, 150, 24, 100
    grid = TileGrid(h, w, th, tw)
    grid.indices = [(int(torch.randint(0, h - th + 1, (1,), generator=g)), int(torch.randint(0, w - tw + 1, (1,), generator=g)))
                    for _ in range(90)] + [(0, 0), (h - th, w - tw), (0, w - tw), (h - th, 0)]
    grid._idx_dev = {}
    outs = torch.rand(len(grid.indices), 2, th, tw, generator=g)
    preds, cnt = torch.zeros(1, 2, h, w), torch.zeros(1, 1, h, w)
    for t, (i, j) in enumerate(grid.indices):
        preds[0, :, i:i + th, j:j + tw] += outs[t]
        cnt[0, 0, i:i + th, j:j + tw] += 1.0
    got = grid.merge(outs.to(dev))
    assert torch.equal(got.cpu(), preds / cnt)
