"""Golden vectors for the decode / normalise step, from the UNMODIFIED reference code
(build container only):

    python tests/golden/make_golden_ingest.py

* `SonySIDLMDBDataset.__getitem__` of /root/reference/NAFNet_base/basicsr/data/sony_sid_lmdb_dataset.py runs as it is;
  only its storage is replaced: the module's imports of LMDB-side helpers are satisfied by stubs, the instance is
  created without `__init__` (which opens the LMDB) and `_fetch_png` returns the seeded uint16 arrays.  `img2tensor`
  is the reference's own (basicsr/utils/img_util.py, loaded by path).  The crop runs through the reference's
  `_maybe_random_crop` (square `patch_size` windows); its `rng` is a stand-in that returns the case's top / left.
* `_linear_to_srgb`, `_norm_to_lpips_range`, `_compute_channelwise_psnr_sse`, `_sobel_magnitude` are executed from the
  source text of /root/reference/metrics/evaluation.py (the module itself needs torchmetrics at import time): each
  `def` is cut out by `ast` and executed unchanged.

Writes tests/golden/ingest_vectors.npz.
"""
from __future__ import annotations

import ast
import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
REF = "/root/reference"
BASE = f"{REF}/NAFNet_base/basicsr"

from tests.golden_inputs import INGEST_CASES, ingest_inputs, srgb_inputs  # noqa: E402


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


for pkg in ("basicsr", "basicsr.data", "basicsr.utils"):
    m = types.ModuleType(pkg)
    m.__path__ = []
    sys.modules[pkg] = m
sys.modules["basicsr.data.transforms"] = types.SimpleNamespace(paired_random_crop=None)
sys.modules["basicsr.utils.sid_paths"] = types.SimpleNamespace(expand_with_sid_root=lambda *a, **k: None)
img_util = _load("basicsr.utils.img_util", f"{BASE}/utils/img_util.py")
sys.modules["basicsr.utils"].img2tensor = img_util.img2tensor
sys.modules["basicsr.utils"].FileClient = object
ds_mod = _load("basicsr.data.sony_sid_lmdb_dataset", f"{BASE}/data/sony_sid_lmdb_dataset.py")

vec = {}
for case in INGEST_CASES:
    short, long = ingest_inputs(case)
    ds = object.__new__(ds_mod.SonySIDLMDBDataset)
    ds.samples_per_pair = 1
    ds.entries = [dict(short_key="s", long_key="l", exposure_ratio=case["ratio"], pair_id="p0")]
    ds.return_metadata = False
    ds.phase = "train" if case["crop"] is not None else "val"
    ds.random_crop = True
    ds.patch_size = None
    if case["crop"] is not None:
        top, left, ch, cw = case["crop"]
        assert ch == cw, "the reference crops square patches"
        ds.patch_size = ch
        ds.rng = types.SimpleNamespace(integers=lambda lo, hi, _it=iter((top, left)): next(_it))
    ds._fetch_png = lambda key, client_key, s=short, l=long: s if client_key == "short" else l
    sample = ds[0]
    for k in ("lq", "gt", "short_raw", "long_raw", "short_obs", "expo_ratio"):
        vec[f"{case['name']}_{k}"] = sample[k].numpy()

src = open(f"{REF}/metrics/evaluation.py").read()
ns = {"torch": torch}
wanted = {"_linear_to_srgb", "_norm_to_lpips_range", "_compute_channelwise_psnr_sse", "_sobel_magnitude"}
for node in ast.parse(src).body:
    if isinstance(node, ast.FunctionDef) and node.name in wanted:
        exec(compile(ast.Module([node], []), f"{REF}/metrics/evaluation.py", "exec"), ns)
ns["kornia_sobel"] = None  # kornia is not installed: the reference takes its manual-Sobel branch
x = srgb_inputs()
s = ns["_linear_to_srgb"](x)
vec["srgb_out"] = s.numpy()
vec["lpips_out"] = ns["_norm_to_lpips_range"](s).numpy()
y = (x + 0.05 * torch.randn(x.shape, generator=torch.Generator().manual_seed(312))).clamp(0, 1)
vec["sse_out"] = ns["_compute_channelwise_psnr_sse"](x.clamp(0, 1), y).numpy()
vec["sobel_out"] = ns["_sobel_magnitude"](s.mean(dim=1, keepdim=True)).numpy()
np.savez_compressed(os.path.join(HERE, "ingest_vectors.npz"), **vec)
print({k: v.shape for k, v in vec.items()})
