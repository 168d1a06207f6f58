"""Golden scalars for the BasicSR pixel losses, from the UNMODIFIED reference file
/root/reference/NAFNet_base/basicsr/models/losses/losses.py (build container only):

    python tests/golden/make_golden_basicsr.py

The reference package's `__init__` pulls in data/arch code with dependencies that are not installed, so the two loss
modules and the registry are loaded by file path under their own dotted names; nothing in them is modified.  Writes
tests/golden/basicsr_scalars.json: per case the loss value and three checksums of d loss / d pred (sum, sum of
squares, and a dot product with a seeded probe), float64-printed float32 results.
"""
from __future__ import annotations

import importlib.util
import json
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
BASE = "/root/reference/NAFNet_base/basicsr"


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


for pkg in ("basicsr", "basicsr.models", "basicsr.models.losses", "basicsr.utils"):
    m = types.ModuleType(pkg)
    m.__path__ = []
    sys.modules[pkg] = m
_load("basicsr.utils.registry", f"{BASE}/utils/registry.py")
_load("basicsr.models.losses.loss_util", f"{BASE}/models/losses/loss_util.py")
ref = _load("basicsr.models.losses.losses", f"{BASE}/models/losses/losses.py")

from tests.golden_inputs import PIXEL_LOSS_CASES, pixel_loss_inputs  # noqa: E402

torch.set_num_threads(4)
pred, tgt, weights = pixel_loss_inputs()
probe = torch.randn(pred.shape, generator=torch.Generator().manual_seed(72))
out = {}
for case in PIXEL_LOSS_CASES:
    p = pred.clone().requires_grad_(True)
    if case["loss"] == "psnr":
        loss = ref.PSNRLoss(loss_weight=0.5, toY=case["toY"])(p, tgt)
    else:
        cls = {"l1": ref.L1Loss, "mse": ref.MSELoss, "charbonnier": ref.CharbonnierLoss}[case["loss"]]
        kw = {"eps": case["eps"]} if "eps" in case else {}
        loss = cls(loss_weight=0.5, reduction=case["reduction"], **kw)(p, tgt, weights[case["weight"]])
    (g,) = torch.autograd.grad(loss, p)
    out[case["name"]] = dict(value=float(loss.detach()), grad_sum=float(g.double().sum()),
                             grad_sq=float((g.double() ** 2).sum()), grad_probe=float((g.double() * probe).sum()))
with open(os.path.join(HERE, "basicsr_scalars.json"), "w") as f:
    json.dump(out, f, indent=1, sort_keys=True)
print(json.dumps(out, indent=1))
