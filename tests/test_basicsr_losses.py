"""BasicSR pixel losses: oracle vs the reference's golden scalars on CPU; the CUDA mirror
(`lowlight_image_enhancement_b200.basicsr_losses`) vs the golden scalars and vs the oracle's autograd on the GPU.
Tolerance: 1e-5 relative on values, 1e-5 of the gradient scale on gradients (float32 path)."""
from __future__ import annotations

import json
import os

import pytest
import torch

from oracle import pixel_losses as O
from tests.golden_inputs import PIXEL_LOSS_CASES, pixel_loss_inputs

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "basicsr_scalars.json")))
RTOL = 1e-5


def _probe(shape):
    return torch.randn(shape, generator=torch.Generator().manual_seed(72))


def _oracle_loss(case, p, tgt, weights):
    if case["loss"] == "psnr":
        return O.psnr_loss(p, tgt, loss_weight=0.5, toY=case["toY"])
    fn = {"l1": O.l1_loss, "mse": O.mse_loss, "charbonnier": O.charbonnier_loss}[case["loss"]]
    kw = {"eps": case["eps"]} if "eps" in case else {}
    return 0.5 * fn(p, tgt, weights[case["weight"]], reduction=case["reduction"], **kw)


def _mirror_loss(M, case, p, tgt, weights):
    if case["loss"] == "psnr":
        return M.PSNRLoss(loss_weight=0.5, toY=case["toY"])(p, tgt)
    cls = {"l1": M.L1Loss, "mse": M.MSELoss, "charbonnier": M.CharbonnierLoss}[case["loss"]]
    kw = {"eps": case["eps"]} if "eps" in case else {}
    return cls(loss_weight=0.5, reduction=case["reduction"], **kw)(p, tgt, weights[case["weight"]])


def _check_against_gold(name, loss, g):
    gold = GOLD[name]
    probe = _probe(g.shape)
    gd = g.detach().double().cpu()
    scale = max(abs(gold["grad_sq"]) ** 0.5, 1e-30)
    assert abs(float(loss) - gold["value"]) <= RTOL * max(abs(gold["value"]), 1e-12)
    assert abs(float((gd ** 2).sum()) - gold["grad_sq"]) <= 4 * RTOL * gold["grad_sq"]
    assert abs(float(gd.sum()) - gold["grad_sum"]) <= RTOL * scale * gd.numel() ** 0.5
    assert abs(float((gd * probe).sum()) - gold["grad_probe"]) <= RTOL * scale * gd.numel() ** 0.5


@pytest.mark.parametrize("case", PIXEL_LOSS_CASES, ids=[c["name"] for c in PIXEL_LOSS_CASES])
def test_oracle_matches_reference_golden(case):
    pred, tgt, weights = pixel_loss_inputs()
    p = pred.clone().requires_grad_(True)
    loss = _oracle_loss(case, p, tgt, weights)
    (g,) = torch.autograd.grad(loss, p)
    _check_against_gold(case["name"], loss.detach(), g)


def test_oracle_none_reduction_keeps_weighted_map():
    pred, tgt, weights = pixel_loss_inputs()
    m = O.l1_loss(pred, tgt, weights[1], reduction="none")
    assert m.shape == pred.shape
    assert torch.equal(m, (pred - tgt).abs() * weights[1])


@pytest.mark.gpu
@pytest.mark.parametrize("case", PIXEL_LOSS_CASES, ids=[c["name"] for c in PIXEL_LOSS_CASES])
def test_mirror_matches_reference_golden(case):
    from lowlight_image_enhancement_b200 import basicsr_losses as M
    pred, tgt, weights = pixel_loss_inputs()
    dev = torch.device("cuda:0")
    w = {k: (v.to(dev) if v is not None else None) for k, v in weights.items()}
    p = pred.to(dev).requires_grad_(True)
    loss = _mirror_loss(M, case, p, tgt.to(dev), w)
    (g,) = torch.autograd.grad(loss, p)
    _check_against_gold(case["name"], loss.detach().cpu(), g)


@pytest.mark.gpu
@pytest.mark.parametrize("case", PIXEL_LOSS_CASES, ids=[c["name"] for c in PIXEL_LOSS_CASES])
@pytest.mark.parametrize("shape", [(2, 3, 64, 48), (1, 3, 33, 31), (4, 1, 16, 40)], ids=["even", "ragged", "mono"])
def test_mirror_matches_oracle_elementwise(case, shape):
    """Element-wise gradient parity incl. planes with H*W % 4 != 0 (composed path) and C == 1."""
    from lowlight_image_enhancement_b200 import basicsr_losses as M
    if case["loss"] == "psnr" and case["toY"] and shape[1] != 3:
        pytest.skip("toY needs RGB")
    pred, tgt, _ = pixel_loss_inputs(shape)
    g = torch.Generator().manual_seed(73)
    weights = {None: None, 1: torch.rand(shape[0], 1, *shape[2:], generator=g), 3: torch.rand(*shape, generator=g)}
    p64 = pred.double().requires_grad_(True)
    ref = _oracle_loss(case, p64, tgt.double(), {k: (v.double() if v is not None else None) for k, v in weights.items()})
    (gref,) = torch.autograd.grad(ref, p64)
    dev = torch.device("cuda:0")
    p = pred.to(dev).requires_grad_(True)
    loss = _mirror_loss(M, case, p, tgt.to(dev), {k: (v.to(dev) if v is not None else None) for k, v in weights.items()})
    (gm,) = torch.autograd.grad(loss, p)
    assert abs(float(loss.detach()) - float(ref.detach())) <= RTOL * max(abs(float(ref)), 1e-12)
    err = (gm.double().cpu() - gref).abs()
    if case["loss"] == "l1":  # sign(d) flips where |d| is at float32 rounding level
        err = err[(pred.double() - tgt.double()).abs() > 1e-6]
    assert float(err.max()) <= RTOL * float(gref.abs().max())


@pytest.mark.gpu
def test_mirror_none_reduction_and_host_inputs():
    from lowlight_image_enhancement_b200 import basicsr_losses as M
    pred, tgt, weights = pixel_loss_inputs()
    m = M.l1_loss(pred, tgt, weights[1], reduction="none")      # host tensors in, host tensor out
    assert m.device.type == "cpu" and m.shape == pred.shape
    torch.testing.assert_close(m, (pred - tgt).abs() * weights[1], rtol=1e-6, atol=1e-7)
    with pytest.raises(ValueError):
        M.L1Loss(reduction="median")
    with pytest.raises(ValueError):
        M.mse_loss(pred, tgt[:, :1])


@pytest.mark.gpu
def test_pixel_loss_flags_nonfinite_and_accumulates():
    """C-ABI level: non-finite inputs raise the flag bits; accumulate=1 adds into an existing gradient."""
    from lowlight_image_enhancement_b200 import _native as nat
    dev = torch.device("cuda:0")
    x = torch.rand(2, 3, 16, 16, device=dev)
    y = torch.rand(2, 3, 16, 16, device=dev)
    sums = torch.zeros(1, dtype=torch.float64, device=dev)
    flags = torch.zeros(1, dtype=torch.int32, device=dev)
    grad = torch.ones_like(x)

    def call(xx, acc):
        rc = nat.lib().llie_pixel_loss_grad(1, nat.ptr(xx), nat.ptr(y), None, 0, None, 2, 3, 256, 0.0, 0.25, acc,
                                            nat.ptr(sums), nat.ptr(grad), nat.ptr(flags), nat.ptr(nat.workspace(dev)),
                                            nat.stream_ptr(dev))
        nat.check(rc, "llie_pixel_loss_grad")

    call(x, 1)
    torch.testing.assert_close(grad, 1.0 + 0.5 * (x - y), rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(sums[0].float(), ((x - y) ** 2).sum(), rtol=1e-5, atol=0)
    assert int(flags) == 0
    xb = x.clone()
    xb[1, 2, 3, 4] = float("nan")
    call(xb, 0)
    assert int(flags) & 1
