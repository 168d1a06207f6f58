"""GPU parity: fused phys-consistency kernels and the NewBP operator vs the CPU oracle and the
reference-generated golden vectors.  Tolerances from BASELINE.json: 1e-5 relative for fp32 losses
and gradients."""
import warnings

import numpy as np
import pytest
import torch

from oracle import psf_ops
from tests.golden_inputs import PHYS_CASES, dense_from_spec, depthwise_from_spec, loss_inputs, phys_inputs

pytestmark = pytest.mark.gpu
RTOL = 1e-5


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _rel(a, b):
    a = a.detach().double().cpu()
    b = torch.as_tensor(b).double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _run_case(case, psf, dev):
    from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw, phys_cons_srgb
    pred, obs, rho = phys_inputs(case)
    fn = phys_cons_srgb if case["srgb"] else phys_cons_raw
    kw = dict(reduction=case["reduction"], padding=case["padding"], crop=case["crop"], robust=case["robust"],
              eps=case["eps"], return_map=True, normalize_psf=case.get("normalize_psf", True))
    if case["srgb"]:
        kw["clamp01"] = case["clamp01"]
    rho_d = rho.to(dev) if torch.is_tensor(rho) else rho
    return fn(pred.to(dev), obs.to(dev), psf.to(dev), rho_d, **kw)


@pytest.mark.parametrize("i", range(len(PHYS_CASES)))
def test_phys_metric_vs_golden_dense(golden, dev, i):
    """Dense [3,3,kh,kw] PSFs exactly as the reference accepts them."""
    val, mp = _run_case(PHYS_CASES[i], dense_from_spec(PHYS_CASES[i]["psf"]), dev)
    assert _rel(val, golden[f"phys{i}_val"]) < RTOL
    np.testing.assert_allclose(mp.cpu().numpy(), golden[f"phys{i}_map"], atol=2e-6, rtol=0)


@pytest.mark.parametrize("i", [k for k, c in enumerate(PHYS_CASES) if c["psf"][0] == "diag"])
def test_phys_metric_depthwise_form(golden, dev, i):
    """Same cases with the PSF given in the depthwise [3,1,kh,kw] form (superset of the reference)."""
    val, mp = _run_case(PHYS_CASES[i], depthwise_from_spec(PHYS_CASES[i]["psf"][1]), dev)
    assert _rel(val, golden[f"phys{i}_val"]) < RTOL
    np.testing.assert_allclose(mp.cpu().numpy(), golden[f"phys{i}_map"], atol=2e-6, rtol=0)


@pytest.mark.parametrize("padding", ["reflect", "replicate", "zeros"])
@pytest.mark.parametrize("crop", ["valid", "same"])
@pytest.mark.parametrize("robust,eps", [("none", 1e-12), ("charbonnier", 1e-3)])
@pytest.mark.parametrize("shape", [(2, 3, 64, 128), (1, 3, 70, 252), (3, 3, 33, 8)])
def test_phys_tile_path_fwd_bwd(dev, padding, crop, robust, eps, shape):
    """Tile path (depthwise 3x3, W % 4 == 0): forward value, map and the adjoint gradient vs autograd of
    the oracle.  Shapes straddle tile edges (120-wide tiles, 32-high)."""
    from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_srgb
    g = torch.Generator().manual_seed(sum(shape) + len(padding))
    pred = torch.rand(shape, generator=g)
    k = depthwise_from_spec("B2")
    rho = torch.tensor([0.9, 1.1, 1.3][: shape[0]])
    obs = (psf_ops.crosstalk_forward(pred, k) * rho.view(-1, 1, 1, 1) + 0.05 * torch.randn(shape, generator=g)).clamp(0, 1.2)
    dense = psf_ops.embed_dense(k)
    kw = dict(padding=padding, crop=crop, robust=robust, eps=eps, reduction="mean")
    xr = pred.clone().requires_grad_(True)
    with torch.enable_grad():
        ref = _oracle_diff(xr, obs, dense, rho, clamp01=True, **kw)
        ref.backward()
    x = pred.to(dev).requires_grad_(True)
    val, mp = phys_cons_srgb(x.detach(), obs.to(dev), k.to(dev), rho.to(dev), return_map=True, **kw)
    assert _rel(val, ref) < RTOL
    out = phys_cons_srgb(x, obs.to(dev), k.to(dev), rho.to(dev), differentiable=True, **kw)
    assert out.requires_grad
    out.backward()
    assert _rel(out, ref) < RTOL
    if robust == "none":
        assert _rel(x.grad, xr.grad) < RTOL
    else:
        # Charbonnier with eps=1e-3 is ill-conditioned where |r| ~ eps (d psi/dr = 1/eps): fp32 eager is
        # itself ~1e-4 away from an fp64 run of the same oracle.  Judge both against fp64.
        x64 = pred.double().requires_grad_(True)
        _oracle_diff(x64, obs.double(), dense.double(), rho.double(), clamp01=True, dtype=torch.float64, **kw).backward()
        ours, eager = _rel(x.grad, x64.grad), _rel(xr.grad, x64.grad)
        assert ours < max(RTOL, 3.0 * eager), (ours, eager)


def _oracle_diff(pred, obs, dense, rho, *, padding, crop, robust, eps, reduction, clamp01, dtype=torch.float32):
    """Differentiable restatement (the oracle's phys_metric detaches like the reference does)."""
    import torch.nn.functional as F
    k = psf_ops.prepare_dense_psf(dense, 3, 3, dtype, True, False, eps)
    ph, pw = k.shape[-2] // 2, k.shape[-1] // 2
    if padding == "zeros":
        y = F.conv2d(pred, k, padding=(ph, pw))
    else:
        y = F.conv2d(F.pad(pred, (pw, pw, ph, ph), mode=padding), k)
    y = y * psf_ops.expand_rho(rho, y)
    if clamp01:
        y = y.clamp(0.0, 1.0)
    o = obs
    if crop == "valid":
        y, o = y[..., ph:-ph, pw:-pw], o[..., ph:-ph, pw:-pw]
    d = y - o
    lm = torch.sqrt(d * d + eps * eps) if robust == "charbonnier" else d.abs()
    per = lm.flatten(1).mean(1)
    return per.mean(0) if reduction == "mean" else per.sum(0)


@pytest.mark.parametrize("padding", ["reflect", "replicate", "zeros"])
@pytest.mark.parametrize("crop", ["valid", "same"])
@pytest.mark.parametrize("ksize", [3, 5, 7])
@pytest.mark.parametrize("shape", [(2, 3, 8, 7), (1, 3, 19, 515), (2, 3, 37, 50), (1, 3, 9, 257), (3, 3, 64, 6)])
def test_phys_generic_kernels_interior_and_border(dev, padding, crop, ksize, shape):
    """Generic kernels (W % 4 != 0 keeps the row ring out): the interior sweep in row batches (3x3, 5x5), the
    per-row interior path of partial units, the compact border phase and the run-time-size body (7x7) against autograd
    of the oracle -- value, map and gradient; planes smaller than one unit, more than two strips, H not a multiple of
    the unit height, planes that are all border."""
    from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw
    if crop == "valid" and min(shape[-2:]) <= 2 * (ksize // 2):
        pytest.skip("valid crop leaves nothing")
    g = torch.Generator().manual_seed(sum(shape) * 7 + ksize + len(padding))
    pred = torch.rand(shape, generator=g)
    k = torch.rand(3, 1, ksize, ksize, generator=g) + 0.05
    k = k / k.sum((-2, -1), keepdim=True)
    rho = torch.tensor([0.9, 1.1, 1.3][: shape[0]])
    dense = psf_ops.embed_dense(k)
    obs = (pred * rho.view(-1, 1, 1, 1) + 0.05 * torch.randn(shape, generator=g)).clamp(0, 1.2)
    kw = dict(padding=padding, crop=crop, robust="none", eps=1e-12, reduction="mean")
    xr = pred.clone().requires_grad_(True)
    ref = _oracle_diff(xr, obs, dense, rho, clamp01=False, **kw)
    ref.backward()
    x = pred.to(dev).requires_grad_(True)
    val, mp = phys_cons_raw(x.detach(), obs.to(dev), k.to(dev), rho.to(dev), return_map=True, **kw)
    assert _rel(val, ref) < RTOL
    ph = ksize // 2
    import torch.nn.functional as F
    kd = psf_ops.prepare_dense_psf(dense, 3, 3, torch.float32, True, False, 1e-12)
    y = F.conv2d(pred, kd, padding=ph) if padding == "zeros" else F.conv2d(F.pad(pred, (ph, ph, ph, ph), mode=padding), kd)
    want = (y * rho.view(-1, 1, 1, 1) - obs).abs()
    if crop == "valid":
        want = want[..., ph:-ph, ph:-ph]
    assert mp.shape == want.shape and float((mp.cpu() - want).abs().max()) < 2e-6
    out = phys_cons_raw(x, obs.to(dev), k.to(dev), rho.to(dev), differentiable=True, **kw)
    out.backward()
    assert _rel(out, ref) < RTOL
    # |r| is sign-like: a residual at the fp32 rounding level of the tap sum may take the other sign in the oracle
    bad = ((x.grad.cpu() - xr.grad).abs() > 2e-5 * float(xr.grad.abs().max())).float().mean()
    assert float(bad) < 2e-3, float(bad)


@pytest.mark.parametrize("case_i", [5, 6, 7, 8])
def test_phys_generic_path_gradient(dev, case_i):
    """Generic path (dense / 5x5 / 3x5 / spatial rho): gradient vs oracle autograd."""
    from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw, phys_cons_srgb
    case = PHYS_CASES[case_i]
    pred, obs, rho = phys_inputs(case)
    dense = dense_from_spec(case["psf"])
    xr = pred.clone().requires_grad_(True)
    red = "mean" if case["reduction"] == "none" else case["reduction"]
    ref = _oracle_diff(xr, obs, dense, rho, padding=case["padding"], crop=case["crop"], robust=case["robust"],
                       eps=case["eps"], reduction=red, clamp01=case["clamp01"])
    ref.backward()
    x = pred.to(dev).requires_grad_(True)
    fn = phys_cons_srgb if case["srgb"] else phys_cons_raw
    kw = dict(padding=case["padding"], crop=case["crop"], robust=case["robust"], eps=case["eps"], reduction=red,
              differentiable=True)
    if case["srgb"]:
        kw["clamp01"] = case["clamp01"]
    out = fn(x, obs.to(dev), dense.to(dev), rho.to(dev) if torch.is_tensor(rho) else rho, **kw)
    out.backward()
    assert _rel(out, ref) < RTOL
    assert _rel(x.grad, xr.grad) < 2e-5


@pytest.mark.parametrize("mode", ["mono", "rgb"])
def test_crosstalk_newbp_and_losses_vs_golden(golden, dev, mode):
    from lowlight_image_enhancement_b200.NewBP_model.losses import PhysicalConsistencyLossSRGB, PhysicsConsistencyLoss
    from lowlight_image_enhancement_b200.NewBP_model.newbp_layer import CrosstalkPSF, NewBPFunction, build_psf_kernels
    k = build_psf_kernels(mode, "P2" if mode == "mono" else "B2")
    np.testing.assert_array_equal(k.numpy(), golden[f"kern_{mode}"])
    psf = CrosstalkPSF(mode, k).to(dev)
    np.testing.assert_allclose(psf.kernel.cpu().numpy(), golden[f"kern_{mode}_norm"], rtol=1e-7)
    bhat, a, rho = loss_inputs(mode)
    gy = torch.randn(bhat.shape, generator=torch.Generator().manual_seed(77))
    x = bhat.to(dev).requires_grad_(True)
    y = psf(x)
    y.backward(gy.to(dev))
    assert _rel(y, golden[f"xtalk_{mode}_y"]) < RTOL
    assert _rel(x.grad, golden[f"xtalk_{mode}_gx"]) < RTOL
    kd = psf.kernel.expand(3, 1, 3, 3).contiguous()
    x2 = bhat.to(dev).requires_grad_(True)
    y2 = NewBPFunction.apply(x2, kd, 1, 3)
    y2.backward(gy.to(dev))
    assert _rel(y2, golden[f"newbp_{mode}_fwd"]) < RTOL
    assert _rel(x2.grad, golden[f"newbp_{mode}_gx"]) < RTOL
    # fused sRGB physics loss: value and gradient
    x3 = bhat.to(dev).requires_grad_(True)
    ls = PhysicalConsistencyLossSRGB(psf)(x3, a.to(dev), rho.to(dev))
    ls.backward()
    assert _rel(ls, golden[f"physloss_srgb_{mode}_val"]) < RTOL
    assert _rel(x3.grad, golden[f"physloss_srgb_{mode}_grad"]) < RTOL
    # unfused route (forward hook present) must agree
    calls = []
    h = psf.register_forward_hook(lambda *_: calls.append(1))
    x4 = bhat.to(dev).requires_grad_(True)
    ls2 = PhysicalConsistencyLossSRGB(psf)(x4, a.to(dev), rho.to(dev))
    ls2.backward()
    h.remove()
    assert len(calls) == 1
    assert _rel(ls2, golden[f"physloss_srgb_{mode}_val"]) < RTOL
    assert _rel(x4.grad, golden[f"physloss_srgb_{mode}_grad"]) < RTOL
    # RAW physics loss, 3x3 and 5x5
    for tag, kk in (("k3", k), ("k5", torch.ones(k.shape[0], 1, 5, 5) / 25.0)):
        x5 = bhat.to(dev).requires_grad_(True)
        lr = PhysicsConsistencyLoss(kk, device=dev)(x5, a.to(dev), rho.view(-1).to(dev))
        lr.backward()
        assert _rel(lr, golden[f"physloss_raw_{mode}_{tag}_val"]) < RTOL
        assert _rel(x5.grad, golden[f"physloss_raw_{mode}_{tag}_grad"]) < RTOL


def test_newbp_double_precision_and_gradcheck(dev):
    """core_tests/test_physics_loss_grad.py semantics: fp64 adjoint == conv_transpose2d; gradcheck."""
    import torch.nn.functional as F
    from lowlight_image_enhancement_b200.NewBP_model.newbp_layer import CrosstalkPSF
    torch.manual_seed(0)
    for mode in ("mono", "rgb"):
        k = torch.rand(1 if mode == "mono" else 3, 1, 3, 3, dtype=torch.double)
        psf = CrosstalkPSF(mode, k).to(dev).to(torch.double)
        z = torch.randn(1, 3, 6, 6, dtype=torch.double, device=dev, requires_grad=True)
        ref = torch.randn(1, 3, 6, 6, dtype=torch.double, device=dev)
        assert torch.autograd.gradcheck(lambda t: (psf(t) - ref).abs().sum(), (z,), eps=1e-6, atol=1e-4)
        y = psf(z)
        ga = torch.randn_like(y)
        (gz,) = torch.autograd.grad(y, z, ga)
        kd = psf.kernel.expand(3, -1, -1, -1).contiguous().cpu()
        want = F.conv_transpose2d(ga.cpu(), kd, padding=1, groups=3)
        assert torch.allclose(gz.cpu(), want, atol=1e-12)


def test_reference_contract_errors_and_flags(dev):
    """test_phys_consistency.py of the reference: ValueErrors, zero-PSF warning, NaN detection, shapes."""
    from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw, phys_cons_srgb
    x = torch.rand(1, 3, 24, 24, device=dev)
    with pytest.raises(ValueError):
        phys_cons_raw(x, x, torch.ones(3, 3, 4, 4, device=dev), 1.0)
    with pytest.raises(ValueError):
        phys_cons_raw(x, x, torch.ones(2, 3, 3, 3, device=dev), 1.0)
    with pytest.warns(RuntimeWarning):
        v = phys_cons_raw(x, x, torch.zeros(3, 3, 3, 3, device=dev), 1.0)
    assert torch.isfinite(v)
    bad = x.clone()
    bad[0, 1, 0, 0] = float("nan")  # on the cropped border: still must be detected
    with pytest.raises(ValueError, match="pred"):
        phys_cons_raw(bad, x, dense_from_spec(("delta", 3)).to(dev), 1.0)
    with pytest.raises(ValueError, match="obs"):
        phys_cons_raw(x, bad, dense_from_spec(("delta", 3)).to(dev), 1.0)
    # delta PSF, obs = rho * gt -> ~0
    v = phys_cons_raw(x, 0.7 * x, dense_from_spec(("delta", 3)).to(dev), 0.7, crop="same")
    assert abs(float(v)) <= 1e-7
    _, m = phys_cons_raw(x, x, dense_from_spec(("delta", 5)).to(dev), 1.0, crop="valid", padding="zeros", return_map=True)
    assert tuple(m.shape) == (1, 3, 20, 20)
    # clamp caps the residual
    p9, o8 = torch.full((1, 3, 24, 24), 0.9, device=dev), torch.full((1, 3, 24, 24), 0.8, device=dev)
    _, m = phys_cons_srgb(p9, o8, dense_from_spec(("delta", 3)).to(dev), 1.5, crop="same", return_map=True)
    assert float(m.max()) <= 0.25 + 1e-6
    # host tensors are uploaded and the result comes back on the host
    out = phys_cons_raw(x.cpu(), x.cpu(), dense_from_spec(("delta", 3)), 1.0)
    assert out.device.type == "cpu"


def test_full_size_properties(dev):
    """BASELINE C2 size (8x3x512x512): linearity in rho and the delta-PSF identity, size-independent."""
    from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw
    g = torch.Generator(device=dev).manual_seed(102)
    pred = torch.rand(8, 3, 512, 512, device=dev, generator=g)
    k = depthwise_from_spec("B2").to(dev)
    zero = torch.zeros_like(pred)
    rho = torch.tensor([100., 250., 300., 100., 250., 300., 100., 250.], device=dev)
    a = phys_cons_raw(pred / 300, zero, k, rho, reduction="none", crop="same", padding="zeros")
    b = phys_cons_raw(pred / 300, zero, k, 2 * rho, reduction="none", crop="same", padding="zeros")
    assert torch.allclose(2 * a, b, rtol=1e-6)
    ident = torch.zeros(3, 1, 3, 3, device=dev)
    ident[:, 0, 1, 1] = 1
    v = phys_cons_raw(pred, pred * 0.5, ident, 0.5, crop="same")
    assert float(v) <= 1e-7
    # deterministic: bitwise identical on repeat
    a2 = phys_cons_raw(pred / 300, zero, k, rho, reduction="none", crop="same", padding="zeros")
    assert torch.equal(a, a2)


@pytest.mark.parametrize("padding", ["reflect", "replicate", "zeros"])
@pytest.mark.parametrize("crop", ["valid", "same"])
@pytest.mark.parametrize("robust,clip", [(False, (False, False)), (True, (True, False)), (False, (False, True))])
@pytest.mark.parametrize("shape", [(2, 3, 64, 128), (1, 3, 70, 252), (3, 3, 33, 8), (8, 3, 512, 512), (2, 3, 37, 50)])
def test_phys_value_and_grad_equals_two_passes(dev, padding, crop, robust, clip, shape):
    """llie_phys_value_and_grad (one fused pass) returns exactly what llie_phys_forward + llie_phys_backward do:
    the loss vector to 1e-6 relative (different but fixed reduction grouping), the gradient bit for bit."""
    from lowlight_image_enhancement_b200 import _functional as Fn
    from lowlight_image_enhancement_b200 import _native as nat
    g = torch.Generator().manual_seed(sum(shape))
    n = shape[0]
    pred = torch.rand(shape, generator=g).to(dev)
    obs = (pred * 0.9 + 0.05 * torch.randn(shape, generator=g).to(dev)).clamp(0, 1.1).contiguous()
    k = depthwise_from_spec("B2").to(dev)
    psf = Fn.PreparedPSF(k.contiguous(), Fn.PSF_DEPTHWISE, 3, 3, False)
    rho = torch.tensor([0.9, 1.1, 1.3, 1.0, 0.8, 1.2, 0.7, 1.05][:n], device=dev)
    cfg = Fn.PhysCfg(nat.PAD[padding], crop == "valid", robust, 1e-3, 0 if clip[0] else 1, clip[0], clip[1])
    gout = torch.rand(n, generator=g).to(dev)
    vec, flags, _ = Fn.phys_forward(pred, obs, psf, rho, 0.0, (1, 0, 0, 0), cfg, False)
    grad = Fn.phys_backward(pred, obs, psf, rho, 0.0, (1, 0, 0, 0), cfg, gout, 0.37)
    vec2, flags2, grad2 = Fn.phys_value_and_grad(pred, obs, psf, rho, 0.0, (1, 0, 0, 0), cfg, gout, 0.37)
    assert torch.equal(grad, grad2)
    assert torch.allclose(vec, vec2, rtol=1e-6, atol=0)
    assert int(flags) == int(flags2)


@pytest.mark.parametrize("shape", [(8, 3, 512, 512), (2, 3, 96, 640), (4, 3, 130, 252)])
def test_phys_stream_overlap_is_bit_identical(dev, shape):
    """llie_phys_desc.stream_overlap, set per call by `with Fn.phys_stream_overlap():` (launch i + 1 waits for launch i only before it touches the shared workspace, and work
    items grow to ~128 rows): a mixed back-to-back sequence of fused / forward / backward launches over rotating,
    disjoint buffer sets must return what fully serialised launches return -- gradients bit for bit, loss vectors to
    1e-6 relative (the item partition, hence the grouping of the float32 partial sums, differs between the modes) --
    on a stream and replayed from a CUDA graph."""
    from lowlight_image_enhancement_b200 import _functional as Fn
    from lowlight_image_enhancement_b200 import _native as nat
    g = torch.Generator().manual_seed(7 + sum(shape))
    n = shape[0]
    pool = 5
    preds = [torch.rand(shape, generator=g).to(dev) for _ in range(pool)]
    obss = [(p * 0.9 + 0.05 * torch.randn(shape, generator=g).to(dev)).clamp(0, 1.1).contiguous() for p in preds]
    grads = [torch.empty(shape, device=dev) for _ in range(pool)]
    psf = Fn.PreparedPSF(depthwise_from_spec("B2").to(dev).contiguous(), Fn.PSF_DEPTHWISE, 3, 3, False)
    rho = torch.tensor([0.9, 1.1, 1.3, 1.0, 0.8, 1.2, 0.7, 1.05][:n], device=dev)
    cfg = Fn.PhysCfg(nat.PAD["reflect"], True, True, 1e-3, 0, True, False)
    gout = torch.rand(n, generator=g).to(dev)
    st = (1, 0, 0, 0)

    def sequence(rounds):
        outs = []
        for i in range(rounds * pool):
            b = i % pool
            kind = i % 3
            if kind == 0:
                vec, flags, gr = Fn.phys_value_and_grad(preds[b], obss[b], psf, rho, 0.0, st, cfg, gout, 0.37, out=grads[b])
                outs.append((vec.clone(), gr.clone()))
            elif kind == 1:
                vec, flags, _ = Fn.phys_forward(preds[b], obss[b], psf, rho, 0.0, st, cfg, False)
                outs.append((vec.clone(),))
            else:
                gr = Fn.phys_backward(preds[b], obss[b], psf, rho, 0.0, st, cfg, gout, 0.37, out=grads[b])
                outs.append((gr.clone(),))
        return outs

    def same(a, b):  # loss vectors are 1-D [n + 2], gradients 4-D
        return torch.equal(a, b) if a.dim() == 4 else torch.allclose(a, b, rtol=1e-6, atol=0)

    ref = sequence(2)
    torch.cuda.synchronize(dev)
    want = {}
    for i in range(pool):
        v, _, gr = Fn.phys_value_and_grad(preds[i], obss[i], psf, rho, 0.0, st, cfg, gout, 0.37)
        want[i] = (v.clone(), gr.clone())
    torch.cuda.synchronize(dev)
    keep = []

    def burst():
        for i in range(4 * pool):
            b = i % pool
            vec, _, _ = Fn.phys_value_and_grad(preds[b], obss[b], psf, rho, 0.0, st, cfg, gout, 0.37, out=grads[b])
            keep.append(vec)

    with Fn.phys_stream_overlap():
        assert Fn._desc(cfg, preds[0], shape[1], psf, 0.0, st).stream_overlap == 1
        # (the .clone() kernels sit between the phys launches here: they read what the launch before them wrote, and
        # they are ordinary launches, so every phys kernel still waits for them -- this leg checks the flag's plumbing)
        got = sequence(2)
        torch.cuda.synchronize(dev)
        for r, o in zip(ref, got):
            assert all(same(a, b) for a, b in zip(r, o))
        # back-to-back phys launches only (the overlap proper): stream, then graph replay
        burst()
        torch.cuda.synchronize(dev)
        first = [v.clone() for v in keep]
        for i, v in enumerate(keep):
            assert same(v, want[i % pool][0]), i
        for b in range(pool):
            assert torch.equal(grads[b], want[b][1]), b
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            keep.clear()
            burst()  # warm-up on the capture stream (workspace of that stream)
            side.synchronize()
            keep.clear()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                burst()
            for b in range(pool):
                grads[b].zero_()
            for _ in range(3):
                graph.replay()
        side.synchronize()
        for i, v in enumerate(keep):
            assert torch.equal(v, first[i]), i  # overlapped launches are reproducible bit for bit
        for b in range(pool):
            assert torch.equal(grads[b], want[b][1]), b
    # the scope is per call: nothing leaks out of the `with` block
    assert Fn._desc(cfg, preds[0], shape[1], psf, 0.0, st).stream_overlap == 0


def test_phys_stream_overlap_scope_is_per_thread_and_skips_autograd(dev, monkeypatch):
    """The overlap promise is a field of each call's descriptor, set from a thread-local scope: another thread and the
    autograd engine's backward threads (whose upstream gradient comes from the previous kernel) never carry it."""
    import threading
    from lowlight_image_enhancement_b200 import _functional as Fn
    from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw
    seen = []
    real = Fn._desc

    def spy(*a, **k):
        d = real(*a, **k)
        seen.append((threading.current_thread() is threading.main_thread(), d.stream_overlap))
        return d

    monkeypatch.setattr(Fn, "_desc", spy)
    g = torch.Generator().manual_seed(3)
    pred = torch.rand(2, 3, 64, 64, generator=g).to(dev).requires_grad_(True)
    obs = torch.rand(2, 3, 64, 64, generator=g).to(dev)
    psf = psf_ops.canonical_kernels("rgb").to(dev)
    with Fn.phys_stream_overlap():
        out = phys_cons_raw(pred, obs, psf, 1.2, differentiable=True)
        out.backward()
        t = threading.Thread(target=lambda: phys_cons_raw(pred.detach(), obs, psf, 1.2))
        t.start()
        t.join()
    assert (True, 1) in seen                       # forward in the scope's thread
    assert all(ov == 0 for main, ov in seen if not main) and any(not main for main, _ in seen)
    assert pred.grad is not None and torch.isfinite(pred.grad).all()


@pytest.mark.parametrize("shape", [(2, 3, 64, 256), (1, 3, 13, 8), (4, 3, 50, 516), (3, 1, 24, 260)])
def test_newbp_depthwise3_fast_path_equals_generic(dev, shape):
    """The four-column depthwise 3x3 kernel (16-byte aligned planes, W % 4 == 0) against the generic one-thread-per-output
    kernel, reached through a view whose base address breaks the alignment: forward and adjoint bit for bit, and both
    against F.conv2d / F.conv_transpose2d (newbp_layer.py:9-21)."""
    import torch.nn.functional as F
    from lowlight_image_enhancement_b200.NewBP_model.newbp_layer import NewBPFunction
    g = torch.Generator().manual_seed(sum(shape))
    n, c, h, w = shape
    x = torch.randn(shape, generator=g).to(dev)
    k = (torch.rand(c, 1, 3, 3, generator=g) - 0.3).to(dev)
    gy = torch.randn(shape, generator=g).to(dev)
    store = torch.empty(x.numel() + 1, device=dev)
    x_un = store[1:].view(shape)           # 4-byte aligned only: the generic kernel takes it
    x_un.copy_(x)
    outs = []
    for inp in (x, x_un):
        t = inp.detach().requires_grad_(True)
        y = NewBPFunction.apply(t, k, 1, c)
        y.backward(gy)
        outs.append((y.detach().clone(), t.grad.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    ref = F.conv2d(x.double(), k.double(), padding=1, groups=c)
    rgx = F.conv_transpose2d(gy.double(), k.double(), padding=1, groups=c)
    assert _rel(outs[0][0], ref.cpu()) < 1e-6 and _rel(outs[0][1], rgx.cpu()) < 1e-6
