"""GPU parity: DeltaE00Loss / SSIMLoss / HybridLossPlus (value + analytic gradient) vs autograd through the
CPU oracle.  Tolerance from BASELINE.json: 1e-5 relative for fp32 losses and gradients; a gradient is
compared max-norm-wise against the float64 run of the oracle.  dE00 gradients are ill-conditioned where
the two colours are close (dE ~ 1: every term is a difference of nearly equal fp32 numbers) or neutral
(1 / (a^2 + b^2) amplification); there fp32 autograd of the oracle itself sits 6e-6 .. 5e-4 from float64,
and the kernel has to stay within a small factor (5x) of that noise: its Lab values carry ~1.5 ulp
(cbrtf 1 ulp, x^2.4 < 1 ulp, FMA-contracted matrix products) against glibc's correctly rounded ~1 ulp, and
a = 500 (f(X) - f(Y)) turns one ulp of f into 1.5e-5 of a, which dC = C2' - C1' ~ 0.3 amplifies to 1e-4
relative on single pixels (measured: mean |error| equal to fp32 autograd's, 99.9th percentile 1.6x)."""
import pytest
import torch

from oracle import colour, hybrid, quality

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _rel(a, b):
    return float((a.detach().double().cpu() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def _oracle_grad(fn, x, y, dt):
    xx = x.to(dt).clone().requires_grad_(True)
    v = fn(xx, y.to(dt))
    v.backward()
    return v.detach(), xx.grad


def _images(shape, seed, spread=0.1):
    g = torch.Generator().manual_seed(seed)
    tgt = torch.rand(shape, generator=g)
    gen = tgt + spread * torch.randn(shape, generator=g)     # leaves [0,1]: exercises the clamp masks
    return gen, tgt


def _check_grad(mine, g32, g64, tol=1e-5):
    e_mine, e_ref = _rel(mine, g64), _rel(g32, g64)
    assert e_mine <= max(tol, 5.0 * e_ref), (e_mine, e_ref)
    return e_mine, e_ref


@pytest.mark.parametrize("shape", [(2, 3, 37, 53), (1, 3, 128, 96), (3, 3, 8, 8)])
def test_deltae_loss_value_and_grad(dev, shape):
    from lowlight_image_enhancement_b200.NewBP_model.losses import DeltaE00Loss
    gen, tgt = _images(shape, shape[2])
    gen[0, :, :3, :3] = 0.5                                    # neutral grey pixels
    gen[0, :, 3:5, :3] = tgt[0, :, 3:5, :3]                    # identical pixels
    v32, g32 = _oracle_grad(colour.de00_loss, gen, tgt, torch.float32)
    v64, g64 = _oracle_grad(colour.de00_loss, gen, tgt, torch.float64)
    x = gen.to(dev).requires_grad_(True)
    v = DeltaE00Loss()(x, tgt.to(dev))
    assert v.dtype == torch.float32 and v.ndim == 0 and v.requires_grad
    (3.0 * v).backward()
    assert abs(float(v) - float(v64)) <= 1e-5 * float(v64)
    _check_grad(x.grad / 3.0, g32, g64)
    # pixels pushed outside [0,1] get exactly zero gradient on the clamped channels
    outside = (gen < 0) | (gen > 1)
    assert outside.any() and float(x.grad.cpu()[outside].abs().max()) == 0.0
    # no-grad call takes the value-only kernel and agrees
    with torch.no_grad():
        v2 = DeltaE00Loss()(gen.to(dev), tgt.to(dev))
    assert abs(float(v2) - float(v)) <= 1e-6 * float(v)


def test_deltae_loss_contracts(dev):
    """core_tests/test_losses_domain_norm.py:143-146: identical images give ~sqrt(eps), below 0.01."""
    from lowlight_image_enhancement_b200.NewBP_model.losses import DeltaE00Loss
    x = torch.rand(2, 3, 32, 32, device=dev)
    v = DeltaE00Loss()(x, x.clone())
    assert 0.0 < float(v) < 0.01 and abs(float(v) - 1e-3) < 1e-5
    with pytest.raises(ValueError):
        DeltaE00Loss()(x[:, :2], x[:, :2])
    host = torch.rand(1, 3, 16, 16, requires_grad=True)       # host tensors: result and grad come back to the host
    out = DeltaE00Loss()(host, torch.rand(1, 3, 16, 16))
    out.backward()
    assert out.device.type == "cpu" and host.grad is not None and host.grad.device.type == "cpu"
    # black pixels sit on atan2's singular point (a = b = 0): ATen's backward returns 0 there, so the
    # gradient stays finite and equal to the oracle's
    blk = torch.zeros(1, 3, 4, 4)
    tg = torch.rand(1, 3, 4, 4)
    _, gref = _oracle_grad(colour.de00_loss, blk, tg, torch.float32)
    xb = blk.to(dev).requires_grad_(True)
    DeltaE00Loss()(xb, tg.to(dev)).backward()
    assert torch.isfinite(gref).all() and torch.isfinite(xb.grad).all()
    assert _rel(xb.grad, gref) < 1e-5


@pytest.mark.parametrize("shape,ws", [((2, 3, 37, 53), 11), ((1, 3, 128, 96), 11), ((2, 1, 11, 11), 11),
                                      ((1, 3, 33, 130), 7), ((1, 2, 70, 64), 3), ((1, 1, 16, 15), 15),
                                      ((1, 3, 12, 200), 5),
                                      # windows above 15 taps: runtime-window tile kernels (forward and adjoint)
                                      ((1, 3, 40, 97), 17), ((1, 2, 70, 140), 21), ((1, 1, 64, 66), 31),
                                      ((1, 1, 80, 130), 63)])
def test_ssim_loss_value_and_grad(dev, shape, ws):
    from lowlight_image_enhancement_b200.NewBP_model.losses import SSIMLoss

    def f(a, b):
        return quality.ssim_loss_kornia(a, b, ws, 1.0)

    gen, tgt = _images(shape, shape[3] + ws)
    v32, g32 = _oracle_grad(f, gen, tgt, torch.float32)
    v64, g64 = _oracle_grad(f, gen, tgt, torch.float64)
    x = gen.to(dev).requires_grad_(True)
    v = SSIMLoss(window_size=ws)(x, tgt.to(dev))
    (0.5 * v).backward()
    assert abs(float(v) - float(v64)) <= 1e-5 * float(v64)
    _check_grad(2.0 * x.grad, g32, g64)
    outside = (gen < 0) | (gen > 1)
    assert float(x.grad.cpu()[outside].abs().max()) == 0.0


def test_ssim_loss_reductions_and_identity(dev):
    from lowlight_image_enhancement_b200.NewBP_model.losses import SSIMLoss
    gen, tgt = _images((2, 3, 40, 44), 5)
    xg = gen.clamp(0, 1)
    smap = torch.clamp((1 - quality.kornia_ssim_map(xg.double(), tgt.double())) / 2, 0, 1)
    x = gen.to(dev).requires_grad_(True)
    m = SSIMLoss(reduction="none")(x, tgt.to(dev))
    assert m.shape == gen.shape and float((m.cpu().double() - smap).abs().max()) < 2e-6
    wgt = torch.rand(gen.shape, generator=torch.Generator().manual_seed(1))
    (m * wgt.to(dev)).sum().backward()
    xr = gen.double().clone().requires_grad_(True)
    (torch.clamp((1 - quality.kornia_ssim_map(xr.clamp(0, 1), tgt.double())) / 2, 0, 1) * wgt.double()).sum().backward()
    assert _rel(x.grad, xr.grad) < 1e-5
    s = SSIMLoss(reduction="sum")(gen.to(dev), tgt.to(dev))
    assert abs(float(s) - float(smap.sum())) < 1e-5 * float(smap.sum())
    ident = SSIMLoss()(tgt.to(dev), tgt.to(dev))
    assert abs(float(ident)) < 1e-6
    with pytest.raises(ValueError):
        SSIMLoss(window_size=4)
    with pytest.raises(RuntimeError):
        SSIMLoss()(torch.rand(1, 3, 5, 40, device=dev), torch.rand(1, 3, 5, 40, device=dev))


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("mode", ["raw_kernel", "srgb_psf"])
def test_hybrid_loss_plus_vs_oracle(dev, mode, fused):
    """losses.py:318-372 with the perceptual / LPIPS terms absent (out of scope): total, logs, gradient."""
    from lowlight_image_enhancement_b200.NewBP_model.losses import HybridLossPlus
    from lowlight_image_enhancement_b200.NewBP_model.newbp_layer import CrosstalkPSF, build_psf_kernels
    g = torch.Generator().manual_seed(31)
    n = 2
    B = torch.rand(n, 3, 48, 56, generator=g)
    Bhat = (B + 0.08 * torch.randn(B.shape, generator=g))
    rho = torch.tensor([100.0, 250.0]).view(n, 1, 1, 1)
    A = (B / rho + 0.0005 * torch.randn(B.shape, generator=g)).clamp_min(0)
    kern = build_psf_kernels("rgb", "B2")
    kw = dict(w_l1=1.0, w_de=0.02, w_ssim=0.05, w_phys=0.05)
    if mode == "raw_kernel":
        ora_kw = dict(physics_kernel=kern)
        mod = HybridLossPlus(device="cuda", w_phys=0.05, physics_kernel=kern, fused=fused, perceptual_module=False)
    else:
        psf = CrosstalkPSF("rgb", kern)
        ora_kw = dict(psf_kernel=psf.kernel.clone())
        mod = HybridLossPlus(device="cuda", w_phys=0.05, physics_psf_module=psf, fused=fused, perceptual_module=False)

    def ora(dt):
        x = Bhat.to(dt).clone().requires_grad_(True)
        xs = x.clamp(0, 1)
        kws = {k: v.to(dt) for k, v in ora_kw.items()}
        tot, logs = hybrid.hybrid_loss(Bhat_raw=x, B_raw=B.to(dt), A_raw=A.to(dt), expo_ratio=rho.to(dt), Bhat_srgb01=xs,
                                       B_srgb01=B.to(dt), A_srgb01=A.to(dt), **kw, **kws)
        tot.backward()
        return tot.detach(), logs, x.grad

    t32, l32, g32 = ora(torch.float32)
    t64, l64, g64 = ora(torch.float64)
    x = Bhat.to(dev).requires_grad_(True)
    xs = x.clamp(0, 1)                                             # irm:289-300: srgb tensor = clamp of the raw one
    tot, logs = mod(Bhat_raw=x, B_raw=B.to(dev), A_raw=A.to(dev), expo_ratio=rho.to(dev), Bhat_srgb01=xs,
                    B_srgb01=B.to(dev), A_srgb01=A.to(dev))
    tot.backward()
    assert set(logs) >= {"L1_raw", "Perc", "DeltaE", "SSIM", "Phys", "Total"}
    for k in ("L1_raw", "DeltaE", "SSIM", "Phys", "Total"):
        assert abs(float(logs[k]) - float(l64[k])) <= 1e-5 * abs(float(l64[k])), k
        assert not logs[k].requires_grad
    assert float(logs["Perc"]) == 0.0
    assert abs(float(tot) - float(t64)) <= 1e-5 * float(t64)
    _check_grad(x.grad, g32, g64)


def test_hybrid_loss_plus_behaviour(dev):
    from lowlight_image_enhancement_b200.NewBP_model.losses import HybridLossPlus
    g = torch.Generator().manual_seed(3)
    B = torch.rand(1, 3, 32, 32, generator=g).to(dev)
    x = (B + 0.05).clone().requires_grad_(True)
    rho = torch.tensor([100.0], device=dev)
    kern = torch.tensor([[.01, .02, .01], [.02, .88, .02], [.01, .02, .01]]).view(1, 1, 3, 3)
    args = dict(B_raw=B, A_raw=B / 100.0, expo_ratio=rho, B_srgb01=B)
    # uncertainty weighting: val * exp(-2 s) + s, learnable log-sigmas receive gradients (losses.py:308-316)
    mod = HybridLossPlus(device="cuda", use_uncertainty=True, physics_kernel=kern, perceptual_module=False).to(dev)
    tot, logs = mod(Bhat_raw=x, Bhat_srgb01=x.clamp(0, 1), **args)
    tot.backward()
    assert all(mod.log_sigma[k].grad is not None for k in ("l1", "de", "ssim", "phys"))
    ref = sum(float(logs[k]) for k in ("L1_raw", "DeltaE", "SSIM", "Phys"))
    assert abs(float(tot) - ref) < 1e-5 * ref                         # s = 0 -> plain sum
    # switches
    mod2 = HybridLossPlus(device="cuda", use_deltaE=False, use_ssim=False, use_phys=False, perceptual_module=False)
    tot2, logs2 = mod2(Bhat_raw=x, Bhat_srgb01=x, **args)
    assert "DeltaE" not in logs2 and "SSIM" not in logs2 and "Phys" not in logs2
    assert abs(float(tot2) - float(logs2["L1_raw"])) < 1e-7
    # injected perceptual module is used and weighted
    class Half(torch.nn.Module):
        def forward(self, a, b):
            return (a - b).pow(2).mean()
    mod3 = HybridLossPlus(device="cuda", use_deltaE=False, use_ssim=False, use_phys=False, w_perc=0.5,
                          perceptual_module=Half())
    tot3, logs3 = mod3(Bhat_raw=x, Bhat_srgb01=x, **args)
    assert abs(float(tot3) - (float(logs3["L1_raw"]) + 0.5 * float(logs3["Perc"]))) < 1e-7 and float(logs3["Perc"]) > 0
    # non-finite term -> RuntimeError naming it (losses.py:298-306)
    bad = x.detach().clone()
    bad[0, 0, 0, 0] = float("nan")
    with pytest.raises(RuntimeError, match="non-finite values in term 'L1_raw'"):
        HybridLossPlus(device="cuda", physics_kernel=kern, perceptual_module=False)(Bhat_raw=bad, Bhat_srgb01=bad, **args)


def test_legacy_hybrid_loss(dev):
    """losses.py:72-89: (total, l1, perceptual) with an injected perceptual module; L1 gradient = sign / numel."""
    from lowlight_image_enhancement_b200.NewBP_model.losses import HybridLoss
    g = torch.Generator().manual_seed(2)
    a, b = torch.rand(2, 3, 33, 47, generator=g), torch.rand(2, 3, 33, 47, generator=g)

    class Perc(torch.nn.Module):
        def forward(self, x, y):
            return (x - y).pow(2).mean()

    x = a.to(dev).requires_grad_(True)
    tot, l1, pv = HybridLoss(1.0, 0.1, "cuda", perceptual_module=Perc())(x, b.to(dev))
    tot.backward()
    xr = a.clone().requires_grad_(True)
    ref = torch.nn.functional.l1_loss(xr, b) + 0.1 * (xr - b).pow(2).mean()
    ref.backward()
    assert abs(float(tot) - float(ref)) < 1e-6 and abs(float(l1) - float(torch.nn.functional.l1_loss(a, b))) < 1e-7
    assert _rel(x.grad, xr.grad) < 1e-6
    tot2, _, p2 = HybridLoss(device="cuda", perceptual_module=False)(x.detach(), b.to(dev))
    assert float(p2) == 0.0 and abs(float(tot2) - float(l1)) < 1e-7


def test_upstream_scalar_is_applied_in_place_and_survives_retained_graphs(dev):
    """The gradient buffers the value + gradient kernels fill in forward are scaled by autograd's upstream scalar in
    place (a no-op launch for 1.0): any scalar, and repeated backward passes through a retained graph, must give
    scale x the unit gradient -- for the single-term losses and for the fused HybridLossPlus node (two buffers)."""
    from lowlight_image_enhancement_b200.NewBP_model.losses import DeltaE00Loss, HybridLossPlus
    g = torch.Generator().manual_seed(91)
    b = torch.rand(2, 3, 40, 48, generator=g).to(dev)
    x0 = (b + 0.05 * torch.randn(b.shape, generator=g).to(dev)).clamp(0, 1)

    def unit_and_scaled(make_loss, scales):
        x = x0.clone().requires_grad_(True)
        make_loss(x).backward()
        unit = x.grad.clone()
        x = x0.clone().requires_grad_(True)
        loss = make_loss(x)
        got = []
        for i, s in enumerate(scales):
            x.grad = None
            (loss * s).backward(retain_graph=i + 1 < len(scales))
            got.append(x.grad.clone())
        return unit, got

    de = DeltaE00Loss()
    kern = torch.tensor([[.01, .02, .01], [.02, .88, .02], [.01, .02, .01]], device=dev).view(1, 1, 3, 3)
    hyb = HybridLossPlus(device=dev, w_phys=0.05, physics_kernel=kern, check_finite=False, perceptual_module=False)
    rho = torch.ones(2, 1, 1, 1, device=dev)
    for make in (lambda x: de(x, b),
                 lambda x: hyb(Bhat_raw=x, B_raw=b, A_raw=b * 0.9, expo_ratio=rho, Bhat_srgb01=x, B_srgb01=b)[0]):
        scales = [1.0, 2.5, -0.75, 1.0]
        unit, got = unit_and_scaled(make, scales)
        assert float(unit.abs().max()) > 0
        for s, gr in zip(scales, got):
            # (the hybrid node returns two buffers that autograd adds: a*s + b*s against (a + b)*s)
            assert float((gr - unit * s).abs().max()) <= 2e-6 * abs(s) * float(unit.abs().max()), s


@pytest.mark.parametrize("shape", [(2, 3, 64, 300), (1, 3, 257, 513), (1, 1, 40, 11), (1, 3, 33, 47), (16, 3, 512, 512),
                                   (1, 2, 12, 260), (2, 1, 23, 1024), (1, 3, 700, 264)])
def test_two_column_dssim_kernels_vs_one_column(dev, shape):
    """SSIMLoss forward (coefficient planes) + backward on the default kernels for the 11-tap window (TMA row ring +
    packed FP32 when W % 4 == 0, else the two-column cp.async kernels) and on the two-column kernels (path 3) against
    the one-column streaming kernels (LLIE_TUNE_SSIM_PATH = 2): same loss to 1e-6 relative, same gradient to 2e-6 of
    its scale (four moment maps instead of five in forward; identical tap order in backward).  The small-H shapes put
    the top and bottom mirror folds into one row segment, W = 260 / 264 a partial strip next to a full one."""
    from lowlight_image_enhancement_b200 import _native as nat
    from lowlight_image_enhancement_b200.NewBP_model.losses import SSIMLoss
    g = torch.Generator().manual_seed(sum(shape))
    b = torch.rand(shape, generator=g).to(dev)
    x0 = (b + 0.08 * torch.randn(shape, generator=g).to(dev)).clamp(-0.05, 1.05)
    crit = SSIMLoss()
    out = {}
    try:
        for path in (2, 3, 0):
            nat.check(nat.lib().llie_tune(3, path), "llie_tune")
            x = x0.clone().requires_grad_(True)
            loss = crit(x, b)
            loss.backward()
            out[path] = (loss.detach().clone(), x.grad.clone())
    finally:
        nat.check(nat.lib().llie_tune(3, 0), "llie_tune")
    l2, g2 = out[2]
    for path in (3, 0):
        l0, g0 = out[path]
        assert abs(float(l0) - float(l2)) <= 1e-6 * abs(float(l2)), path
        assert float((g0 - g2).abs().max()) <= 2e-6 * float(g2.abs().max()), path
    assert float(g2.abs().max()) > 0


def test_hybrid_fused_terms_and_weights_are_read_per_call(dev):
    """the fused node must not freeze the set of present terms or the weights on the first call.  The
    reference allows A_srgb01=None on some batches (image_restoration_model.py:300) and reads self.w on every forward
    (losses.py:308-316)."""
    from lowlight_image_enhancement_b200.NewBP_model.losses import HybridLossPlus
    from lowlight_image_enhancement_b200.NewBP_model.newbp_layer import create_crosstalk_psf
    g = torch.Generator().manual_seed(77)
    B = torch.rand(2, 3, 40, 48, generator=g).to(dev)
    x = (B + 0.05 * torch.randn(B.shape, generator=g).to(dev)).clamp(0, 1)
    A = (B * 0.9).clamp(0, 1)
    rho = torch.ones(2, 1, 1, 1, device=dev)
    mod = HybridLossPlus(device="cuda", w_phys=0.10, physics_psf_module=create_crosstalk_psf("rgb", "B2"),
                         perceptual_module=False)
    kw = dict(Bhat_raw=x, B_raw=B, A_raw=A, expo_ratio=rho, Bhat_srgb01=x, B_srgb01=B)

    def expect(logs, with_phys):
        t = mod.w["l1"] * float(logs["L1_raw"]) + mod.w["de"] * float(logs["DeltaE"]) + mod.w["ssim"] * float(logs["SSIM"])
        return t + (mod.w["phys"] * float(logs["Phys"]) if with_phys else 0.0)

    for with_phys in (True, False, True, False):                      # alternate: present set changes between calls
        tot, logs = mod(**kw, A_srgb01=A if with_phys else None)
        assert ("Phys" in logs) == with_phys
        assert torch.isfinite(tot) and abs(float(tot) - expect(logs, with_phys)) <= 1e-6 * abs(float(tot))
    mod.w["ssim"] = 0.5                                               # a weight schedule edits the dict in place
    mod.w["phys"] = 0.0
    tot, logs = mod(**kw, A_srgb01=A)
    assert abs(float(tot) - expect(logs, True)) <= 1e-6 * abs(float(tot))
    # gradient follows the new weights too: d total / dx with w_ssim 0.5 differs from the w_ssim 0.05 one
    xg = x.clone().requires_grad_(True)
    mod(**{**kw, "Bhat_raw": xg, "Bhat_srgb01": xg}, A_srgb01=A)[0].backward()
    g_new = xg.grad.clone()
    mod.w["ssim"] = 0.05
    xg = x.clone().requires_grad_(True)
    mod(**{**kw, "Bhat_raw": xg, "Bhat_srgb01": xg}, A_srgb01=A)[0].backward()
    assert float((g_new - xg.grad).abs().max()) > 0


def test_retained_graph_after_zero_upstream_fails_loudly(dev):
    """backward twice through a retained graph with an upstream scalar of 0 first must not hand out inf / NaN
    gradients (the in-place scaled buffer is gone): RuntimeError.  Double backward is refused (once_differentiable)."""
    from lowlight_image_enhancement_b200.NewBP_model.losses import DeltaE00Loss
    g = torch.Generator().manual_seed(5)
    b = torch.rand(1, 3, 24, 24, generator=g).to(dev)
    x = (b + 0.05).clamp(0, 1).requires_grad_(True)
    loss = DeltaE00Loss()(x, b)
    (loss * 0.0).backward(retain_graph=True)
    assert float(x.grad.abs().max()) == 0.0
    with pytest.raises(RuntimeError, match="retained graph"):
        (loss * 2.0).backward()
    x2 = (b + 0.05).clamp(0, 1).requires_grad_(True)
    (g1,) = torch.autograd.grad(DeltaE00Loss()(x2, b), x2, create_graph=True)
    with pytest.raises(RuntimeError):
        g1.sum().backward()


@pytest.mark.parametrize("in_dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_losses_under_autocast(dev, in_dtype):
    """BASELINE configs[4] / image_restoration_model.py:255-310: the reference calls the hybrid loss inside
    `autocast`.  Every loss class must (i) compute in fp32 whatever the autocast state, (ii) return an fp32 scalar,
    (iii) equal the plain fp32 call on the up-cast inputs to 1e-6, (iv) hand back a gradient in the input's dtype."""
    from lowlight_image_enhancement_b200.NewBP_model.losses import (DeltaE00Loss, HybridLossPlus,
                                                                   PhysicalConsistencyLossSRGB, PhysicsConsistencyLoss,
                                                                   SSIMLoss)
    from lowlight_image_enhancement_b200.NewBP_model.newbp_layer import build_psf_kernels, create_crosstalk_psf
    g = torch.Generator().manual_seed(123)
    B = torch.rand(2, 3, 48, 64, generator=g).to(dev)
    X = (B + 0.05 * torch.randn(B.shape, generator=g).to(dev)).clamp(0, 1).to(in_dtype)
    A = (B * 0.8).clamp(0, 1)
    rho = torch.ones(2, 1, 1, 1, device=dev)
    kern = build_psf_kernels("rgb", "B2")
    hyb = HybridLossPlus(device="cuda", w_phys=0.05, physics_kernel=kern, perceptual_module=False)
    cases = {
        "de": lambda x: DeltaE00Loss()(x, B),
        "ssim": lambda x: SSIMLoss()(x, B),
        "phys_raw": lambda x: PhysicsConsistencyLoss(kern)(x, A, rho),
        "phys_srgb": lambda x: PhysicalConsistencyLossSRGB(create_crosstalk_psf("rgb", "B2").to(dev))(x, A, rho),
        "hybrid": lambda x: hyb(Bhat_raw=x, B_raw=B, A_raw=A, expo_ratio=rho, Bhat_srgb01=x.clamp(0, 1), B_srgb01=B)[0],
    }
    for name, fn in cases.items():
        xa = X.clone().requires_grad_(True)
        with torch.autocast("cuda", dtype=torch.bfloat16 if in_dtype != torch.float16 else torch.float16):
            va = fn(xa)
        va.backward()
        xf = X.float().clone().requires_grad_(True)
        vf = fn(xf)
        vf.backward()
        assert va.dtype == torch.float32 and va.ndim == 0, name
        assert abs(float(va) - float(vf)) <= 1e-6 * abs(float(vf)), (name, float(va), float(vf))
        assert xa.grad.dtype == in_dtype, name
        # a half-precision leaf accumulates its two incoming gradients (raw + sRGB branch of the hybrid loss) in ITS
        # dtype, so the comparison with the rounded fp32 gradient allows two ulps of that dtype
        ref = xf.grad.to(in_dtype).float()
        tol = {torch.float32: 1e-6, torch.bfloat16: 2.0 ** -7, torch.float16: 2.0 ** -9}[in_dtype]
        assert float((xa.grad.float() - ref).abs().max()) <= tol * float(ref.abs().max()) + 1e-12, name


@pytest.mark.gpu
@pytest.mark.parametrize("present", [(True, True, True), (False, True, False), (True, False, True), (False, False, False)])
def test_hybrid_combine_equals_the_eager_expression(present):
    """llie_hybrid_combine (term values + weighted total of the fused hybrid loss in one launch) against the scalar
    tensor expression it replaces; absent terms are NaN in the term vector and left out of the total."""
    from lowlight_image_enhancement_b200 import _native as nat
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(11)
    use_de, use_ssim, use_phys = present
    l1 = torch.rand(1, generator=g, dtype=torch.float64).to(dev) * 1e5
    des = (torch.rand(37, generator=g, dtype=torch.float64) * 1e6).to(dev)
    sss = (torch.rand(37 * 3, generator=g, dtype=torch.float64) * 1e5).to(dev)
    vec = torch.rand(39, generator=g).to(dev)
    w = [1.0, 0.02, 0.05, 0.07]
    counts = (12345678.0, 4194304.0, 12582912.0)
    terms = torch.empty(4, device=dev)
    total = torch.empty((), device=dev)
    wv = (nat.C.c_float * 4)(*w)
    pv = vec[37:38]
    with torch.cuda.device(dev):
        rc = nat.lib().llie_hybrid_combine(nat.ptr(l1), counts[0], nat.ptr(des) if use_de else None, 37 if use_de else 0,
                                           counts[1], nat.ptr(sss) if use_ssim else None, 111 if use_ssim else 0,
                                           counts[2], nat.ptr(pv) if use_phys else None, wv, nat.ptr(terms),
                                           nat.ptr(total), nat.stream_ptr(dev))
    nat.check(rc, "llie_hybrid_combine")
    want = [(l1[0] / counts[0]).float(), (des.sum() / counts[1]).float(), (sss.sum() / counts[2]).float(), vec[37]]
    have = [True, use_de, use_ssim, use_phys]
    t = terms.cpu()
    ref_total = torch.zeros((), dtype=torch.float32)
    for i in range(4):
        if have[i]:
            assert abs(float(t[i]) - float(want[i])) <= 1.2e-7 * abs(float(want[i]))
            ref_total = ref_total + want[i].cpu() * torch.tensor(w[i], dtype=torch.float32)
        else:
            assert torch.isnan(t[i])
    assert abs(float(total) - float(ref_total)) <= 2.4e-7 * abs(float(ref_total))
