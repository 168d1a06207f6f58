"""llie_psnr_stats_static (include/llie_b200.h): the allocation-free variant of llie_psnr_stats whose scratch is a
file-scope `static __device__` array (no counterpart in the reference).
Called through the C ABI and compared with the existing entry point on two calls in a row (the array keeps the
previous call's contents).  Named test_zz_* so that pytest collects it LAST."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _call(static, desc, x, y, dev, nat):
    n, c = desc.n, desc.c
    out = torch.full((n + 2, 2 * c + 3), float("nan"), dtype=torch.float64, device=dev)
    mm = torch.zeros(2, dtype=x.dtype, device=dev)
    flags = torch.full((1,), -1, dtype=torch.int32, device=dev)
    st = nat.stream_ptr(dev)
    with torch.cuda.device(dev):
        if static:
            rc = nat.lib().llie_psnr_stats_static(desc, nat.ptr(x), nat.ptr(y), nat.ptr(out), nat.ptr(mm),
                                                  nat.ptr(flags), st)
        else:
            rc = nat.lib().llie_psnr_stats(desc, nat.ptr(x), nat.ptr(y), nat.ptr(out), nat.ptr(mm), nat.ptr(flags),
                                           nat.ptr(nat.workspace(dev)), st)
    nat.check(rc, "llie_psnr_stats_static" if static else "llie_psnr_stats")
    torch.cuda.synchronize(dev)
    return out.cpu(), mm.cpu(), int(flags.item())


def test_static_scratch_psnr_matches_workspace_entry_point_twice():
    from lowlight_image_enhancement_b200 import _native as nat
    assert torch.cuda.is_available()
    dev = torch.device("cuda:0")
    nat.ensure_device(dev)
    g = torch.Generator().manual_seed(20260)
    cases = [((2, 3, 96, 128), torch.float32), ((1, 3, 301, 517), torch.float32), ((3, 2, 40, 52), torch.float64)]
    for rep in range(2):                                   # two calls in a row per case: the array is reused as it was left
        for shape, dt in cases:
            n, c, h, w = shape
            x = torch.rand(shape, generator=g, dtype=dt).to(dev)
            y = (x + 0.05 * torch.randn(shape, generator=g, dtype=dt).to(dev)).contiguous()
            desc = nat.PsnrDesc(n, c, h * w, 0 if dt == torch.float32 else 1, 1, 0, 0.0, 1.0, 1e-12)
            ref, mm_ref, f_ref = _call(False, desc, x, y, dev, nat)
            got, mm_got, f_got = _call(True, desc, x, y, dev, nat)
            assert f_ref == f_got == 0
            assert torch.equal(mm_ref, mm_got)
            assert torch.isfinite(got).all()
            # same kernel, same grid (these sizes fit the static slot budget): float64 sums agree to rounding; the
            # suite's tolerance for PSNR values is 1e-5 relative, asserted here at 1e-12
            assert torch.allclose(got, ref, rtol=1e-12, atol=0), (shape, rep, (got - ref).abs().max())
    # non-finite input is flagged the same way
    xb = x.clone()
    xb[0, 0, 1, 1] = float("inf")
    assert _call(True, desc, xb, y, dev, nat)[2] & 1


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    from lowlight_image_enhancement_b200 import _native as nat
    dev = torch.device("cuda:0")
    n, c = 4100, 2                                          # 8200 planes > LLIE_PSNR_STATIC_MAX_SLOTS = 8192
    x = torch.zeros((n, c, 2, 2), device=dev)
    desc = nat.PsnrDesc(n, c, 4, 0, 1, 0, 0.0, 1.0, 1e-12)
    out = torch.zeros((n + 2, 2 * c + 3), dtype=torch.float64, device=dev)
    flags = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = nat.lib().llie_psnr_stats_static(desc, nat.ptr(x), nat.ptr(x), nat.ptr(out), None, nat.ptr(flags),
                                              nat.stream_ptr(dev))
    assert rc != 0
    msg = nat.lib().llie_last_error().decode()
    assert "exceed the static scratch" in msg and "8192" in msg
