"""Accuracy of the dE00 metric kernel against a float64 evaluation of the reference formulas (oracle/colour.py in double,
on the GPU), next to the error of the float32 eager path itself.  Cases: the C4 synthetic pair (8-bit gt, noisy pred),
uniform random pairs, near-neutral pairs, dark pairs.  Prints mean / p99.9 / max abs error and the share above 1e-4.
Usage: PYTHONPATH=. python tools/de_accuracy.py [H W]"""
import sys
import torch, bench
from lowlight_image_enhancement_b200 import _functional as Fn
from oracle import colour

dev = torch.device("cuda:0")
H, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2848, 4256)


def de64(p, t):
    return colour.de00_metric_lab(colour.rgb_to_lab(p.double()), colour.rgb_to_lab(t.double()))


def de32(p, t):
    return colour.de00_metric_lab(colour.rgb_to_lab(p), colour.rgb_to_lab(t))


def stats(name, err, mask=None):
    e = err.flatten() if mask is None else err[mask]
    if e.numel() == 0:
        return
    k = max(1, int(e.numel() * 0.999))
    p999 = e.float().kthvalue(k).values.item()
    print(f"  {name:28s} n={e.numel():9d} mean {e.mean().item():.3e}  p99.9 {p999:.3e}  max {e.max().item():.3e}  "
          f">1e-4: {100.0 * (e > 1e-4).double().mean().item():.4f} %")


g = torch.Generator(device=dev).manual_seed(7)
cases = {}
cases["C4 synthetic pair"] = bench.synth_eval_pair(4000, dev, (1, 3, H, W))
a = torch.rand(1, 3, H, W, device=dev, generator=g)
cases["uniform random pair"] = (a, torch.rand(1, 3, H, W, device=dev, generator=g))
grey = torch.rand(1, 1, H, W, device=dev, generator=g).expand(1, 3, H, W)
cases["near-neutral pair"] = ((grey + 0.004 * torch.randn(1, 3, H, W, device=dev, generator=g)).clamp(0, 1).contiguous(),
                              (grey + 0.004 * torch.randn(1, 3, H, W, device=dev, generator=g)).clamp(0, 1).contiguous())
cases["dark pair (x 0.05)"] = (a * 0.05, (a * 0.05 + 0.002 * torch.randn(1, 3, H, W, device=dev, generator=g)).clamp(0, 1))
for name, (p, t) in cases.items():
    p, t = p.contiguous(), t.contiguous()
    ref = de64(p, t)
    _, _, dmap, lpl = Fn.deltae(p, t, variant=0, clamp_in=True, want_map=True, want_l=True)
    eager = de32(p, t)
    lab64 = colour.rgb_to_lab(p.double())
    chroma = torch.minimum(lab64[:, 1:].pow(2).sum(1).sqrt(), colour.rgb_to_lab(t.double())[:, 1:].pow(2).sum(1).sqrt())
    well = chroma > 1.0
    print(f"{name}: mean dE {ref.mean().item():.4f}")
    stats("kernel vs f64 (all)", (dmap.double() - ref).abs())
    stats("kernel vs f64 (chroma > 1)", (dmap.double() - ref).abs(), well)
    stats("eager f32 vs f64 (all)", (eager.double() - ref).abs())
    stats("eager f32 vs f64 (chroma > 1)", (eager.double() - ref).abs(), well)
    stats("kernel vs eager f32 (all)", (dmap - eager).abs().double())
    stats("L plane kernel vs f64", (lpl.double() - lab64[:, 0]).abs())
    print(f"  mean of map: kernel {dmap.double().mean().item():.7f}  f64 {ref.mean().item():.7f}  eager {eager.double().mean().item():.7f}")
    del ref, eager, dmap, lab64, chroma, well
    torch.cuda.empty_cache()
t = bench.timed_warm(lambda i: Fn.deltae(cases["C4 synthetic pair"][0], cases["C4 synthetic pair"][1], variant=0, clamp_in=True), 20, dev) / 20
t2 = bench.timed_warm(lambda i: Fn.deltae(cases["C4 synthetic pair"][0], cases["C4 synthetic pair"][1], variant=0, clamp_in=True, want_map=True, want_l=True, sqerr_out=[]), 20, dev) / 20
print(f"deltae_kernel: sums only {t*1e3:.1f} us, map + L + sqerr {t2*1e3:.1f} us per {H}x{W} pair")
