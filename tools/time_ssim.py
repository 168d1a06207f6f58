"""Time the SSIM kernels of the loaded library build: statistics on one full-resolution C4 pair (torchmetrics policy and
the five-map ssim_linear policy) and the DSSIM loss forward + backward at C3, for LLIE_TUNE_SSIM_PATH = 0 (TMA row ring +
packed FP32), 3 (round-1 two-column cp.async kernel) and 2 (one-column streaming)."""
import torch, bench
from lowlight_image_enhancement_b200 import _functional as Fn
from lowlight_image_enhancement_b200 import _native as nat
from lowlight_image_enhancement_b200.NewBP_model.losses import SSIMLoss

dev = torch.device("cuda:0")
pool = [bench.synth_eval_pair(4000 + i, dev) for i in range(3)]
win = Fn.window_1d("torchmetrics", 11, 1.5, dev)
x3 = [torch.rand(16, 3, 512, 512, device=dev) for _ in range(6)]
crit = SSIMLoss()


def stats(i, **kw):
    return Fn.ssim_stats(pool[i % 3][0], pool[i % 3][1], win, pad_mode=nat.PAD["reflect"], c1=1e-4, c2=9e-4, **kw)


def loss_fb(i):
    a = x3[i % 3].requires_grad_(True)
    crit(a, x3[3 + i % 3]).backward()
    a.grad = None


def loss_f(i):
    with torch.no_grad():
        crit(x3[i % 3], x3[3 + i % 3])


for path in (0, 3, 2):
    nat.check(nat.lib().llie_tune(3, path), "llie_tune")
    t_tm = bench.timed_warm(lambda i: stats(i, eps=0.0, interior=True), 30, dev) / 30
    t_lin = bench.timed_warm(lambda i: stats(i, eps=1e-12, clamp_var=True), 30, dev) / 30
    t_fb = bench.timed_warm(loss_fb, 30, dev) / 30
    t_f = bench.timed_warm(loss_f, 30, dev) / 30
    v = stats(0, eps=0.0, interior=True)[0].sum().item()
    print(f"path {path}: stats(4 maps) {t_tm*1e3:.1f} us  stats(5 maps) {t_lin*1e3:.1f} us  "
          f"dssim fwd(no grad) {t_f*1e3:.1f} us  dssim fwd+bwd {t_fb*1e3:.1f} us  sum {v:.6f}")
nat.check(nat.lib().llie_tune(3, 0), "llie_tune")
