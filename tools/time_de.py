"""Time the dE00 metric / loss-gradient kernels and the select passes of the loaded library build (A/B variants)."""
import torch, bench
from lowlight_image_enhancement_b200 import _functional as Fn
from lowlight_image_enhancement_b200.eval import ShardedMeans, accumulate_image_metrics
dev = torch.device("cuda:0")
pool = [bench.synth_eval_pair(4000 + i, dev) for i in range(3)]
t = bench.timed_warm(lambda i: Fn.deltae(pool[i % 3][0], pool[i % 3][1], variant=0, clamp_in=True), 30, dev) / 30
names = ["linear_psnr", "linear_ssim", "deltae2000_mean", "edge_deltae2000_mean"]
acc = ShardedMeans(names, dev)
ta = bench.timed_warm(lambda i: accumulate_image_metrics(pool[i % 3][0], pool[i % 3][1], names, acc), 30, dev) / 30
x = [torch.rand(16, 3, 512, 512, device=dev) for _ in range(6)]
def lg(i):
    a = x[i % 3].requires_grad_(True)
    Fn.DeltaE00LossFn.apply(a, x[3 + i % 3], 1e-12)  # value + gradient in the forward launch


tg = bench.timed_warm(lg, 30, dev) / 30
xo = [torch.rand(16, 3, 511, 513, device=dev) for _ in range(6)]    # odd plane: the one-pixel-per-thread kernels


def lgo(i):
    a = xo[i % 3].requires_grad_(True)
    Fn.DeltaE00LossFn.apply(a, xo[3 + i % 3], 1e-12)


tgo = bench.timed_warm(lgo, 30, dev) / 30
print(f"loss_grad scalar path (16x3x511x513) {tgo*1e3:.1f} us")
print(f"deltae {t*1e3:.1f} us  all4 {ta*1e3:.1f} us  loss_grad {tg*1e3:.1f} us")
