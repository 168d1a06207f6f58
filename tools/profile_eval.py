"""One call of each eval / loss kernel at bench shapes, for `ncu` captures (no timing here)."""
import sys

import torch

from lowlight_image_enhancement_b200 import _functional as Fn
from lowlight_image_enhancement_b200.eval import image_metrics
from lowlight_image_enhancement_b200.NewBP_model.losses import HybridLossPlus

dev = torch.device("cuda:0")
H, W = (2848, 4256) if len(sys.argv) < 3 else (int(sys.argv[1]), int(sys.argv[2]))
g = torch.Generator(device=dev).manual_seed(4000)
gt = torch.round(255 * torch.rand(1, 3, H, W, device=dev, generator=g)) / 255
gt = torch.round(255 * torch.nn.functional.avg_pool2d(gt, 5, 1, 2)) / 255
pred = (gt + 0.03 * torch.randn(gt.shape, device=dev, generator=g)).clamp(0, 1)
names = ["linear_psnr", "linear_ssim", "deltae2000_mean", "deltae2000_p95", "edge_deltae2000_mean"]
for _ in range(2):
    v = image_metrics(pred, gt, names)
torch.cuda.synchronize()
print(dict(zip(names, v.tolist())))
n = 16
B = torch.rand(n, 3, 512, 512, device=dev)
x = (B + 0.05 * torch.randn_like(B)).clamp(0, 1).requires_grad_(True)
kern = torch.tensor([[.01, .02, .01], [.02, .88, .02], [.01, .02, .01]], device=dev).view(1, 1, 3, 3)
hyb = HybridLossPlus(device="cuda", w_phys=0.05, physics_kernel=kern, perceptual_module=False)
for _ in range(2):
    tot, logs = hyb(Bhat_raw=x, B_raw=B, A_raw=B * 0.9, expo_ratio=torch.ones(n, 1, 1, 1, device=dev), Bhat_srgb01=x,
                    B_srgb01=B)
    tot.backward()
torch.cuda.synchronize()
print({k: float(v) for k, v in logs.items()})
