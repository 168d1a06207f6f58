"""Run one eval function a few times for ncu: python -m tools.profile_one ssim|de|edge [H W]."""
import sys

import torch

from lowlight_image_enhancement_b200.eval import image_metrics

what = sys.argv[1]
H, W = (2848, 4256) if len(sys.argv) < 4 else (int(sys.argv[2]), int(sys.argv[3]))
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(4000)
gt = torch.round(255 * torch.rand(1, 3, H, W, device=dev, generator=g)) / 255
gt = torch.round(255 * torch.nn.functional.avg_pool2d(gt, 5, 1, 2)) / 255
pred = (gt + 0.03 * torch.randn(gt.shape, device=dev, generator=g)).clamp(0, 1)
names = {"ssim": ["linear_ssim"], "de": ["deltae2000_mean"], "edge": ["edge_deltae2000_mean"],
         "psnr": ["linear_psnr"]}[what]
for _ in range(3):
    v = image_metrics(pred, gt, names)
torch.cuda.synchronize()
print(v.tolist())
