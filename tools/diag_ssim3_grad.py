"""SSIMLoss gradient with a 3-tap window: deviation from the float64 oracle of this library and of the float32
evaluation of the same oracle (conditioning of sigma^2 = E[x^2] - mu^2 over nine pixels)."""
import torch
from lowlight_image_enhancement_b200.NewBP_model.losses import SSIMLoss
from oracle import quality

dev = torch.device("cuda:0")


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max())


for ws in (3, 5, 11):
    e_ours, e_f32 = [], []
    for it in range(12):
        g = torch.Generator().manual_seed(300 + it)
        n, c, h, w = 3, 1, 137, 256
        tgt = torch.rand(n, c, h, w, generator=g)
        gen = (tgt + 0.1 * torch.randn(n, c, h, w, generator=g)).clamp(-0.05, 1.05)
        x64 = gen.double().requires_grad_(True)
        quality.ssim_loss_kornia(x64, tgt.double(), ws, 1.0).backward()
        x32 = gen.clone().requires_grad_(True)
        quality.ssim_loss_kornia(x32, tgt, ws, 1.0).backward()
        x = gen.to(dev).requires_grad_(True)
        SSIMLoss(window_size=ws)(x, tgt.to(dev)).backward()
        e_ours.append(rel(x.grad, x64.grad))
        e_f32.append(rel(x32.grad, x64.grad))
    print(f"window {ws:2d}: max over 12 seeds of max|grad - grad64| / max|grad64|:  this library {max(e_ours):.2e}   "
          f"float32 oracle {max(e_f32):.2e}")
