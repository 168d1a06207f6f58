"""A few fused value + gradient phys launches in the bench configuration (BASELINE configs[1]: 8x3x512x512, rgb B2 PSF,
per-image rho, MAE, valid crop, reflect) for an `ncu --set full --import-source on` capture."""
import torch
from lowlight_image_enhancement_b200 import _functional as Fn
from lowlight_image_enhancement_b200 import _native as nat
from oracle import psf_ops

dev = torch.device("cuda:0")
n, c, h, w = 8, 3, 512, 512
sets = []
for i in range(3):
    p = torch.rand(n, c, h, w, device=dev)
    rho = torch.tensor([100.0, 250.0, 300.0, 100.0, 250.0, 300.0, 100.0, 250.0], device=dev) / 300.0
    o = (p * rho.view(-1, 1, 1, 1) * 0.9).contiguous()
    sets.append((p, o, rho))
grad = torch.empty(n, c, h, w, device=dev)
psf = Fn.prepare_psf(psf_ops.canonical_kernels("rgb").to(dev), 3, 3, dev, normalize=True, nonneg=False, eps=1e-12)
cfg = Fn.PhysCfg(nat.PAD["reflect"], True, False, 1e-12, 0, False, False)
gout = torch.ones(1, device=dev)
for i in range(8):
    p, o, r = sets[i % 3]
    Fn.phys_value_and_grad(p, o, psf, r, 0.0, (1, 0, 0, 0), cfg, gout, 1.0 / (n * c * (h - 2) * (w - 2)), out=grad)
torch.cuda.synchronize()
print("ok")
