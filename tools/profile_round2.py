"""One call of each round-2 hot kernel at bench shapes, for `ncu` captures (no timing here): the C4 eval chain of one
2848x4256 pair (both SSIM policies, dE00 metric, Sobel, selects, PSNR table), the C3 hybrid loss forward + backward and
one fused C2 phys launch."""
import torch

from lowlight_image_enhancement_b200 import _functional as Fn
from lowlight_image_enhancement_b200 import _native as nat
from lowlight_image_enhancement_b200.eval import C4_METRICS, image_metrics
from lowlight_image_enhancement_b200.NewBP_model.losses import HybridLossPlus
from tests.golden_inputs import depthwise_from_spec

dev = torch.device("cuda:0")
H, W = 2848, 4256
g = torch.Generator(device=dev).manual_seed(4000)
gt = torch.round(255 * torch.rand(1, 3, H, W, device=dev, generator=g)) / 255
gt = torch.round(255 * torch.nn.functional.avg_pool2d(gt, 5, 1, 2)) / 255
pred = (gt + 0.03 * torch.randn(gt.shape, device=dev, generator=g)).clamp(0, 1)
v = image_metrics(pred, gt, list(C4_METRICS))
torch.cuda.synchronize()
print(dict(zip(C4_METRICS, v.tolist())))
n = 16
B = torch.rand(n, 3, 512, 512, device=dev)
x = (B + 0.05 * torch.randn_like(B)).clamp(0, 1).requires_grad_(True)
kern = torch.tensor([[.01, .02, .01], [.02, .88, .02], [.01, .02, .01]], device=dev).view(1, 1, 3, 3)
hyb = HybridLossPlus(device="cuda", w_phys=0.05, physics_kernel=kern, perceptual_module=False)
tot, logs = hyb(Bhat_raw=x, B_raw=B, A_raw=B * 0.9, expo_ratio=torch.ones(n, 1, 1, 1, device=dev), Bhat_srgb01=x,
                B_srgb01=B)
tot.backward()
torch.cuda.synchronize()
print({k: float(v) for k, v in logs.items()})
# C2: fused value + gradient phys launch (serialised geometry)
p2 = torch.rand(8, 3, 512, 512, device=dev)
o2 = (p2 * 0.9).contiguous()
psf = Fn.PreparedPSF(depthwise_from_spec("B2").to(dev).contiguous(), Fn.PSF_DEPTHWISE, 3, 3, False)
rho = torch.full((8,), 1.1, device=dev)
cfg = Fn.PhysCfg(nat.PAD["reflect"], True, True, 1e-3, 0, True, False)
vec, fl, gr = Fn.phys_value_and_grad(p2, o2, psf, rho, 0.0, (1, 0, 0, 0), cfg, torch.ones(8, device=dev), 1.0 / 8)
torch.cuda.synchronize()
print(vec.tolist())
