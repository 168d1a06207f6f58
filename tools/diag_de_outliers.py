"""Largest per-pixel deviations of the dE00 metric map from the float64 oracle on random 8-bit images, next to the
deviation of the float32 evaluation of the same oracle formula at the same pixel: is an outlier the kernel's or the
formula's conditioning (hue of an almost neutral colour)?"""
import sys
import torch
from lowlight_image_enhancement_b200.metrics.color_error import deltaE2000_map
from oracle import colour

dev = torch.device("cuda:0")
trials = int(sys.argv[1]) if len(sys.argv) > 1 else 60
worst = []
for it in range(trials):
    g = torch.Generator().manual_seed(7000 + it)
    n, h, w = 1, 134, 248
    gt8 = torch.round(255 * torch.rand(n, 3, h, w, generator=g)) / 255
    gt8 = torch.round(255 * torch.nn.functional.avg_pool2d(gt8, 3, 1, 1)) / 255
    pr8 = torch.round(255 * (gt8 + 0.05 * torch.randn(n, 3, h, w, generator=g)).clamp(0, 1)) / 255
    m = deltaE2000_map(pr8.to(dev), gt8.to(dev)).cpu().double().view(n, h, w)
    lab1, lab2 = colour.rgb_to_lab(pr8.double()), colour.rgb_to_lab(gt8.double())
    r64 = colour.de00_metric_lab(lab1, lab2).view(n, h, w)
    r32 = colour.de00_metric_lab(colour.rgb_to_lab(pr8), colour.rgb_to_lab(gt8)).double().view(n, h, w)
    d = (m - r64).abs()
    k = int(d.argmax())
    y, x = divmod(k, w)
    c1 = float((lab1[0, 1, y, x] ** 2 + lab1[0, 2, y, x] ** 2).sqrt())
    c2 = float((lab2[0, 1, y, x] ** 2 + lab2[0, 2, y, x] ** 2).sqrt())
    worst.append((float(d.max()), float((r32 - r64).abs()[0, y, x]), float((r32 - r64).abs().max()), c1, c2,
                  (pr8[0, :, y, x] * 255).int().tolist(), (gt8[0, :, y, x] * 255).int().tolist(), float(r64[0, y, x])))
worst.sort(reverse=True)
print("kernel max |err|   fp32 oracle |err| at that pixel   fp32 oracle max |err| in the image   chroma pred / gt   rgb pred / gt   dE00")
for wv in worst[:8]:
    print("%.4f   %.4f   %.4f   %.3f / %.3f   %s / %s   %.3f" % wv)
