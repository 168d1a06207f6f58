"""Where an end-to-end edge-mask count differs from the CPU oracle's: is it the L plane (last-bit differences between
this library's sRGB -> Lab and the host libm's) or the Sobel / quantile / count stage?  Random 8-bit batches; per trial
the GPU count, the oracle count on the oracle's own L plane, and the oracle STAGE count on the GPU's L plane."""
import sys
import torch
from lowlight_image_enhancement_b200 import _functional as Fn
from lowlight_image_enhancement_b200.metrics.color_error import edge_deltaE2000_stats
from oracle import colour

dev = torch.device("cuda:0")
trials = int(sys.argv[1]) if len(sys.argv) > 1 else 40
n, h, w = 9, 130, 256
stage_bad = e2e_bad = l_diff_trials = 0
for it in range(trials):
    g = torch.Generator().manual_seed(9000 + it)
    gt8 = torch.round(255 * torch.rand(n, 3, h, w, generator=g)) / 255
    gt8 = torch.round(255 * torch.nn.functional.avg_pool2d(gt8, 3, 1, 1)) / 255
    pr8 = torch.round(255 * (gt8 + 0.05 * torch.randn(n, 3, h, w, generator=g)).clamp(0, 1)) / 255
    es = edge_deltaE2000_stats(pr8.to(dev), gt8.to(dev), q=0.85)
    er = colour.edge_stats(pr8, gt8, q=0.85)
    _, _, _, lpl = Fn.deltae(pr8.to(dev), gt8.to(dev), variant=0, clamp_in=True, want_l=True)
    l_gpu = lpl.cpu().view(n, 1, h, w)
    l_cpu = colour.rgb_to_lab(pr8)[:, 0:1]
    ndiff = int((l_gpu.view(torch.int32) != l_cpu.contiguous().view(torch.int32)).sum())
    _, _, cnt_stage = colour.edge_from_l(l_gpu, q=0.85)
    stage = int(cnt_stage.sum())
    l_diff_trials += ndiff > 0
    if es["count"] != stage:
        stage_bad += 1
    if es["count"] != er["count"]:
        e2e_bad += 1
        print(f"trial {it}: GPU {es['count']}  oracle end-to-end {er['count']}  oracle stage on the GPU's L {stage}  "
              f"L values differing in the last bits: {ndiff} of {l_gpu.numel()}", flush=True)
print(f"{trials} trials of {n}x{h}x{w}: end-to-end count differs in {e2e_bad}, stage count (same L plane) differs in "
      f"{stage_bad}; trials with any last-bit difference in L: {l_diff_trials}")
