"""Sweep LLIE_TUNE_SSIM_SEG_ROWS for the statistics-only SSIM call on a full-size pair (diagnostic)."""
import torch, bench
from lowlight_image_enhancement_b200 import _native as nat
from lowlight_image_enhancement_b200.metrics.ssim import ssim_sums_torchmetrics
dev = torch.device("cuda:0")
pool = [bench.synth_eval_pair(4000 + i, dev) for i in range(3)]
nat.ensure_device(dev)
for rows in (84, 168, 0, 84, 112, 168, 80, 88, 90, 76, 84, 0):
    nat.lib().llie_tune(4, rows)
    t = bench.timed_warm(lambda i: ssim_sums_torchmetrics(pool[i % 3][0], pool[i % 3][1], data_range=1.0, sigma=1.5,
                                                          k1=0.01, k2=0.03), 60, dev) / 60
    print(f"seg_rows {rows:4d}: {t*1e3:.1f} us")
nat.lib().llie_tune(4, 0)
