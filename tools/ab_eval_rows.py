"""A/B of LLIE_TUNE_SSIM_SEG_ROWS inside the concurrent four-metric evaluation of 2848x4256 pairs (diagnostic)."""
import torch, bench
from lowlight_image_enhancement_b200 import _native as nat
from lowlight_image_enhancement_b200.eval import ShardedMeans, accumulate_image_metrics
dev = torch.device("cuda:0")
nat.ensure_device(dev)
pool = [bench.synth_eval_pair(4000 + i, dev) for i in range(3)]
names = ["linear_psnr", "linear_ssim", "deltae2000_mean", "edge_deltae2000_mean"]
acc = ShardedMeans(names, dev)
for rows in (0, 168, 84, 112, 0, 168, 84, 112, 126):
    nat.lib().llie_tune(4, rows)
    t = bench.timed_warm(lambda i: accumulate_image_metrics(pool[i % 3][0], pool[i % 3][1], names, acc), 48, dev) / 48
    ts = bench.timed_warm(lambda i: accumulate_image_metrics(pool[i % 3][0], pool[i % 3][1], names, acc,
                                                             concurrent=False), 48, dev) / 48
    print(f"seg_rows {rows:4d}: concurrent {t*1e3:.1f} us  one stream {ts*1e3:.1f} us")
nat.lib().llie_tune(4, 0)
