"""torchmetrics.image.StructuralSimilarityIndexMeasure restated in oracle/quality.py (unpinned)."""
import torch

from oracle import quality


class StructuralSimilarityIndexMeasure(torch.nn.Module):
    def __init__(self, gaussian_kernel: bool = True, sigma=1.5, kernel_size=11, reduction="elementwise_mean",
                 data_range=None, k1: float = 0.01, k2: float = 0.03, **_):
        super().__init__()
        if not gaussian_kernel:
            raise NotImplementedError("the restatement covers the Gaussian window only")
        self.cfg = dict(sigma=float(sigma), kernel_size=int(kernel_size), k1=k1, k2=k2)
        # public attributes of the real class that metrics/ssim.py:236-241, 320 reads back
        self.kernel_size, self.sigma, self.k1, self.k2 = kernel_size, sigma, k1, k2
        self.data_range, self.reduction = data_range, reduction
        self.reset()

    def reset(self):
        self._scores = []

    def update(self, preds, target):
        dr = self.data_range
        if dr is None:
            dr = float(max(preds.max() - preds.min(), target.max() - target.min()))
        self._scores.append(quality.ssim_torchmetrics(preds, target, data_range=float(dr), **self.cfg))

    def compute(self):
        s = torch.cat([v.reshape(-1) for v in self._scores])
        if self.reduction == "elementwise_mean":
            return s.mean()
        if self.reduction == "sum":
            return s.sum()
        return s

    def forward(self, preds, target):
        s = quality.ssim_torchmetrics(preds, target, data_range=float(self.data_range or 1.0), **self.cfg)
        self._scores.append(s)
        if self.reduction == "elementwise_mean":
            return s.mean()
        if self.reduction == "sum":
            return s.sum()
        return s
