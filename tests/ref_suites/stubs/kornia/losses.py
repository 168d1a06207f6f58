"""kornia.losses.SSIMLoss restated in oracle/quality.py (unpinned)."""
import torch

from oracle import quality


class SSIMLoss(torch.nn.Module):
    def __init__(self, window_size: int, max_val: float = 1.0, eps: float = 1e-12, reduction: str = "mean",
                 padding: str = "same"):
        super().__init__()
        self.window_size, self.max_val, self.eps, self.reduction = window_size, max_val, eps, reduction

    def forward(self, img1, img2):
        loss = torch.clamp((1.0 - quality.kornia_ssim_map(img1, img2, self.window_size, self.max_val, self.eps)) / 2,
                           0, 1)
        if self.reduction == "mean":
            return loss.mean()
        if self.reduction == "sum":
            return loss.sum()
        return loss
