"""Oracle (TEST INFRASTRUCTURE ONLY): sRGB -> Lab(D65/2deg) and the two CIEDE2000 variants, the
Sobel-on-L edge mask with torch.quantile thresholds, on CPU torch.

Follows (paths relative to /root/reference):
  * metrics/color_error.py:105-210  _deltaE00_lab_map   (METRIC variant: eps=1e-12 in sqrt/denominators,
                                     hue NOT wrapped, valid-mask piecewise dh / hbar, sin(2*dtheta), clamp>=0)
  * metrics/color_error.py:213-231  _compute_percentiles (torch.quantile, linear interpolation)
  * metrics/color_error.py:235-293  deltaE2000_map / deltaE2000_summary
  * metrics/color_error.py:296-344  _sobel_magnitude / edge_deltaE2000
  * NewBP_model/losses.py:99-143    DeltaE00Loss._ciede2000 (LOSS variant: eps=1e-6 everywhere,
                                     hue mod 2pi, boolean-mask wraps, sin(dtheta), +eps under final sqrt)

Third-party arithmetic (kornia==0.6.12 color/lab.py::rgb_to_lab / lab_to_rgb, color/rgb.py::rgb_to_linear_rgb /
linear_rgb_to_rgb, color/xyz.py::rgb_to_xyz / xyz_to_rgb — not vendored, not installed, probed absent on the GPU
image too): restated in `rgb_to_lab` / `lab_to_rgb`; parity unpinned.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

PI = math.pi


def rgb_to_lab(img: torch.Tensor) -> torch.Tensor:
    """kornia 0.6.12 rgb_to_lab: sRGB EOTF -> XYZ (D65) -> normalise by white -> f() -> Lab."""
    lin = torch.where(img > 0.04045, torch.pow((img + 0.055) / 1.055, 2.4), img / 12.92)
    r, g, b = lin[..., 0, :, :], lin[..., 1, :, :], lin[..., 2, :, :]
    x = 0.412453 * r + 0.357580 * g + 0.180423 * b
    y = 0.212671 * r + 0.715160 * g + 0.072169 * b
    z = 0.019334 * r + 0.119193 * g + 0.950227 * b
    xyz = torch.stack([x, y, z], dim=-3)
    white = torch.tensor([0.95047, 1.0, 1.08883], dtype=xyz.dtype, device=xyz.device)[..., :, None, None]
    t = torch.div(xyz, white)
    thr = 0.008856
    f = torch.where(t > thr, torch.pow(t.clamp(min=thr), 1 / 3.0), 7.787 * t + 4.0 / 29.0)
    fx, fy, fz = f[..., 0, :, :], f[..., 1, :, :], f[..., 2, :, :]
    return torch.stack([116.0 * fy - 16.0, 500.0 * (fx - fy), 200.0 * (fy - fz)], dim=-3)


def lab_to_rgb(lab: torch.Tensor, clip: bool = True) -> torch.Tensor:
    """kornia 0.6.12 lab_to_rgb (only the reference's TESTS call it, standard_tests/test_color_error.py:191-192):
    Lab -> f -> XYZ (times the D65 white) -> linear RGB (inverse matrix) -> sRGB OETF, clipped to [0,1]."""
    L, a, b = lab[..., 0, :, :], lab[..., 1, :, :], lab[..., 2, :, :]
    fy = (L + 16.0) / 116.0
    fx = (a / 500.0) + fy
    fz = fy - (b / 200.0)
    if clip:
        fz = fz.clamp(min=0.0)
    f = torch.stack([fx, fy, fz], dim=-3)
    xyz = torch.where(f > 0.2068966, torch.pow(f, 3.0), (f - 4.0 / 29.0) / 7.787)
    white = torch.tensor([0.95047, 1.0, 1.08883], dtype=xyz.dtype, device=xyz.device)[..., :, None, None]
    xyz = xyz * white
    x, y, z = xyz[..., 0, :, :], xyz[..., 1, :, :], xyz[..., 2, :, :]
    r = 3.2404813432005266 * x + -1.5371515162713185 * y + -0.4985363261688878 * z
    g = -0.9692549499965682 * x + 1.8759900014898907 * y + 0.0415559265582928 * z
    bl = 0.0556466391351772 * x + -0.2040413383665112 * y + 1.0573110696453443 * z
    lin = torch.stack([r, g, bl], dim=-3)
    thr = 0.0031308
    rgb = torch.where(lin > thr, 1.055 * torch.pow(lin.clamp(min=thr), 1 / 2.4) - 0.055, 12.92 * lin)
    if clip:
        rgb = rgb.clamp(0.0, 1.0)
    return rgb


def _rad(deg: float) -> float:
    return float(deg) * PI / 180.0


def de00_metric_lab(lab1, lab2, *, kL=1.0, kC=1.0, kH=1.0, eps=1e-12):
    """METRIC variant on Lab tensors [N,3,H,W] -> [N,H,W]."""
    L1, a1, b1 = lab1[:, 0], lab1[:, 1], lab1[:, 2]
    L2, a2, b2 = lab2[:, 0], lab2[:, 1], lab2[:, 2]
    c1 = torch.sqrt(a1 * a1 + b1 * b1 + eps)
    c2 = torch.sqrt(a2 * a2 + b2 * b2 + eps)
    cb = 0.5 * (c1 + c2)
    cb7 = cb.pow(7)
    k25 = torch.tensor(25.0 ** 7, dtype=lab1.dtype)
    g = 0.5 * (1.0 - torch.sqrt(cb7 / (cb7 + k25 + eps)))
    a1p, a2p = (1.0 + g) * a1, (1.0 + g) * a2
    c1p = torch.sqrt(a1p * a1p + b1 * b1 + eps)
    c2p = torch.sqrt(a2p * a2p + b2 * b2 + eps)
    h1, h2 = torch.atan2(b1, a1p), torch.atan2(b2, a2p)
    dL, dC = L2 - L1, c2p - c1p
    ok = (c1p * c2p) != 0.0
    df = h2 - h1
    dh = torch.zeros_like(h1)
    dh = torch.where(ok & (df.abs() <= PI), df, dh)
    dh = torch.where(ok & (df > PI), df - 2.0 * PI, dh)
    dh = torch.where(ok & (df < -PI), df + 2.0 * PI, dh)
    dH = 2.0 * torch.sqrt(c1p * c2p + eps) * torch.sin(dh / 2.0)
    Lb, Cb = 0.5 * (L1 + L2), 0.5 * (c1p + c2p)
    hs, ad = h1 + h2, (h1 - h2).abs()
    hb = torch.where(~ok, hs,
                     torch.where(ad <= PI, 0.5 * hs,
                                 torch.where(hs < 2.0 * PI, 0.5 * (hs + 2.0 * PI), 0.5 * (hs - 2.0 * PI))))
    t = (1.0 - 0.17 * torch.cos(hb - _rad(30.0)) + 0.24 * torch.cos(2.0 * hb)
         + 0.32 * torch.cos(3.0 * hb + _rad(6.0)) - 0.20 * torch.cos(4.0 * hb - _rad(63.0)))
    dth = _rad(30.0) * torch.exp(-(((hb * 180.0 / PI) - 275.0) / 25.0) ** 2)
    rc = 2.0 * torch.sqrt(Cb.pow(7) / (Cb.pow(7) + k25 + eps))
    rt = -torch.sin(2.0 * dth) * rc
    sL = 1.0 + (0.015 * (Lb - 50.0) ** 2) / torch.sqrt(20.0 + (Lb - 50.0) ** 2 + eps)
    sC = 1.0 + 0.045 * Cb
    sH = 1.0 + 0.015 * Cb * t
    lt, ct, ht = dL / (kL * sL + eps), dC / (kC * sC + eps), dH / (kH * sH + eps)
    return torch.sqrt(torch.clamp(lt * lt + ct * ct + ht * ht + rt * ct * ht, min=0.0))


def de00_loss_lab(lab1, lab2, eps=1e-6):
    """LOSS variant on Lab tensors [N,3,H,W] -> [N,H,W] (differentiable)."""
    L1, a1, b1 = lab1[:, 0], lab1[:, 1], lab1[:, 2]
    L2, a2, b2 = lab2[:, 0], lab2[:, 1], lab2[:, 2]
    c1 = torch.sqrt(a1 * a1 + b1 * b1 + eps)
    c2 = torch.sqrt(a2 * a2 + b2 * b2 + eps)
    cb = 0.5 * (c1 + c2)
    g = 0.5 * (1 - torch.sqrt((cb ** 7) / ((cb ** 7) + (25 ** 7) + eps)))
    a1p, a2p = (1 + g) * a1, (1 + g) * a2
    c1p = torch.sqrt(a1p * a1p + b1 * b1 + eps)
    c2p = torch.sqrt(a2p * a2p + b2 * b2 + eps)
    h1 = torch.atan2(b1, a1p) % (2 * PI)
    h2 = torch.atan2(b2, a2p) % (2 * PI)
    dL, dC = L2 - L1, c2p - c1p
    dh = h2 - h1
    dh = dh - (2 * PI) * (dh > PI) + (2 * PI) * (dh < -PI)
    dH = 2 * torch.sqrt(c1p * c2p + eps) * torch.sin(dh / 2)
    Lb, Cb = 0.5 * (L1 + L2), 0.5 * (c1p + c2p)
    hs = h1 + h2
    hb = hs / 2 - PI * ((h1 - h2).abs() > PI) + (2 * PI) * (hs < 0)
    d30 = torch.deg2rad(torch.tensor(30.0))
    d6 = torch.deg2rad(torch.tensor(6.0))
    d63 = torch.deg2rad(torch.tensor(63.0))
    t = (1 - 0.17 * torch.cos(hb - d30) + 0.24 * torch.cos(2 * hb)
         + 0.32 * torch.cos(3 * hb + d6) - 0.20 * torch.cos(4 * hb - d63))
    dro = 30 * torch.exp(-((torch.rad2deg(hb) - 275) / 25) ** 2)
    rc = 2 * torch.sqrt((Cb ** 7) / ((Cb ** 7) + (25 ** 7) + eps))
    sL = 1 + (0.015 * ((Lb - 50) ** 2)) / torch.sqrt(20 + (Lb - 50) ** 2 + eps)
    sC = 1 + 0.045 * Cb
    sH = 1 + 0.015 * Cb * t
    rt = -torch.sin(torch.deg2rad(dro)) * rc
    return torch.sqrt((dL / sL) ** 2 + (dC / sC) ** 2 + (dH / sH) ** 2 + rt * (dC / sC) * (dH / sH) + eps)


def de00_loss(gen, tgt, eps=1e-6):
    """losses.py:138-143."""
    return de00_loss_lab(rgb_to_lab(gen.clamp(0, 1)), rgb_to_lab(tgt.clamp(0, 1)), eps).mean()


def de00_map(pred, target, *, kL=1.0, kC=1.0, kH=1.0, eps=1e-12):
    """color_error.py:235-266 -> [N,H,W] (or [H,W] for 3-D inputs)."""
    squeeze = pred.ndim == 3
    if squeeze:
        pred, target = pred[None], target[None]
    out = de00_metric_lab(rgb_to_lab(pred.detach().to(torch.float32)),
                          rgb_to_lab(target.detach().to(torch.float32)), kL=kL, kC=kC, kH=kH, eps=eps)
    return out.squeeze(0) if squeeze else out


def de00_summary(pred, target, percentiles=(50.0, 95.0), **kw):
    """color_error.py:270-293: mean of per-image means; percentiles over ALL pixels jointly."""
    m = de00_map(pred, target, **kw)
    flat = m.reshape(m.shape[0], -1)
    out = {"mean": float(flat.mean(dim=1).mean())}
    qs = torch.tensor([float(p) / 100.0 for p in percentiles], dtype=flat.dtype, device=flat.device)
    qv = torch.quantile(flat.reshape(-1), qs)
    for i, p in enumerate(percentiles):
        out[f"p{int(p)}"] = float(qv[i])
    return out


SOBEL_X = torch.tensor([[-1.0, 0.0, 1.0], [-2.0, 0.0, 2.0], [-1.0, 0.0, 1.0]]).view(1, 1, 3, 3)
SOBEL_Y = torch.tensor([[-1.0, -2.0, -1.0], [0.0, 0.0, 0.0], [1.0, 2.0, 1.0]]).view(1, 1, 3, 3)


def sobel_magnitude(l_plane):
    """color_error.py:296-301 on [N,1,H,W]."""
    gx = F.conv2d(l_plane, SOBEL_X.to(l_plane), padding=1)
    gy = F.conv2d(l_plane, SOBEL_Y.to(l_plane), padding=1)
    return torch.sqrt(gx * gx + gy * gy + 1e-12)


def edge_stats(pred, target, q=0.85, **kw):
    """color_error.py:305-344 -> dict(mean, p95, count[int], thresholds[N])."""
    de = de00_map(pred, target, **kw)
    if de.ndim == 3:
        de = de.unsqueeze(1)
    g = sobel_magnitude(rgb_to_lab(pred.detach())[:, 0:1])
    thr = torch.quantile(g.view(g.shape[0], -1), q, dim=1, keepdim=True)
    mask = g >= thr.view(-1, 1, 1, 1)
    vals = de[mask.expand_as(de)]
    if vals.numel() == 0:
        return {"mean": float("nan"), "p95": float("nan"), "count": 0, "thresholds": thr.view(-1)}
    return {"mean": float(vals.mean()), "p95": float(torch.quantile(vals, 0.95)),
            "count": int(mask.sum()), "count_per_image": mask.flatten(1).sum(dim=1),
            "thresholds": thr.view(-1)}


def edge_from_l(l_plane, q=0.85):
    """Stage-level oracle: Sobel magnitude of a given L plane [N,1,H,W], per-image quantile
    threshold, mask count.  Used for the bit-exact count tests (no Lab conversion involved)."""
    g = sobel_magnitude(l_plane)
    thr = torch.quantile(g.view(g.shape[0], -1), q, dim=1, keepdim=True)
    mask = g >= thr.view(-1, 1, 1, 1)
    return g, thr.view(-1), mask.flatten(1).sum(dim=1)
