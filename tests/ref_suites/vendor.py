"""Bring the reference's own pytest suites next to the mirror (TEST INFRASTRUCTURE, build container only).

The reference's test files are the real pin of the drop-in boundary ("missing" item 3).  They are NOT
committed: this recipe copies them, byte for byte, from where they lie under /root/reference into
``tests/ref_suites/_ref/`` — a directory listed in ``.gitignore`` (so no reference source enters the history) but not
in ``.gpurunignore`` (so the copies travel to the GPU box with the built ``.so`` files, where /root/reference does not
exist).  ``__graft_entry__.build()`` calls :func:`vendor` whenever /root/reference is present.

What is copied (paths relative to the reference root) and why each file is in the list:

  standard_tests/test_phys_consistency.py   R5/R6  phys_cons_raw / phys_cons_srgb
  standard_tests/test_linear_metrics.py     R15/R16 psnr_linear / ssim_linear
  standard_tests/test_channelwise.py        R17    rgb_psnr / cpsnr_rgb / rgb_ssim
  standard_tests/test_color_error.py        R12-R14 deltaE2000_* / edge_deltaE2000 (imports kornia -> stub, "unpinned")
  standard_tests/test_ssim_wrapper.py       R19    SSIMEvaluator & co. (imports torchmetrics -> stub, "unpinned")
  standard_tests/data/ciede2000_pairs.json  the 18 Lab pairs test_color_error.py reads
  core_tests/test_psf_depthwise_and_kernels.py  R1-R3 CrosstalkPSF vs F.conv2d, AMP smoke
  core_tests/test_physics_loss_grad.py      R1/R2/R7 gradcheck (float64), adjoint identity
  core_tests/test_mode_switch_and_state.py  R2     buffers / state_dict / device moves (NAFNet = stock torch stand-in)
  core_tests/test_losses_domain_norm.py     R9/R10 DeltaE00Loss / SSIMLoss domain tests (VGG tests need weights)
  core_tests/test_integration_forward_amp.py R7    PhysicalConsistencyLossSRGB under fp16 autocast + GradScaler
  core_tests/test_scenarioB_no_double_k.py  R2/R7  PSF only in the loss branch (forward-hook call counts)
  core_tests/test_model_and_metrics.py      R11/R18/R19 HybridLoss(Plus), calculate_psnr / calculate_ssim (VGG / LPIPS need weights)
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
REFERENCE_ROOT = os.environ.get("LLIE_REFERENCE_ROOT", "/root/reference")

FILES = [
    "standard_tests/test_phys_consistency.py",
    "standard_tests/test_linear_metrics.py",
    "standard_tests/test_channelwise.py",
    "standard_tests/test_color_error.py",
    "standard_tests/test_ssim_wrapper.py",
    "standard_tests/data/ciede2000_pairs.json",
    "core_tests/test_psf_depthwise_and_kernels.py",
    "core_tests/test_physics_loss_grad.py",
    "core_tests/test_mode_switch_and_state.py",
    "core_tests/test_losses_domain_norm.py",
    "core_tests/test_integration_forward_amp.py",
    "core_tests/test_scenarioB_no_double_k.py",
    "core_tests/test_model_and_metrics.py",
]


def vendor(verbose: bool = False) -> dict:
    """Copy FILES into tests/ref_suites/_ref/ (unmodified) and write MANIFEST.json with their sha256."""
    if not os.path.isdir(REFERENCE_ROOT):
        raise FileNotFoundError(f"{REFERENCE_ROOT} is not present: the reference suites can only be vendored in the "
                                "build container")
    manifest = {}
    for rel in FILES:
        src = os.path.join(REFERENCE_ROOT, rel)
        dst = os.path.join(DEST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        with open(dst, "rb") as fh:
            manifest[rel] = hashlib.sha256(fh.read()).hexdigest()
        if verbose:
            print(f"[ref_suites] {rel}")
    with open(os.path.join(DEST, "MANIFEST.json"), "w") as fh:
        json.dump({"source": REFERENCE_ROOT, "files": manifest}, fh, indent=1, sort_keys=True)
    return manifest


def verify() -> bool:
    """True when every vendored file still has the recorded hash (nobody edited the copies)."""
    path = os.path.join(DEST, "MANIFEST.json")
    if not os.path.exists(path):
        return False
    with open(path) as fh:
        files = json.load(fh)["files"]
    for rel, digest in files.items():
        p = os.path.join(DEST, rel)
        if not os.path.exists(p):
            return False
        with open(p, "rb") as fh:
            if hashlib.sha256(fh.read()).hexdigest() != digest:
                return False
    return True


if __name__ == "__main__":
    vendor(verbose=True)
