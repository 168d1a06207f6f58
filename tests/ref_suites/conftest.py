"""Run the reference's OWN pytest suites against the B200 mirror (VERDICT r1 "missing" item 3).

``vendor.py`` copies the reference's test files, unmodified, into ``tests/ref_suites/_ref/`` (git-ignored; they travel to
the GPU box with the built library).  This conftest performs the ``sys.modules`` swap INTEGRATION.md §1 describes, so
that the files' ``from metrics.… import`` / ``from NewBP_model.… import`` lines resolve to
``lowlight_image_enhancement_b200``:

    metrics, metrics.{phys_consistency, linear, channelwise, psnr, ssim, color_error, evaluation}
    NewBP_model, NewBP_model.{newbp_layer, losses, newbp_net_arch}

and installs stand-ins for the two absent third-party packages (``stubs/kornia``, ``stubs/torchmetrics``: oracle
restatements, UNPINNED).  Every collected item gets the ``gpu`` marker (the mirror has no CPU path; the files' ``cpu``
device legs hand HOST tensors to the mirror, which uploads them, computes on the B200 and returns host results).

Markers added here:
  * ``unpinned``  — the assertion compares the mirror with a stand-in of kornia / torchmetrics (restatement, not
                    the real package): it passes or fails on the restatement's terms only.
  * skip          — needs pretrained VGG19 / LPIPS weights (a download; there is no network on either box).

``LLIE_REF_SUITES_TARGET=reference`` (build container only) runs the same files against /root/reference itself on CPU —
the sanity check that the stand-ins and this harness do not bend the tests.
"""
from __future__ import annotations

import importlib
import importlib.util
import os
import sys
import types

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_DIR = os.path.join(HERE, "_ref")
STUBS = os.path.join(HERE, "stubs")
TARGET = os.environ.get("LLIE_REF_SUITES_TARGET", "mirror")

NEEDS_WEIGHTS = {
    # NewBP_model/losses.py:41 builds torchvision VGG19 with IMAGENET1K_V1 weights (download)
    "test_perceptual_loss_accepts_srgb_in_unit_interval", "test_perceptual_loss_imagenet_normalization_math",
    "test_perceptual_loss_rejects_non_three_channel_inputs",
    "test_hybrid_loss",        # asserts perc.item() > 0: needs the VGG term
    "test_lpips",              # metrics/lpips_metric.py: LPIPS weights
}
UNPINNED_FILES = {"test_color_error.py"}
UNPINNED_TESTS = {"test_matches_torchmetrics_reference", "test_rgb_to_lab_and_delta_e_sanity"}


def _alias(name: str, module) -> None:
    sys.modules[name] = module


def _install_mirror() -> None:
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    if STUBS not in sys.path:
        sys.path.insert(0, STUBS)
    pkg = "lowlight_image_enhancement_b200"
    for sub, mods in (("metrics", ("phys_consistency", "linear", "channelwise", "psnr", "ssim", "color_error",
                                   "evaluation")),
                      ("NewBP_model", ("newbp_layer", "losses", "newbp_net_arch"))):
        _alias(sub, importlib.import_module(f"{pkg}.{sub}"))
        for m in mods:
            _alias(f"{sub}.{m}", importlib.import_module(f"{pkg}.{sub}.{m}"))
    # metrics/lpips_metric.py is LPIPS (out of scope, needs downloaded weights): a placeholder so that
    # core_tests/test_model_and_metrics.py imports; its test_lpips cases are skipped (NEEDS_WEIGHTS)
    lp = types.ModuleType("metrics.lpips_metric")

    class LPIPSEvaluator:  # noqa: D401
        def __init__(self, *a, **k):
            raise RuntimeError("LPIPS is outside the hot path and needs pretrained weights (no network)")

    lp.LPIPSEvaluator = LPIPSEvaluator
    _alias("metrics.lpips_metric", lp)


def _install_reference() -> None:
    ref = os.environ.get("LLIE_REFERENCE_ROOT", "/root/reference")
    for p in (STUBS, ROOT, os.path.join(ref, "NAFNet_base"), ref):
        if p not in sys.path:
            sys.path.insert(0, p)
    # pin the reference's own top-level packages NOW: some of its test files put NAFNet_base/basicsr on sys.path, after
    # which a fresh `import metrics` would find basicsr/metrics instead of <reference>/metrics
    for name in ("metrics", "metrics.phys_consistency", "metrics.linear", "metrics.channelwise", "metrics.psnr",
                 "metrics.ssim", "metrics.color_error", "NewBP_model", "NewBP_model.newbp_layer", "NewBP_model.losses"):
        try:
            importlib.import_module(name)
        except Exception:  # a module the container cannot import stays unresolved; its tests then fail on their own
            pass


if os.path.isdir(REF_DIR):
    if TARGET == "reference":
        _install_reference()
    else:
        _install_mirror()


def pytest_configure(config):
    config.addinivalue_line("markers", "unpinned: compares with an oracle stand-in of kornia / torchmetrics")


def _reference_cannot_import(path: str) -> bool:
    """Sanity mode only: files whose imports pull the BasicSR package in (`basicsr.utils` imports lmdb at import time)
    cannot be collected against the reference in a container without lmdb; a failed import there also leaves
    half-initialised `metrics` / `NewBP_model` packages behind and breaks the files collected after it."""
    if TARGET != "reference" or importlib.util.find_spec("lmdb") is not None:
        return False
    return os.path.basename(path) in {"test_model_and_metrics.py", "test_scenarioB_no_double_k.py"}


def pytest_ignore_collect(collection_path, config):
    # nothing vendored (fresh clone outside the build container): no reference suites to run
    if not os.path.isdir(REF_DIR) and str(collection_path).startswith(REF_DIR):
        return True
    if _reference_cannot_import(str(collection_path)):
        return True
    return None


def pytest_collection_modifyitems(config, items):
    for item in items:
        path = str(item.fspath)
        if not path.startswith(REF_DIR):
            continue
        if TARGET != "reference":
            item.add_marker(pytest.mark.gpu)
        name = item.originalname if hasattr(item, "originalname") and item.originalname else item.name
        if name in NEEDS_WEIGHTS:
            item.add_marker(pytest.mark.skip(reason="needs pretrained VGG19 / LPIPS weights (download; no network)"))
        if os.path.basename(path) in UNPINNED_FILES or name in UNPINNED_TESTS:
            item.add_marker(pytest.mark.unpinned)
