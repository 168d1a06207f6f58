"""CPU-only checks of the boundary: the library builds/loads and exports every symbol the header
declares; the product never imports the oracle; missing CUDA raises instead of falling back."""
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "llie_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(llie_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from lowlight_image_enhancement_b200 import _native
    from lowlight_image_enhancement_b200.build import build_library
    build_library()
    lib = _native.lib()
    syms = header_symbols()
    assert len(syms) >= 8
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/llie_b200.h but not exported"
    assert sorted(_native.declared_symbols()) == syms, "ctypes signatures out of sync with the header"
    assert lib.llie_abi_version() == 3
    assert lib.llie_workspace_bytes() >= 1 << 20


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "lowlight_image_enhancement_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports the oracle"


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cuda_raises_loudly():
    from lowlight_image_enhancement_b200._native import NativeLibraryError
    from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw
    x = torch.rand(1, 3, 8, 8)
    with pytest.raises(NativeLibraryError):
        phys_cons_raw(x, x, torch.ones(3, 3, 3, 3), 1.0)


def test_validation_errors_precede_device_use():
    """Shape / type contracts of the reference raise before any CUDA work (so they hold on CPU too)."""
    from lowlight_image_enhancement_b200.metrics.phys_consistency import phys_cons_raw
    x = torch.rand(1, 3, 8, 8)
    with pytest.raises(TypeError):
        phys_cons_raw([1], x, torch.ones(3, 3, 3, 3), 1.0)
    with pytest.raises(ValueError):
        phys_cons_raw(x, torch.rand(1, 3, 8, 9), torch.ones(3, 3, 3, 3), 1.0)
    with pytest.raises(ValueError):
        phys_cons_raw(x, x, torch.ones(3, 3, 3, 3), 1.0, eps=0.0)
    with pytest.raises(ValueError):
        phys_cons_raw(x, x, torch.ones(3, 3, 3, 3), 1.0, crop="bogus")
