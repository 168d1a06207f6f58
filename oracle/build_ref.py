"""Recipe for ``oracle/_ref`` — the reference's OWN importable modules of the hot path (TEST INFRASTRUCTURE).

The reference is pure Python.  Six of its modules run as they are on CPU torch in this image:

    metrics/phys_consistency.py   metrics/linear.py   metrics/channelwise.py   metrics/psnr.py
    NewBP_model/newbp_layer.py    NewBP_model/losses.py (classes that need neither kornia nor VGG weights)

``build()`` copies them, unmodified, from /root/reference into ``oracle/_ref/`` — git-ignored (no reference source in
the history) but not gpurun-ignored, so the copies travel to the GPU box like a built ``.so``.  There they are the
``cpu_baseline`` / ``--impl reference`` arm of ``bench.py`` (``kind: "reference"``) and an extra checker of the
restatements in ``oracle/`` — never a product path.  Run by ``__graft_entry__.build()`` when /root/reference exists.
"""
from __future__ import annotations

import importlib.util
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
REFERENCE_ROOT = os.environ.get("LLIE_REFERENCE_ROOT", "/root/reference")
MODULES = ["metrics/phys_consistency.py", "metrics/linear.py", "metrics/channelwise.py", "metrics/psnr.py",
           "NewBP_model/newbp_layer.py", "NewBP_model/losses.py"]


def build() -> list:
    if not os.path.isdir(REFERENCE_ROOT):
        raise FileNotFoundError(f"{REFERENCE_ROOT} is not present (build container only)")
    for rel in MODULES:
        dst = os.path.join(DEST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(REFERENCE_ROOT, rel), dst)
    return list(MODULES)


def available() -> bool:
    return all(os.path.exists(os.path.join(DEST, rel)) for rel in MODULES)


def load(rel: str):
    """Import one vendored reference module by its path (e.g. 'metrics/phys_consistency.py') under a private name, so
    that it can never shadow — or be shadowed by — the mirror package's modules of the same name."""
    name = "llie_oracle_ref." + rel[:-3].replace("/", ".")
    if name in sys.modules:
        return sys.modules[name]
    path = os.path.join(DEST, rel)
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} missing: run oracle/build_ref.py in the build container")
    if rel.startswith("NewBP_model/") and "llie_oracle_ref.NewBP_model" not in sys.modules:
        import types
        pkg = types.ModuleType("llie_oracle_ref.NewBP_model")
        pkg.__path__ = [os.path.join(DEST, "NewBP_model")]
        sys.modules.setdefault("llie_oracle_ref", types.ModuleType("llie_oracle_ref"))
        sys.modules["llie_oracle_ref.NewBP_model"] = pkg
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    print(build())
