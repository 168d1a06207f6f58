"""Golden vectors for the validation tiling, from the UNMODIFIED reference code (build container
only):

    python tests/golden/make_golden_tiles.py

`ImageRestorationModel.grids` / `grids_inverse` of
/root/reference/NAFNet_base/basicsr/models/image_restoration_model.py:167-241 are cut out of the source text by `ast`
and executed unchanged (the module itself needs lmdb / the BasicSR registry at import time) as methods of a bare object
that carries the attributes they read: `gt`, `lq`, `opt['val']`, `scale`, `outs`, `device`.

Writes tests/golden/tile_vectors.npz: per case the tile origins the reference produced, the tile batch checksum, and
the merged image `preds / count_mt` for seeded network outputs.
"""
from __future__ import annotations

import ast
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
SRC = "/root/reference/NAFNet_base/basicsr/models/image_restoration_model.py"

from tests.golden_inputs import TILE_CASES, tile_inputs  # noqa: E402


def reference_methods():
    tree = ast.parse(open(SRC).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "ImageRestorationModel")
    fns = [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name in ("grids", "grids_inverse")]
    assert len(fns) == 2
    mod = ast.Module(body=fns, type_ignores=[])
    ns = {"torch": torch}
    exec(compile(mod, SRC, "exec"), ns)
    return ns["grids"], ns["grids_inverse"]


class Bare:
    pass


def main():
    grids, grids_inverse = reference_methods()
    out = {}
    for k, case in enumerate(TILE_CASES):
        h, w, scale, val = case["h"], case["w"], case["scale"], case["val"]
        lq, outs_fn = tile_inputs(k)
        m = Bare()
        m.opt = {"val": dict(val)}
        m.scale = scale
        m.device = torch.device("cpu")
        m.gt = torch.zeros(1, 3, h, w)
        m.lq = lq
        grids(m)
        idx = np.asarray([[d["i"], d["j"]] for d in m.idxes], dtype=np.int64)
        tiles = m.lq
        m.outs = outs_fn(len(m.idxes), tiles)
        grids_inverse(m)
        out[f"idx{k}"] = idx
        out[f"tiles_shape{k}"] = np.asarray(tiles.shape, dtype=np.int64)
        out[f"tiles_sum{k}"] = np.asarray(tiles.double().sum().item())
        out[f"merged{k}"] = m.output.numpy()
        assert torch.equal(m.lq, lq)  # grids_inverse restores the full image
        print(k, case, "tiles", tuple(tiles.shape), "origins", idx.tolist()[:4], "...")
    np.savez_compressed(os.path.join(HERE, "tile_vectors.npz"), **out)
    print("wrote tile_vectors.npz")


if __name__ == "__main__":
    main()
