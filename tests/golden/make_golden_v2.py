"""Generates tests/golden/golden_v2.npz (scope rows 8(f)) by EXECUTING THE REFERENCE'S OWN CODE.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_v2.py

golden_v1.npz (rows 8(a)) is left untouched; this file adds
  * ConvReconstruction ("endtoend", models/EndToEnd.py:36-87): the reference's `encoder` Sequential run the way
    `forward` runs it (:66-78).  The class is instantiated unmodified; only its pretrained external classifier -- which
    `__init__` loads from an absolute path on the authors' machine -- is stubbed (oracle/ref_shim.py), and forward is
    evaluated up to the reshape at :78.
Inputs / weights: tests/golden/recipe.py (numpy PCG64), rebuilt bit-identically by the tests.
"""
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
warnings.filterwarnings("ignore")

from oracle import ref_shim  # noqa: E402
import recipe  # noqa: E402


def t(x):
    return torch.from_numpy(np.ascontiguousarray(x))


def main():
    R = ref_shim.load()
    torch.set_num_threads(8)
    out = {}
    w = recipe.make_weights(recipe.CONVRECON_SHAPES, 8886)
    m = R["ConvReconstruction"]().eval()
    missing = m.load_state_dict({k: t(v) for k, v in w.items()}, strict=False)
    assert not missing.unexpected_keys and not missing.missing_keys, missing
    for name, (seed, B, T) in {"convrecon_a": (15, 2, 36), "convrecon_b": (16, 1, 101)}.items():
        x = t(recipe.make_feats(seed, B, T))
        with torch.no_grad():    # EndToEnd.py:66-78
            y = x.reshape(x.shape[0], x.shape[1] * x.shape[2]).unsqueeze(1)
            y = m.encoder(y).squeeze(1).reshape(x.shape[0], x.shape[1], x.shape[2])
        out[name + "_recon"] = y.numpy()
    path = os.path.join(HERE, "golden_v2.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(out), "arrays")


if __name__ == "__main__":
    main()
