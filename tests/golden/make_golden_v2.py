"""Generates tests/golden/golden_v2.npz (scope rows 8(f)) by EXECUTING THE REFERENCE'S OWN CODE.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_v2.py

golden_v1.npz (rows 8(a)) is left untouched; this file adds
  * ConvReconstruction ("endtoend", models/EndToEnd.py:36-87): the reference's `encoder` Sequential run the way
    `forward` runs it (:66-78).  The class is instantiated unmodified; only its pretrained external classifier -- which
    `__init__` loads from an absolute path on the authors' machine -- is stubbed (oracle/ref_shim.py), and forward is
    evaluated up to the reshape at :78.
  * the external x-vector gender classifier (Xvector + Classifier of evaluator_inference.yaml) on feature batches.
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
    # ---------------- external x-vector classifier (evaluator_inference.yaml:34-48) ------------------
    # = EncoderClassifier.classify_batch_feats (pretrained/interfaces.py:762-776,825-829) with the reference's own
    #   Xvector and Classifier modules; StatisticsPooling's unseeded noise is pinned to a recipe tensor.
    fixed = {}
    R["StatisticsPooling"]._get_gauss_noise = lambda self, shape, device="cpu": fixed["noise"]
    emb_model = R["Xvector"](in_channels=80, activation=torch.nn.LeakyReLU, tdnn_blocks=5,
                             tdnn_channels=[512, 512, 512, 512, 1500], tdnn_kernel_sizes=[5, 3, 3, 1, 1],
                             tdnn_dilations=[1, 2, 3, 1, 1], lin_neurons=128).eval()
    clf = R["XvectorClassifier"](input_shape=[None, None, 128], activation=torch.nn.LeakyReLU, lin_blocks=1,
                                 lin_neurons=128, out_neurons=2).eval()
    w_emb = recipe.make_weights(recipe.XVECTOR_EMB_SHAPES, 1230)
    w_clf = recipe.make_weights(recipe.XVECTOR_CLF_SHAPES, 1230)
    emb_model.load_state_dict({k: t(v) for k, v in w_emb.items()})
    clf.load_state_dict({k: t(v) for k, v in w_clf.items()})
    for name, (seed, B, T, ragged) in {"xvec_a": (21, 3, 101, True), "xvec_b": (22, 2, 300, False)}.items():
        x = t(recipe.make_feats(seed, B, T))
        lens = t(np.linspace(0.6, 1.0, B).astype(np.float32)) if ragged else torch.ones(B)
        fixed["noise"] = t(recipe.make_noise(seed, B, 1500))
        with torch.no_grad():
            emb = emb_model(x, lens)                       # encode_batch_feat
            out_prob = clf(emb).squeeze(1)                 # classify_batch_feats
        out[name + "_emb"] = emb.numpy()
        out[name + "_logp"] = out_prob.numpy()

    path = os.path.join(HERE, "golden_v2.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(out), "arrays")


if __name__ == "__main__":
    main()
