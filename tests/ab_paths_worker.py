"""Runs the small golden cases through the CUDA path under whatever A/B environment switches the parent test set (the library
reads them once per process) and prints one JSON line of deviations from the committed golden vectors."""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
sys.path.insert(0, os.path.dirname(HERE))
import recipe  # noqa: E402
import speech_anonymization_b200 as sab  # noqa: E402


def rel_max(a, b):
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-30))


def main():
    g = np.load(os.path.join(HERE, "golden", "golden_v1.npz"))
    out = {}
    # Fbank (fp32 PCM, 80 and 40 mels, ragged lengths) vs the golden features
    def rel_db(a, ref):        # tests/conftest.py rel_db_err (BASELINE.json's Fbank metric)
        a, ref = np.asarray(a, np.float64), np.asarray(ref, np.float64)
        return float(np.max(np.abs(a - ref) / np.maximum(np.abs(ref), 1.0)))
    fb80, fb40 = sab.Fbank(n_mels=80), sab.Fbank()
    wav1, _ = recipe.make_wav(1, 3, 4321)
    out["fbank_noise80"] = rel_db(fb80(torch.from_numpy(wav1).cuda()).cpu().numpy(), g["fbank_noise80"])
    wav2, _ = recipe.make_wav(2, 2, 1600)
    out["fbank_noise40"] = rel_db(fb40(torch.from_numpy(wav2).cuda()).cpu().numpy(), g["fbank_noise40"])
    wav3, _ = recipe.make_wav(3, 4, 8000, ragged=True)
    out["fbank_ragged80"] = rel_db(fb80(torch.from_numpy(wav3).cuda()).cpu().numpy(), g["fbank_ragged80"])
    # ConvAutoencoder fp16 (two calls: the second runs with prepared weights and the full PDL chain)
    m = sab.ConvAutoencoder()
    w = recipe.make_weights(recipe.CONVAE_SHAPES, 8886)
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
    m = m.cuda().eval()
    m.precision = "fp16"
    for name, (seed, B, T) in {"convae_a": (5, 2, 36), "convae_b": (6, 1, 72)}.items():
        x = torch.from_numpy(recipe.make_feats(seed, B, T)).cuda()
        nz = torch.from_numpy(recipe.make_noise(seed, B, 128)).cuda()
        with torch.no_grad():
            for rep in range(2):
                rec, logp = m(x, pool_noise=nz)
                out["%s_rec_%d" % (name, rep)] = rel_max(rec.cpu().numpy(), g[name + "_recon"])
                out["%s_logp_%d" % (name, rep)] = float(np.max(np.abs(logp.cpu().numpy() - g[name + "_logp"])))
    # FullyConnectedAutoencoder
    f = sab.FullyConnectedAutoencoder(80, 3)
    wf = recipe.make_weights(recipe.FCAE_SHAPES, 8886)
    f.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in wf.items()})
    f = f.cuda().eval()
    with torch.no_grad():
        rec, logp = f(torch.from_numpy(recipe.make_feats(7, 3, 101)).cuda(), pool_noise=torch.from_numpy(recipe.make_noise(7, 3, 40)).cuda())
    out["fcae_rec"] = rel_max(rec.cpu().numpy(), g["fcae_recon"])
    out["fcae_logp"] = float(np.max(np.abs(logp.cpu().numpy() - g["fcae_logp"])))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
