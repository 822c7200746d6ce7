"""Generates tests/golden/golden_v1.npz by EXECUTING THE REFERENCE'S OWN CODE.

Run here (the only place /root/reference exists):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

Inputs and weights come from tests/golden/recipe.py (numpy PCG64, rebuilt
bit-identically by the tests), so the file only stores reference OUTPUTS plus
the few inputs that cannot be regenerated (a real speech excerpt and the real
normaliser statistics shipped in the reference's results/ directory).
"""
import os
import sys
import wave
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

    # ---------------- Fbank (lobes/features.py:126-143) -------------------------
    fb80 = R["Fbank"](n_mels=80)
    fb40 = R["Fbank"]()                                    # reference default n_mels=40
    wav1, _ = recipe.make_wav(1, 3, 4321)
    out["fbank_noise80"] = fb80(t(wav1)).numpy()
    wav2, _ = recipe.make_wav(2, 2, 1600)
    out["fbank_noise40"] = fb40(t(wav2)).numpy()
    wav3, lens3 = recipe.make_wav(3, 4, 8000, ragged=True)
    out["fbank_ragged80"] = fb80(t(wav3)).numpy()
    wav4, _ = recipe.make_wav(4, 2, 160000)                # BASELINE unit of work (10 s)
    f4 = fb80(t(wav4)).numpy()
    out["fbank_10s_shape"] = np.array(f4.shape)
    out["fbank_10s_head"] = f4[:, :6]                      # first/last frames + column sums
    out["fbank_10s_tail"] = f4[:, -6:]
    out["fbank_10s_colsum"] = f4.astype(np.float64).sum(axis=1)
    # real speech excerpt shipped with the reference's SpeechBrain fork
    w = wave.open(os.path.join(ref_shim.REF_ROOT, "speechbrain/samples/audio_samples/example1.wav"))
    pcm = np.frombuffer(w.readframes(w.getnframes()), dtype=np.int16)[8000:20000].copy()
    out["speech_pcm_i16"] = pcm
    out["fbank_speech80"] = fb80(t(pcm.astype(np.float32) / 32768.0)[None]).numpy()
    out["fbank_silence80"] = fb80(torch.zeros(1, 800)).numpy()
    for nm, mod in ((80, fb80), (40, fb40)):              # the filter matrix itself
        flt = mod.compute_fbanks
        fc = flt.f_central.repeat(flt.all_freqs_mat.shape[1], 1).transpose(0, 1)
        bd = flt.band.repeat(flt.all_freqs_mat.shape[1], 1).transpose(0, 1)
        out["mel_matrix%d" % nm] = flt._create_fbank_matrix(fc, bd).numpy()

    # ---------------- InputNormalization (processing/features.py:998-1121) ------
    ckpt = os.path.join(ref_shim.REF_ROOT,
                        "results/5_5_fc/8886/save/CKPT+2022-05-06+15-54-03+00/normalizer.ckpt")
    st = torch.load(ckpt, map_location="cpu", weights_only=False)
    out["real_glob_mean"] = st["glob_mean"].float().numpy()
    out["real_glob_std"] = st["glob_std"].float().numpy()
    out["real_count"] = np.array(int(st["count"]))
    feats3 = t(out["fbank_ragged80"])
    norm = R["InputNormalization"](norm_type="global", update_until_epoch=4)
    norm.train()
    out["norm_train_call1"] = norm(feats3.clone(), t(lens3), epoch=0).numpy()
    out["norm_train_mean1"] = norm.glob_mean.numpy().copy()
    out["norm_train_std1"] = norm.glob_std.numpy().copy()
    feats1 = t(out["fbank_noise80"])
    out["norm_train_call2"] = norm(feats1.clone(), torch.ones(3), epoch=1).numpy()
    out["norm_train_mean2"] = norm.glob_mean.numpy().copy()
    out["norm_train_std2"] = norm.glob_std.numpy().copy()
    norm.eval()
    out["norm_eval_call3"] = norm(feats3.clone(), t(lens3), epoch=9).numpy()
    fresh = R["InputNormalization"](norm_type="global", update_until_epoch=4).eval()
    out["norm_eval_fresh"] = fresh(feats3.clone(), t(lens3)).numpy()   # :1114-1117 fallback
    loaded = R["InputNormalization"](norm_type="global", update_until_epoch=4).eval()
    loaded._load_statistics_dict(st)
    out["norm_eval_realstats"] = loaded(feats3.clone(), t(lens3)).numpy()
    sent = R["InputNormalization"](norm_type="sentence")
    out["norm_sentence"] = sent(feats3.clone(), t(lens3)).numpy()

    # ---------------- anonymizers ------------------------------------------------
    pool = R["StatisticsPooling"]
    fixed = {}
    pool._get_gauss_noise = lambda self, shape, device="cpu": fixed["noise"]

    w_conv = recipe.make_weights(recipe.CONVAE_SHAPES, 8886)
    m = R["ConvAutoencoder"]().eval()
    m.load_state_dict({k: t(v) for k, v in w_conv.items()})
    for name, (seed, B, T) in {"convae_a": (5, 2, 36), "convae_b": (6, 1, 72)}.items():
        x = recipe.make_feats(seed, B, T)
        fixed["noise"] = t(recipe.make_noise(seed, B, 128))
        with torch.no_grad():
            rec, logp = m(t(x))
        out[name + "_recon"] = rec.numpy()
        out[name + "_logp"] = logp.numpy()

    w_fc = recipe.make_weights(recipe.FCAE_SHAPES, 8886)
    fc = R["FullyConnectedAutoencoder"](80, 3).eval()
    fc.load_state_dict({k: t(v) for k, v in w_fc.items()})
    x = recipe.make_feats(7, 3, 101)
    fixed["noise"] = t(recipe.make_noise(7, 3, 40))
    with torch.no_grad():
        rec, logp = fc(t(x))
    out["fcae_recon"] = rec.numpy()
    out["fcae_logp"] = logp.numpy()

    # ---------------- losses -----------------------------------------------------
    g = np.random.Generator(np.random.PCG64(11))
    x1 = g.standard_normal((4, 50, 64)).astype(np.float32)
    x2 = (x1 + 0.5 * g.standard_normal((4, 50, 64))).astype(np.float32)
    out["cos_loss"] = R["CosineSimilarityLoss"]()(t(x1), t(x2)).numpy()
    X = g.standard_normal((32, 16)).astype(np.float32)
    y = np.tile(np.array([0, 1]), 16)
    X[y == 1] += 0.75
    out["mi_dists"] = R["pairwise_dists"](t(X), R["cosine_distance_2d"]).numpy()
    out["mi_value"] = R["ClusterMI"](n_classes=2, k=3)(t(X), t(y)).numpy()
    Xo = g.standard_normal((17, 8)).astype(np.float32)       # odd N exercises the roll tail
    yo = (np.arange(17) % 2)
    out["mi_dists_odd"] = R["pairwise_dists"](t(Xo), R["cosine_distance_2d"]).numpy()
    out["mi_value_odd"] = R["ClusterMI"](n_classes=2, k=3)(t(Xo), t(yo)).numpy()

    path = os.path.join(HERE, "golden_v1.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(out), "arrays")
    print("torch", torch.__version__, "numpy", np.__version__)


if __name__ == "__main__":
    main()
