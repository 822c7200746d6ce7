"""CPU: pins the oracle restatement (oracle/*.py) against golden vectors produced
by executing the reference's own code (tests/golden/make_golden.py) and against
the reference's known-answer tests (speechbrain/tests/unittests/test_features.py)."""
import numpy as np
import torch

import recipe
from conftest import rel_db_err, rel_max
from oracle import fbank_np, normalize_np, anonymizer_ref, losses_np


def test_fbank_oracle_vs_golden(golden):
    wav1, _ = recipe.make_wav(1, 3, 4321)
    assert rel_db_err(fbank_np.fbank(wav1, n_mels=80), golden["fbank_noise80"]) < 2e-5
    wav2, _ = recipe.make_wav(2, 2, 1600)
    assert rel_db_err(fbank_np.fbank(wav2, n_mels=40), golden["fbank_noise40"]) < 2e-5
    wav3, _ = recipe.make_wav(3, 4, 8000, ragged=True)
    assert rel_db_err(fbank_np.fbank(wav3, n_mels=80), golden["fbank_ragged80"]) < 2e-5
    pcm = golden["speech_pcm_i16"].astype(np.float32) / 32768.0
    assert rel_db_err(fbank_np.fbank(pcm[None], n_mels=80), golden["fbank_speech80"]) < 2e-5


def test_fbank_oracle_10s_and_frame_count(golden):
    wav4, _ = recipe.make_wav(4, 2, 160000)
    f = fbank_np.fbank(wav4, n_mels=80)
    assert tuple(golden["fbank_10s_shape"]) == f.shape == (2, 1001, 80)
    assert rel_db_err(f[:, :6], golden["fbank_10s_head"]) < 2e-5
    assert rel_db_err(f[:, -6:], golden["fbank_10s_tail"]) < 2e-5
    assert np.max(np.abs(f.sum(axis=1) - golden["fbank_10s_colsum"])) < 0.05
    for n in (0, 1, 159, 160, 161, 400, 16000, 160000):
        assert fbank_np.num_frames(n) == 1 + n // 160


def test_fbank_known_answers(golden):
    # test_features.py:66-77 -- zero power -> exactly -100 dB (amin, multiplier, flat top_db)
    assert np.all(fbank_np.amplitude_to_db(np.zeros((10, 101, 40))) == -100.0)
    assert np.all(fbank_np.fbank(np.zeros((1, 800)), n_mels=80) == -100.0)
    assert np.all(golden["fbank_silence80"] == -100.0)
    # lobes/features.py:70-75 doctest shape
    assert fbank_np.fbank(np.zeros((10, 16000)), n_mels=40).shape == (10, 101, 40)
    # mel matrix has 387 non-zeros for the 80-mel / 201-bin configuration
    assert int((fbank_np.mel_matrix(80) > 0).sum()) == 387


def test_input_normalization_known_answer():
    # test_features.py:107-119
    norm = normalize_np.InputNormalizationOracle()
    x = np.array([1, 2, 3, 0, 0, 0], dtype=np.float64).reshape(1, 6, 1)
    out = norm(x, np.array([0.5], dtype=np.float32)).squeeze()
    assert np.array_equal(out, np.array([-1, 0, 1, -2, -2, -2], dtype=np.float64))


def test_input_normalization_vs_golden(golden):
    _, lens3 = recipe.make_wav(3, 4, 8000, ragged=True)
    f3, f1 = golden["fbank_ragged80"], golden["fbank_noise80"]
    n = normalize_np.InputNormalizationOracle(norm_type="global", update_until_epoch=4)
    assert rel_max(n(f3, lens3, epoch=0), golden["norm_train_call1"]) < 1e-5
    assert rel_max(n.glob_mean, golden["norm_train_mean1"]) < 1e-5
    assert rel_max(n(f1, np.ones(3, np.float32), epoch=1), golden["norm_train_call2"]) < 1e-5
    assert rel_max(n.glob_mean, golden["norm_train_mean2"]) < 1e-5
    assert rel_max(n.glob_std, golden["norm_train_std2"]) < 1e-5
    n.training = False
    assert rel_max(n(f3, lens3, epoch=9), golden["norm_eval_call3"]) < 1e-5
    fresh = normalize_np.InputNormalizationOracle(norm_type="global")
    fresh.training = False
    assert rel_max(fresh(f3, lens3), golden["norm_eval_fresh"]) < 1e-5
    loaded = normalize_np.InputNormalizationOracle(norm_type="global")
    loaded.training = False
    loaded.glob_mean = golden["real_glob_mean"].astype(np.float64)
    loaded.glob_std = golden["real_glob_std"].astype(np.float64)
    assert rel_max(loaded(f3, lens3), golden["norm_eval_realstats"]) < 1e-5
    sent = normalize_np.InputNormalizationOracle(norm_type="sentence")
    assert rel_max(sent(f3, lens3), golden["norm_sentence"]) < 1e-5


def test_padding_and_masks():
    assert normalize_np.padded_frames(1001) == 1008
    assert normalize_np.padded_frames(1008) == 1008
    assert normalize_np.padded_frames(1) == 36
    # torch.round is round-half-to-even on the fp32 product (features.py:1021)
    assert normalize_np.valid_frames(0.5, 101) == 50
    assert normalize_np.valid_frames(0.5, 103) == 52
    assert normalize_np.valid_frames(1.0, 1001) == 1001
    x = np.ones((2, 5, 3), np.float32)
    p = normalize_np.pad_time(x)
    assert p.shape == (2, 36, 3) and p[:, 5:].sum() == 0 and p[:, :5].sum() == 30


def test_convae_oracle_vs_golden(golden):
    w = recipe.make_weights(recipe.CONVAE_SHAPES, 8886)
    for name, (seed, B, T) in {"convae_a": (5, 2, 36), "convae_b": (6, 1, 72)}.items():
        x = recipe.make_feats(seed, B, T)
        rec, logp = anonymizer_ref.convae_forward(x, w, recipe.make_noise(seed, B, 128))
        assert rel_max(rec.numpy(), golden[name + "_recon"]) < 1e-5
        assert np.max(np.abs(logp.numpy() - golden[name + "_logp"])) < 1e-5


def test_convrecon_oracle_vs_golden(golden2):
    w = recipe.make_weights(recipe.CONVRECON_SHAPES, 8886)
    for name, (seed, B, T) in {"convrecon_a": (15, 2, 36), "convrecon_b": (16, 1, 101)}.items():
        rec = anonymizer_ref.convrecon_forward(recipe.make_feats(seed, B, T), w)
        assert rel_max(rec.numpy(), golden2[name + "_recon"]) < 1e-5


def test_xvector_oracle_vs_golden(golden2):
    from oracle import xvector_ref
    we = recipe.make_weights(recipe.XVECTOR_EMB_SHAPES, 1230)
    wc = recipe.make_weights(recipe.XVECTOR_CLF_SHAPES, 1230)
    for name, (seed, B, T, ragged) in {"xvec_a": (21, 3, 101, True), "xvec_b": (22, 2, 300, False)}.items():
        lens = np.linspace(0.6, 1.0, B).astype(np.float32) if ragged else None
        emb, logp = xvector_ref.classify_batch_feats(recipe.make_feats(seed, B, T), lens, we, wc,
                                                     recipe.make_noise(seed, B, 1500))
        assert emb.shape == (B, 1, 128) and logp.shape == (B, 2)
        assert rel_max(emb.numpy(), golden2[name + "_emb"]) < 1e-5
        assert np.max(np.abs(logp.numpy() - golden2[name + "_logp"])) < 1e-5


def test_fcae_oracle_vs_golden(golden):
    w = recipe.make_weights(recipe.FCAE_SHAPES, 8886)
    x = recipe.make_feats(7, 3, 101)
    rec, logp = anonymizer_ref.fcae_forward(x, w, recipe.make_noise(7, 3, 40))
    assert rel_max(rec.numpy(), golden["fcae_recon"]) < 1e-5
    assert np.max(np.abs(logp.numpy() - golden["fcae_logp"])) < 1e-5


def test_losses_oracle_vs_golden(golden):
    g = np.random.Generator(np.random.PCG64(11))
    x1 = g.standard_normal((4, 50, 64)).astype(np.float32)
    x2 = (x1 + 0.5 * g.standard_normal((4, 50, 64))).astype(np.float32)
    assert abs(losses_np.cosine_similarity_loss(x1, x2) - float(golden["cos_loss"])) < 1e-5
    X = g.standard_normal((32, 16)).astype(np.float32)
    y = np.tile(np.array([0, 1]), 16)
    X[y == 1] += 0.75
    D = losses_np.pairwise_cosine_dists(X)
    assert np.max(np.abs(D - golden["mi_dists"])) < 1e-6
    assert abs(losses_np.cluster_mi(X, y) - float(golden["mi_value"])) < 1e-5
    Xo = g.standard_normal((17, 8)).astype(np.float32)
    yo = np.arange(17) % 2
    assert np.max(np.abs(losses_np.pairwise_cosine_dists(Xo) - golden["mi_dists_odd"])) < 1e-6
    assert abs(losses_np.cluster_mi(Xo, yo) - float(golden["mi_value_odd"])) < 1e-5


def test_mel_matrix_bit_exact(golden):
    # host-side filter construction must reproduce the reference's fp32 matrix exactly
    assert np.array_equal(fbank_np.mel_matrix_f32(80), golden["mel_matrix80"])
    assert np.array_equal(fbank_np.mel_matrix_f32(40), golden["mel_matrix40"])
    assert np.max(np.abs(fbank_np.mel_matrix(80) - golden["mel_matrix80"])) < 2e-5
