"""GPU parity tests: the CUDA path (through the Python mirror -> C ABI) against the golden vectors
generated from the reference's own code, the reference's known-answer tests
(speechbrain/tests/unittests/test_features.py) and the CPU oracle on seeded inputs.

Tolerances (BASELINE.json): Fbank fp32 |a-ref|/max(|ref|,1) <= 1e-4; anonymizer output <= 1e-4
rel. (fp32 mode) / <= 1e-2 rel. (fp16 mode), rel = max|a-ref| / max|ref|; frame counts and
length masks bit-exact."""
import numpy as np
import pytest
import torch

import recipe
from conftest import rel_db_err, rel_max
from oracle import anonymizer_ref, fbank_np, losses_np, normalize_np

import speech_anonymization_b200 as sab

pytestmark = pytest.mark.gpu
DEV = "cuda"
FBANK_TOL = 1e-4


def cu(x):
    return torch.from_numpy(np.ascontiguousarray(x)).to(DEV)


def load_model(cls_args, shapes, seed=8886, precision=None):
    cls, args = cls_args
    m = cls(*args)
    w = recipe.make_weights(shapes, seed)
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
    m = m.to(DEV).eval()
    if precision:
        m.precision = precision
    return m, w


# ------------------------------------------------------------------ Fbank
def test_fbank_vs_golden(golden):
    fb80, fb40 = sab.Fbank(n_mels=80), sab.Fbank()
    wav1, _ = recipe.make_wav(1, 3, 4321)
    assert rel_db_err(fb80(cu(wav1)).cpu().numpy(), golden["fbank_noise80"]) < FBANK_TOL
    wav2, _ = recipe.make_wav(2, 2, 1600)
    assert rel_db_err(fb40(cu(wav2)).cpu().numpy(), golden["fbank_noise40"]) < FBANK_TOL
    wav3, _ = recipe.make_wav(3, 4, 8000, ragged=True)
    assert rel_db_err(fb80(cu(wav3)).cpu().numpy(), golden["fbank_ragged80"]) < FBANK_TOL
    pcm = golden["speech_pcm_i16"].astype(np.float32) / 32768.0
    assert rel_db_err(fb80(cu(pcm[None])).cpu().numpy(), golden["fbank_speech80"]) < FBANK_TOL
    wav4, _ = recipe.make_wav(4, 2, 160000)
    f = fb80(cu(wav4)).cpu().numpy()
    assert f.shape == tuple(golden["fbank_10s_shape"]) == (2, 1001, 80)
    assert rel_db_err(f[:, :6], golden["fbank_10s_head"]) < FBANK_TOL
    assert rel_db_err(f[:, -6:], golden["fbank_10s_tail"]) < FBANK_TOL
    assert np.max(np.abs(f.astype(np.float64).sum(axis=1) - golden["fbank_10s_colsum"])) < 0.05


def test_fbank_known_answers():
    # test_features.py:66-77: zero signal -> exactly -100 dB; doctest shape lobes/features.py:70-75
    fb = sab.Fbank(n_mels=80)
    assert torch.equal(fb(torch.zeros(1, 800, device=DEV)), torch.full((1, 6, 80), -100.0, device=DEV))
    assert sab.Fbank()(torch.zeros(10, 16000, device=DEV)).shape == (10, 101, 40)
    # test_features.py:81-88: top_db is per utterance -> batched result == per-item results
    g = torch.Generator().manual_seed(0)
    x = torch.rand(4, 16000, generator=g).to(DEV) * torch.tensor([1.0, 1e-3, 10.0, 0.1], device=DEV)[:, None]
    batched = fb(x)
    for i in range(4):
        assert torch.sum(torch.abs(fb(x[i:i + 1]) - batched[i:i + 1])) < 8e-5


def test_fbank_vs_oracle_edges():
    fb = sab.Fbank(n_mels=80)
    for seed, B, n in ((11, 1, 1), (12, 2, 159), (13, 3, 161), (14, 1, 400), (15, 2, 2561), (16, 5, 16003)):
        wav, _ = recipe.make_wav(seed, B, n)
        out = fb(cu(wav)).cpu().numpy()
        assert out.shape == (B, 1 + n // 160, 80)                      # frame count bit-exact
        assert rel_db_err(out, fbank_np.fbank(wav, n_mels=80)) < FBANK_TOL
    assert fb(torch.zeros(0, 1600, device=DEV)).shape == (0, 11, 80)   # empty batch
    # non-contiguous rows (row stride > n_samples)
    wav, _ = recipe.make_wav(17, 3, 4000)
    big = cu(np.concatenate([wav, wav], axis=1))[:, :4000]
    assert rel_db_err(fb(big).cpu().numpy(), fbank_np.fbank(wav, n_mels=80)) < FBANK_TOL
    # multi-channel input [B, n, C] -> [B, T, F, C] (processing/features.py:143-146)
    mc = cu(np.stack([wav, 0.5 * wav], axis=2))
    out = fb(mc).cpu().numpy()
    assert out.shape == (3, 26, 80, 2)
    assert rel_db_err(out[..., 1], fbank_np.fbank(0.5 * wav, n_mels=80)) < FBANK_TOL


def test_fbank_generic_dft_path():
    # a configuration outside the reference's (n_fft 200 @ 8 kHz) exercises the direct-DFT kernel
    fb = sab.Fbank(sample_rate=8000, n_fft=200, n_mels=40)
    wav, _ = recipe.make_wav(21, 2, 3000)
    ref = fbank_np.fbank(wav, n_mels=40, sample_rate=8000, n_fft=200)
    assert rel_db_err(fb(cu(wav)).cpu().numpy(), ref) < FBANK_TOL


def test_fbank_full_size_properties():
    # BASELINE size (64 x 10 s): linearity in dB (x -> a*x shifts by 20*log10 a) and idempotent clamp
    fb = sab.Fbank(n_mels=80)
    wav, _ = recipe.make_wav(31, 64, 160000)
    x = cu(wav)
    raw, umax = fb.forward_raw(x)
    raw2, _ = fb.forward_raw(4.0 * x)
    assert float((raw2 - raw - 20.0 * np.log10(4.0)).abs().max()) < 2e-3
    f = fb(x)
    assert f.shape == (64, 1001, 80)
    assert torch.equal(torch.maximum(f, f.amax(dim=(1, 2), keepdim=True) - 80.0), f)
    sub = fbank_np.fbank(wav[5:6], n_mels=80)
    assert rel_db_err(f[5:6].cpu().numpy(), sub) < FBANK_TOL


def test_fbank_int16_wire_format():
    """int16 PCM: sample = x / 32768 like read_audio -> bit-identical to the fp32 path on the converted
    samples; rows of 16001 samples start off the 8-byte vector alignment (scalar-load path) except row 0; both the
    400-point FFT kernel and the generic DFT kernel."""
    rng = np.random.default_rng(5)
    pcm = rng.integers(-32768, 32768, size=(3, 16001), dtype=np.int16)
    pcm[1, 4000:] = 0
    as_f32 = pcm.astype(np.float32) / 32768.0
    for fb in (sab.Fbank(n_mels=80), sab.Fbank(n_mels=40, n_fft=512)):
        raw_i, um_i = fb.forward_raw(torch.from_numpy(pcm).to(DEV))
        raw_f, um_f = fb.forward_raw(cu(as_f32))
        assert torch.equal(raw_i, raw_f) and torch.equal(um_i, um_f)
    ref = fbank_np.fbank(as_f32[:1].astype(np.float64), n_mels=80)
    assert rel_db_err(sab.Fbank(n_mels=80)(torch.from_numpy(pcm[:1]).to(DEV)).cpu().numpy(), ref) < FBANK_TOL


# ------------------------------------------------------------------ InputNormalization
def test_input_normalization_known_answer():
    # test_features.py:107-119
    norm = sab.InputNormalization().to(DEV)
    x = torch.tensor([1, 2, 3, 0, 0, 0], dtype=torch.float32, device=DEV).reshape(1, 6, 1)
    out = norm(x, torch.tensor([0.5], device=DEV)).squeeze()
    assert torch.equal(out.cpu(), torch.tensor([-1.0, 0.0, 1.0, -2.0, -2.0, -2.0]))


def test_input_normalization_vs_golden(golden):
    _, lens3 = recipe.make_wav(3, 4, 8000, ragged=True)
    f3, f1 = cu(golden["fbank_ragged80"]), cu(golden["fbank_noise80"])
    lens = cu(lens3)
    n = sab.InputNormalization(norm_type="global", update_until_epoch=4).to(DEV)
    n.train()
    assert rel_max(n(f3.clone(), lens, epoch=0).cpu().numpy(), golden["norm_train_call1"]) < 1e-5
    assert rel_max(n.glob_mean.cpu().numpy(), golden["norm_train_mean1"]) < 1e-5
    assert rel_max(n.glob_std.cpu().numpy(), golden["norm_train_std1"]) < 1e-5
    assert rel_max(n(f1.clone(), torch.ones(3, device=DEV), epoch=1).cpu().numpy(), golden["norm_train_call2"]) < 1e-5
    assert rel_max(n.glob_mean.cpu().numpy(), golden["norm_train_mean2"]) < 1e-5
    assert rel_max(n.glob_std.cpu().numpy(), golden["norm_train_std2"]) < 1e-5
    assert n.count == 2
    n.eval()
    assert rel_max(n(f3.clone(), lens, epoch=9).cpu().numpy(), golden["norm_eval_call3"]) < 1e-5
    fresh = sab.InputNormalization(norm_type="global", update_until_epoch=4).to(DEV).eval()
    assert rel_max(fresh(f3.clone(), lens).cpu().numpy(), golden["norm_eval_fresh"]) < 1e-5
    loaded = sab.InputNormalization(norm_type="global").to(DEV).eval()
    loaded._load_statistics_dict({"count": int(golden["real_count"]),
                                  "glob_mean": torch.from_numpy(golden["real_glob_mean"]),
                                  "glob_std": torch.from_numpy(golden["real_glob_std"]),
                                  "spk_dict_mean": {}, "spk_dict_std": {}, "spk_dict_count": {}})
    assert rel_max(loaded(f3.clone(), lens).cpu().numpy(), golden["norm_eval_realstats"]) < 1e-5
    sent = sab.InputNormalization(norm_type="sentence").to(DEV)
    assert rel_max(sent(f3.clone(), lens).cpu().numpy(), golden["norm_sentence"]) < 1e-5


def test_length_masks_bit_exact():
    n = sab.InputNormalization().to(DEV)
    rng = np.random.Generator(np.random.PCG64(5))
    for T in (101, 103, 1001, 36):
        lens = np.concatenate([rng.uniform(0.05, 1.0, 29), [0.5, 1.0, 0.25]]).astype(np.float32)
        x = cu(rng.standard_normal((32, T, 8)).astype(np.float32))
        _, _, nv = n.batch_stats(x, cu(lens))
        assert nv.cpu().tolist() == [normalize_np.valid_frames(l, T) for l in lens]
    m, s, nv = n.batch_stats(cu(np.ones((1, 4, 3), np.float32)), cu(np.array([0.25], np.float32)))
    assert int(nv[0]) == 1 and bool(torch.isnan(s).all()) and bool((m == 1).all())   # torch.std of one row = NaN
    feats = cu(rng.standard_normal((2, 1001, 80)).astype(np.float32))
    assert sab.pad_time(feats).shape == (2, 1008, 80)
    assert float(sab.pad_time(feats)[:, 1001:].abs().sum()) == 0.0


# ------------------------------------------------------------------ anonymizers
@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("fp16", 1e-2)])
def test_convae_vs_golden(golden, precision, tol):
    m, _ = load_model((sab.ConvAutoencoder, ()), recipe.CONVAE_SHAPES, precision=precision)
    for name, (seed, B, T) in {"convae_a": (5, 2, 36), "convae_b": (6, 1, 72)}.items():
        x = cu(recipe.make_feats(seed, B, T))
        with torch.no_grad():
            rec, logp = m(x, pool_noise=cu(recipe.make_noise(seed, B, 128)))
        assert rec.shape == (B, T, 80) and logp.shape == (B, 2)
        assert rel_max(rec.cpu().numpy(), golden[name + "_recon"]) < tol
        assert np.max(np.abs(logp.cpu().numpy() - golden[name + "_logp"])) < (1e-4 if precision == "fp32" else 2e-3)


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("fp16", 1e-2)])
def test_convae_vs_oracle_10s(precision, tol):
    # BASELINE unit of work: T' = 1008 (10 s padded), L = 80640; ragged chunking (B=3, chunk=2)
    m, w = load_model((sab.ConvAutoencoder, ()), recipe.CONVAE_SHAPES, precision=precision)
    m.chunk_utts = 2
    x = recipe.make_feats(41, 3, 1008)
    x[:, 1001:] = 0.0
    noise = recipe.make_noise(41, 3, 128)
    with torch.no_grad():
        rec, logp = m(cu(x), pool_noise=cu(noise))
    torch.set_num_threads(8)
    with torch.no_grad():
        r_ref, l_ref = anonymizer_ref.convae_forward(x, w, noise)
    assert rel_max(rec.cpu().numpy(), r_ref.numpy()) < tol
    assert np.max(np.abs(logp.cpu().numpy() - l_ref.numpy())) < (1e-4 if precision == "fp32" else 2e-3)


def test_convae_composed_decoder_edges():
    # The 16-bit path runs decoder.4 o decoder.5 (ConvAutoEncoder.py:166-167) as ONE composed transposed convolution
    # and takes the two rows of decoder.4 that the reference does not have back out of y[0], y[L-2], y[L-1]
    # (convae.cu: dec45_edge_kernel).  Those rows reach the first and last 8 samples of the reconstruction through
    # decoder.8 (k15): their error must be no worse than the interior's, for short and long utterances alike.
    m, w = load_model((sab.ConvAutoencoder, ()), recipe.CONVAE_SHAPES, precision="fp16")
    for seed, B, T in [(51, 3, 1), (52, 2, 2), (53, 2, 5), (54, 1, 127)]:
        x = recipe.make_feats(seed, B, T)
        with torch.no_grad():
            rec, _ = m(cu(x), pool_noise=cu(recipe.make_noise(seed, B, 128)))
            r_ref, _ = anonymizer_ref.convae_forward(x, w, recipe.make_noise(seed, B, 128))
        err = np.abs(rec.cpu().numpy().reshape(B, -1) - r_ref.numpy().reshape(B, -1)) / np.max(np.abs(r_ref.numpy()))
        assert err.max() < 1e-2
        edge = max(err[:, :8].max(), err[:, -8:].max())
        assert edge < max(2.0 * err[:, 8:-8].max(), 2e-3), (T, edge, err[:, 8:-8].max())


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("fp16", 1e-2)])
def test_convrecon_vs_golden_and_oracle(golden2, precision, tol):
    """ConvReconstruction ("endtoend", models/EndToEnd.py:36-87; scope row 8(f)-3)."""
    m, w = load_model((sab.ConvReconstruction, ()), recipe.CONVRECON_SHAPES, precision=precision)
    for name, (seed, B, T) in {"convrecon_a": (15, 2, 36), "convrecon_b": (16, 1, 101)}.items():
        with torch.no_grad():
            rec, logits = m(cu(recipe.make_feats(seed, B, T)))
        assert rec.shape == (B, T, 80) and logits is None
        assert rel_max(rec.cpu().numpy(), golden2[name + "_recon"]) < tol
    # BASELINE unit of work (10 s: T = 1001 frames, L = 80080, no x36 padding needed), ragged chunking
    m.chunk_utts = 2
    x = recipe.make_feats(42, 3, 1001)
    with torch.no_grad():
        rec, _ = m(cu(x))
    torch.set_num_threads(8)
    with torch.no_grad():
        r_ref = anonymizer_ref.convrecon_forward(x, w)
    assert rel_max(rec.cpu().numpy(), r_ref.numpy()) < tol
    # the external classifier hook (EndToEnd.py:81) and the reference's reshape failure for odd T*F
    m.sex_classifier = lambda feats: (feats.mean(dim=(1, 2)), None, None)
    with torch.no_grad():
        rec2, logits = m(cu(x[:1]))
    assert torch.equal(logits, rec2.mean(dim=(1, 2)))
    with pytest.raises(RuntimeError):
        m(cu(np.zeros((1, 3, 5), np.float32)))


def _load_xvector(seed=1230):
    from oracle import xvector_ref  # noqa: F401
    m = sab.XvectorClassifier()
    we = recipe.make_weights(recipe.XVECTOR_EMB_SHAPES, seed)
    wc = recipe.make_weights(recipe.XVECTOR_CLF_SHAPES, seed)
    m.embedding_model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in we.items()})
    m.classifier.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in wc.items()})
    return m.to(DEV).eval(), we, wc


def test_xvector_classifier_vs_golden_and_oracle(golden2):
    """External x-vector gender classifier on features (scope row 8(f)-2): fp16 tensor-core TDNN, <= 1e-2 rel. on the
    x-vector, <= 1e-2 abs. on the log posteriors; valid-frame counts of the pooling bit-exact (ragged lengths)."""
    from oracle import xvector_ref
    m, we, wc = _load_xvector()
    for name, (seed, B, T, ragged) in {"xvec_a": (21, 3, 101, True), "xvec_b": (22, 2, 300, False)}.items():
        lens = cu(np.linspace(0.6, 1.0, B).astype(np.float32)) if ragged else None
        noise = cu(recipe.make_noise(seed, B, 1500))
        x = cu(recipe.make_feats(seed, B, T))
        emb = m.encode_batch_feat(x, lens, pool_noise=noise)
        out_prob, score, index = m.classify_batch_feats(x, lens, pool_noise=noise)
        assert emb.shape == (B, 1, 128) and out_prob.shape == (B, 2) and score.shape == (B,) and index.shape == (B,)
        assert rel_max(emb.cpu().numpy(), golden2[name + "_emb"]) < 1e-2
        assert np.max(np.abs(out_prob.cpu().numpy() - golden2[name + "_logp"])) < 1e-2
        assert torch.equal(index.cpu(), torch.from_numpy(golden2[name + "_logp"]).argmax(dim=-1))
    # BASELINE unit of work (10 s = 1001 frames, last 128-frame tile partly empty), ragged lengths, weights reused
    B, T = 4, 1001
    x = recipe.make_feats(43, B, T)
    lens = np.array([1.0, 0.5, 0.7513, 0.9], np.float32)
    noise = recipe.make_noise(43, B, 1500)
    emb = m.encode_batch_feat(cu(x), cu(lens), pool_noise=cu(noise))
    out_prob, _, _ = m.classify_batch_feats(cu(x), cu(lens), pool_noise=cu(noise))
    torch.set_num_threads(8)
    with torch.no_grad():
        e_ref, l_ref = xvector_ref.classify_batch_feats(x, lens, we, wc, noise)
    assert rel_max(emb.cpu().numpy(), e_ref.numpy()) < 1e-2
    assert np.max(np.abs(out_prob.cpu().numpy() - l_ref.numpy())) < 1e-2
    # an in-place weight update must be seen (packed weights are re-made)
    with torch.no_grad():
        m.classifier.out.w.bias.add_(torch.tensor([0.5, -0.5], device=DEV))
    out2, _, _ = m.classify_batch_feats(cu(x), cu(lens), pool_noise=cu(noise))
    assert float((out2 - out_prob).abs().max()) > 0.1
    # the composed evaluation path of gender_classifier_recon.yaml: anonymizer output -> external classifier
    cr, _ = load_model((sab.ConvReconstruction, ()), recipe.CONVRECON_SHAPES, precision="fp16")
    cr.sex_classifier = m.classify_batch_feats
    with torch.no_grad():
        rec, logits = cr(cu(x[:2]))
    assert logits.shape == (2, 2) and bool(torch.isfinite(logits).all())


def test_convae_config3_batch_is_utterance_independent(golden):
    """BASELINE config 3 size (batch 256 x 10 s through Fbank -> ConvAE, fp16): utterances are independent on the
    inference path, so every utterance's result must not depend on its batch, chunk (60/60/60/60/16) or position."""
    m, _ = load_model((sab.ConvAutoencoder, ()), recipe.CONVAE_SHAPES, precision="fp16")
    pipe = _pipeline(m, golden)
    B = 256
    g = torch.Generator().manual_seed(7)
    wav = (0.1 * torch.randn(B, 160000, generator=g)).to(DEV)
    noise = cu(recipe.make_noise(7, B, 128))
    rec, logp = pipe(wav, pool_noise=noise)
    assert rec.shape == (B, 1008, 80) and logp.shape == (B, 2) and bool(torch.isfinite(rec).all())
    for i in (0, 59, 60, 255):
        r1, l1 = pipe(wav[i:i + 1].contiguous(), pool_noise=noise[i:i + 1].contiguous())
        assert rel_max(rec[i:i + 1].cpu().numpy(), r1.cpu().numpy()) < 1e-5     # (fp64 atomics: summation order)
        assert float((logp[i:i + 1] - l1).abs().max()) < 1e-4
    assert float(rec[:, 1001:].abs().max()) < 50.0 and float((rec[0] - rec[1]).abs().max()) > 0.0


def test_convae_error_behaviour():
    m, _ = load_model((sab.ConvAutoencoder, ()), recipe.CONVAE_SHAPES)
    with pytest.raises(RuntimeError):           # the reference's reshape fails when T*F % 4 != 0
        with torch.no_grad():
            m(torch.zeros(1, 37, 81, device=DEV))
    with pytest.raises(NotImplementedError):    # backward pass is a later scope row
        m(torch.zeros(1, 36, 80, device=DEV))
    with torch.no_grad():
        rec, logp = m(torch.zeros(0, 36, 80, device=DEV))
    assert rec.shape == (0, 36, 80) and logp.shape == (0, 2)


def test_fcae_vs_golden_and_oracle(golden):
    m, w = load_model((sab.FullyConnectedAutoencoder, (80, 3)), recipe.FCAE_SHAPES)
    x = recipe.make_feats(7, 3, 101)
    with torch.no_grad():
        rec, logp = m(cu(x), pool_noise=cu(recipe.make_noise(7, 3, 40)))
    assert rel_max(rec.cpu().numpy(), golden["fcae_recon"]) < 1e-4
    assert np.max(np.abs(logp.cpu().numpy() - golden["fcae_logp"])) < 1e-4
    # BASELINE config 2 shape (64 x 10 s, T = 1001: frame tiles straddle utterances)
    x = recipe.make_feats(8, 64, 1001)
    noise = recipe.make_noise(8, 64, 40)
    with torch.no_grad():
        rec, logp = m(cu(x), pool_noise=cu(noise))
        r_ref, l_ref = anonymizer_ref.fcae_forward(x, w, noise)
    assert rel_max(rec.cpu().numpy(), r_ref.numpy()) < 1e-4
    assert np.max(np.abs(logp.cpu().numpy() - l_ref.numpy())) < 1e-4


# ------------------------------------------------------------------ composed path
def _pipeline(model, golden):
    fb = sab.Fbank(n_mels=80)
    norm = sab.InputNormalization(norm_type="global", update_until_epoch=4)
    norm._load_statistics_dict({"count": int(golden["real_count"]),
                                "glob_mean": torch.from_numpy(golden["real_glob_mean"]),
                                "glob_std": torch.from_numpy(golden["real_glob_std"]),
                                "spk_dict_mean": {}, "spk_dict_std": {}, "spk_dict_count": {}})
    norm = norm.to(DEV).eval()
    return sab.AnonymizerPipeline(fb, norm, model)


@pytest.mark.parametrize("kind", ["convae", "fcae"])
def test_pipeline_fused_equals_modular_equals_oracle(golden, kind):
    if kind == "convae":
        m, w = load_model((sab.ConvAutoencoder, ()), recipe.CONVAE_SHAPES, precision="fp32")
        nz = 128
    else:
        m, w = load_model((sab.FullyConnectedAutoencoder, (80, 4)), recipe.FCAE_SHAPES)
        nz = 40
    pipe = _pipeline(m, golden)
    wav, lens = recipe.make_wav(51, 4, 16000, ragged=True)
    noise = recipe.make_noise(51, 4, nz)
    rec_f, logp_f = pipe(cu(wav), cu(lens), pool_noise=cu(noise))          # fused clamp/norm/pad
    feats = pipe.normalize(pipe.compute_features(cu(wav)), cu(lens))
    if kind == "convae":
        feats = sab.pad_time(feats)
        assert feats.shape[1] == 108
    with torch.no_grad():
        rec_m, logp_m = m(feats, pool_noise=cu(noise))                        # module-by-module
    assert rel_max(rec_f.cpu().numpy(), rec_m.cpu().numpy()) < 1e-5
    assert float((logp_f - logp_m).abs().max()) < 1e-5
    # oracle end to end (fp64 Fbank -> normalise with the real statistics -> pad -> anonymizer)
    f = fbank_np.fbank(wav, n_mels=80)
    f = (f - golden["real_glob_mean"].astype(np.float64)) / golden["real_glob_std"].astype(np.float64)
    if kind == "convae":
        f = normalize_np.pad_time(f)
        r_ref, l_ref = anonymizer_ref.convae_forward(f.astype(np.float32), w, noise)
    else:
        r_ref, l_ref = anonymizer_ref.fcae_forward(f.astype(np.float32), w, noise)
    assert rel_max(rec_f.cpu().numpy(), r_ref.numpy()) < 1e-4
    assert np.max(np.abs(logp_f.cpu().numpy() - l_ref.numpy())) < 1e-4
    # host-buffer C entry point == device path
    pcm = torch.from_numpy(wav).pin_memory()
    rec_h = torch.empty(rec_f.shape, dtype=torch.float32).pin_memory()
    logp_h = torch.empty((4, 2), dtype=torch.float32).pin_memory()
    ws = pipe.host_workspace(4, 16000, torch.device(DEV))
    pipe.anonymize_host(pcm, rec_h, logp_h, ws, pool_noise=cu(noise))
    assert rel_max(rec_h.numpy(), rec_f.cpu().numpy()) < 1e-6
    assert float((logp_h - logp_f.cpu()).abs().max()) < 1e-6
    # streaming pipe (three streams, two batches in flight, weights packed once per slot) == the one-shot call,
    # including ragged batch sizes and slot reuse
    torch.cuda.synchronize()
    sp = pipe.stream_pipe(4, 16000, torch.device(DEV), depth=2)
    outs = []
    for i, nb in enumerate([4, 3, 4, 1, 4]):
        r = torch.full((nb,) + tuple(rec_f.shape[1:]), float("nan"), dtype=torch.float32).pin_memory()
        l = torch.full((nb, 2), float("nan"), dtype=torch.float32).pin_memory()
        sp.submit(pcm[:nb], r, l, pool_noise=cu(noise)[:nb].contiguous())
        outs.append((nb, r, l))
    sp.drain()
    for nb, r, l in outs:
        assert rel_max(r.numpy(), rec_h.numpy()[:nb]) < 1e-6      # (fp64 atomics: summation order may differ)
        assert float(np.abs(l.numpy() - logp_h.numpy()[:nb]).max()) < 1e-6
    sp.close()
    # int16 PCM on the wire (half the H2D bytes) == fp32 path on x / 32768, one-shot and streaming
    pcm16 = torch.from_numpy(np.clip(np.round(wav * 32768.0), -32768, 32767).astype(np.int16)).pin_memory()
    pcm32 = (pcm16.to(torch.float32) / 32768.0).pin_memory()
    r32 = torch.empty(rec_f.shape, dtype=torch.float32).pin_memory()
    r16 = torch.empty(rec_f.shape, dtype=torch.float32).pin_memory()
    l32 = torch.empty((4, 2), dtype=torch.float32).pin_memory()
    l16 = torch.empty((4, 2), dtype=torch.float32).pin_memory()
    pipe.anonymize_host(pcm32, r32, l32, ws, pool_noise=cu(noise))
    pipe.anonymize_host(pcm16, r16, l16, ws, pool_noise=cu(noise))
    assert rel_max(r16.numpy(), r32.numpy()) < 1e-6 and float((l16 - l32).abs().max()) < 1e-6
    sp = pipe.stream_pipe(4, 16000, torch.device(DEV), depth=2)
    r16s = torch.full(rec_f.shape, float("nan"), dtype=torch.float32).pin_memory()
    sp.submit(pcm16, r16s, l16, pool_noise=cu(noise))
    sp.drain()
    sp.close()
    assert rel_max(r16s.numpy(), r32.numpy()) < 1e-6


def test_graphed_step_equals_direct_launches(golden):
    """AnonymizerPipeline.capture: replaying the captured launch sequence gives the direct call's result, also after the
    input buffer has been overwritten in place (fp16 mode: the path bench.py times)."""
    m, _ = load_model((sab.ConvAutoencoder, ()), recipe.CONVAE_SHAPES, precision="fp16")
    pipe = _pipeline(m, golden)
    wav, _ = recipe.make_wav(61, 3, 16000)
    wav2, _ = recipe.make_wav(62, 3, 16000)
    noise = cu(recipe.make_noise(61, 3, 128))
    buf = cu(wav).clone()
    rec_d, logp_d = pipe(buf, pool_noise=noise)
    rec_d2, logp_d2 = pipe(cu(wav2), pool_noise=noise)
    g = pipe.capture(buf, pool_noise=noise)
    assert g.launches >= 19
    r, l = g.replay()
    torch.cuda.synchronize()
    assert rel_max(r.cpu().numpy(), rec_d.cpu().numpy()) < 1e-6 and float((l - logp_d).abs().max()) < 1e-6
    buf.copy_(cu(wav2))
    r, l = g.replay()
    torch.cuda.synchronize()
    assert rel_max(r.cpu().numpy(), rec_d2.cpu().numpy()) < 1e-6 and float((l - logp_d2).abs().max()) < 1e-6


def test_convae_sees_weight_updates():
    """The module keeps its workspace (with the packed weights) between calls; an in-place parameter update must
    invalidate that cache (key = storage pointer + version counter of every parameter and buffer)."""
    m, w = load_model((sab.ConvAutoencoder, ()), recipe.CONVAE_SHAPES)
    x = cu(recipe.make_feats(31, 2, 36))
    noise = cu(recipe.make_noise(31, 2, 128))
    with torch.no_grad():
        r0, _ = m(x, pool_noise=noise)
        r1, _ = m(x, pool_noise=noise)                    # second call reuses the packed weights
        assert rel_max(r1.cpu().numpy(), r0.cpu().numpy()) < 1e-6
        m.decoder[8].weight.mul_(2.0)                     # recon = conv(...) + bias: doubles the conv part
        r2, _ = m(x, pool_noise=noise)
        bias = float(m.decoder[8].bias[0])
        assert rel_max((r2 - bias).cpu().numpy(), (2.0 * (r0 - bias)).cpu().numpy()) < 2e-3
        m.sex_classifier.tdnn[2].running_mean.add_(0.5)   # a buffer (folded BatchNorm) changes the log-probs only
        r3, l3 = m(x, pool_noise=noise)
        assert rel_max(r3.cpu().numpy(), r2.cpu().numpy()) < 1e-6


# ------------------------------------------------------------------ losses
def test_cosine_loss(golden):
    g = np.random.Generator(np.random.PCG64(11))
    x1 = g.standard_normal((4, 50, 64)).astype(np.float32)
    x2 = (x1 + 0.5 * g.standard_normal((4, 50, 64))).astype(np.float32)
    loss_fn = sab.CosineSimilarityLoss()
    a, b = cu(x1).requires_grad_(True), cu(x2).requires_grad_(True)
    loss = loss_fn(a, b)
    assert abs(float(loss) - float(golden["cos_loss"])) < 1e-4
    assert abs(float(loss) - losses_np.cosine_similarity_loss(x1, x2)) < 1e-4
    (2.0 * loss).backward()
    ta, tb = torch.from_numpy(x1).double().requires_grad_(True), torch.from_numpy(x2).double().requires_grad_(True)
    ref = 2.0 * torch.sum(1 - torch.nn.functional.cosine_similarity(ta, tb, dim=2, eps=1e-6)) / 50
    ref.backward()
    assert rel_max(a.grad.cpu().numpy(), ta.grad.numpy()) < 1e-5
    assert rel_max(b.grad.cpu().numpy(), tb.grad.numpy()) < 1e-5
    z = torch.zeros(2, 3, 8, device=DEV)        # zero vectors: cos = 0 -> loss = B (eps clamp, no NaN)
    assert abs(float(loss_fn(z, z)) - 2.0) < 1e-6


def test_knn_mi(golden):
    g = np.random.Generator(np.random.PCG64(11))
    g.standard_normal((4, 50, 64)); g.standard_normal((4, 50, 64))     # same stream position as make_golden.py
    X = g.standard_normal((32, 16)).astype(np.float32)
    y = np.tile(np.array([0, 1]), 16)
    X[y == 1] += 0.75
    mi, m_i, d = sab.knn_mi(cu(X), y, None, 2, 3, return_dists=True)
    assert np.max(np.abs(d[0].cpu().numpy() - golden["mi_dists"])) < 1e-6
    assert abs(float(mi[0]) - float(golden["mi_value"])) < 1e-5
    m_ref, _ = losses_np.knn_mi_counts(golden["mi_dists"].astype(np.float64), y, 3)
    assert m_i[0].cpu().tolist() == m_ref.tolist()                      # integer counts bit-exact
    assert abs(float(sab.ClusterMI(2, 3)(cu(X), cu(y))) - float(golden["mi_value"])) < 1e-5
    Xo = g.standard_normal((17, 8)).astype(np.float32)
    yo = np.arange(17) % 2
    mi, _, d = sab.knn_mi(cu(Xo), yo, None, 2, 3, return_dists=True)
    assert np.max(np.abs(d[0].cpu().numpy() - golden["mi_dists_odd"])) < 1e-6
    assert abs(float(mi[0]) - float(golden["mi_value_odd"])) < 1e-5
    # GroupSamplingMI: 100 index sets in one launch == the oracle evaluated per set
    groups = [i % 4 for i in range(32)]
    np.random.seed(7)
    gmi = sab.GroupSamplingMI(n_samples=4, n_classes=2, n_iterations=5)
    vals, mean, std = gmi(cu(X), cu(y), groups)
    np.random.seed(7)
    gs = sab.GroupSampler(groups, 4)
    for v in vals:
        idx = gs.sample_groups()
        assert abs(float(v) - losses_np.cluster_mi(X[idx], y[idx])) < 1e-4
    assert len(sab.MILoss()(cu(X), cu(y), groups, 4)) == 100
