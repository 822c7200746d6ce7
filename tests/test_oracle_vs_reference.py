"""CPU, only where /root/reference is mounted (this container, never the GPU box): the oracle restatement against
the UNMODIFIED reference executed live through oracle/ref_shim.py, on fresh random inputs the golden file does
not contain.  Skipped when the reference tree is absent."""
import numpy as np
import pytest
import torch

import recipe
from conftest import rel_db_err, rel_max
from oracle import anonymizer_ref, fbank_np, losses_np, normalize_np, ref_shim

pytestmark = pytest.mark.skipif(not ref_shim.available(), reason="reference tree not mounted")


@pytest.fixture(scope="module")
def R():
    torch.set_num_threads(4)
    return ref_shim.load()


def t(x):
    return torch.from_numpy(np.ascontiguousarray(x))


@pytest.mark.parametrize("seed,B,n,n_mels", [(101, 2, 16000, 80), (102, 3, 2777, 40), (103, 1, 160, 80)])
def test_fbank_live(R, seed, B, n, n_mels):
    wav, _ = recipe.make_wav(seed, B, n)
    ref = R["Fbank"](n_mels=n_mels)(t(wav)).numpy()
    assert ref.shape == (B, fbank_np.num_frames(n), n_mels)
    assert rel_db_err(fbank_np.fbank(wav, n_mels=n_mels), ref) < 2e-5


def test_input_normalization_live(R):
    wav, lens = recipe.make_wav(104, 4, 8000, ragged=True)
    feats = fbank_np.fbank(wav, n_mels=80).astype(np.float32)
    for kind in ("sentence", "batch", "global"):
        ref_mod = R["InputNormalization"](norm_type=kind, update_until_epoch=4)
        ora = normalize_np.InputNormalizationOracle(norm_type=kind, update_until_epoch=4)
        for epoch in (0, 1):                       # two training calls exercise the running average
            ref = ref_mod(t(feats.copy()), t(lens), epoch=epoch).numpy()
            assert rel_max(ora(feats, lens, epoch=epoch), ref) < 1e-5
        ref_mod.eval()
        ora.training = False
        assert rel_max(ora(feats, lens, epoch=9), ref_mod(t(feats.copy()), t(lens), epoch=9).numpy()) < 1e-5
    for T in (1, 35, 36, 37, 1001):
        assert normalize_np.padded_frames(T) == (T + 35) // 36 * 36


def test_convae_live(R):
    w = recipe.make_weights(recipe.CONVAE_SHAPES, 777)
    m = R["ConvAutoencoder"]().eval()
    m.load_state_dict({k: t(v) for k, v in w.items()})
    x = recipe.make_feats(105, 2, 36)
    noise = recipe.make_noise(105, 2, 128)
    R["StatisticsPooling"]._get_gauss_noise = lambda self, shape, device="cpu": t(noise)
    with torch.no_grad():
        rec, logp = m(t(x))
    r2, l2 = anonymizer_ref.convae_forward(x, w, noise)
    assert rel_max(r2.numpy(), rec.numpy()) < 1e-5
    assert np.abs(l2.numpy() - logp.numpy()).max() < 1e-5


def test_convrecon_live(R):
    """models/EndToEnd.py:66-78 executed from the reference tree vs the oracle restatement."""
    w = recipe.make_weights(recipe.CONVRECON_SHAPES, 779)
    m = R["ConvReconstruction"]().eval()
    m.load_state_dict({k: t(v) for k, v in w.items()}, strict=False)
    x = recipe.make_feats(106, 2, 37)
    with torch.no_grad():
        y = t(x).reshape(2, 37 * 80).unsqueeze(1)
        y = m.encoder(y).squeeze(1).reshape(2, 37, 80)
    assert rel_max(anonymizer_ref.convrecon_forward(x, w).numpy(), y.numpy()) < 1e-5


def test_xvector_live(R):
    """The reference's Xvector + Classifier (evaluator_inference.yaml) executed from the reference tree vs the oracle."""
    from oracle import xvector_ref
    we = recipe.make_weights(recipe.XVECTOR_EMB_SHAPES, 780)
    wc = recipe.make_weights(recipe.XVECTOR_CLF_SHAPES, 780)
    em = R["Xvector"](in_channels=80, activation=torch.nn.LeakyReLU, tdnn_blocks=5, tdnn_channels=[512, 512, 512, 512, 1500],
                      tdnn_kernel_sizes=[5, 3, 3, 1, 1], tdnn_dilations=[1, 2, 3, 1, 1], lin_neurons=128).eval()
    cl = R["XvectorClassifier"](input_shape=[None, None, 128], activation=torch.nn.LeakyReLU, lin_blocks=1,
                                lin_neurons=128, out_neurons=2).eval()
    em.load_state_dict({k: t(v) for k, v in we.items()})
    cl.load_state_dict({k: t(v) for k, v in wc.items()})
    x = recipe.make_feats(107, 2, 57)
    lens = np.array([0.75, 1.0], np.float32)
    noise = recipe.make_noise(107, 2, 1500)
    R["StatisticsPooling"]._get_gauss_noise = lambda self, shape, device="cpu": t(noise)
    with torch.no_grad():
        emb = em(t(x), t(lens))
        logp = cl(emb).squeeze(1)
    e2, l2 = xvector_ref.classify_batch_feats(x, lens, we, wc, noise)
    assert rel_max(e2.numpy(), emb.numpy()) < 1e-5
    assert np.abs(l2.numpy() - logp.numpy()).max() < 1e-5


def test_fcae_live(R):
    w = recipe.make_weights(recipe.FCAE_SHAPES, 778)
    m = R["FullyConnectedAutoencoder"](80, 2).eval()
    m.load_state_dict({k: t(v) for k, v in w.items()})
    x = recipe.make_feats(106, 2, 57)
    noise = recipe.make_noise(106, 2, 40)
    R["StatisticsPooling"]._get_gauss_noise = lambda self, shape, device="cpu": t(noise)
    with torch.no_grad():
        rec, logp = m(t(x))
    r2, l2 = anonymizer_ref.fcae_forward(x, w, noise)
    assert rel_max(r2.numpy(), rec.numpy()) < 1e-5
    assert np.abs(l2.numpy() - logp.numpy()).max() < 1e-5


def test_losses_live(R):
    g = np.random.Generator(np.random.PCG64(107))
    x1 = g.standard_normal((3, 20, 32)).astype(np.float32)
    x2 = g.standard_normal((3, 20, 32)).astype(np.float32)
    assert abs(float(R["CosineSimilarityLoss"]()(t(x1), t(x2))) - losses_np.cosine_similarity_loss(x1, x2)) < 1e-4
    X = g.standard_normal((24, 12)).astype(np.float32)
    y = np.arange(24) % 2
    X[y == 1] += 0.5
    d_ref = R["pairwise_dists"](t(X), R["cosine_distance_2d"]).numpy()
    assert np.abs(losses_np.pairwise_cosine_dists(X) - d_ref).max() < 1e-5
    mi_ref = float(R["ClusterMI"](n_classes=2, k=3)(t(X), t(y)))
    mi = losses_np.cluster_mi(X, y, n_classes=2, k=3)
    assert abs(float(mi[0] if isinstance(mi, tuple) else mi) - mi_ref) < 1e-4
