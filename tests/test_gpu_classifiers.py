"""Standalone forwards of the classifier modules (callable on their own in the reference):
TDNNSexClassifier.forward (models/ConvAutoEncoder.py:57-69), FullyConnSexClassifier.forward (models/FullyConnected.py:88-104),
Xvector.forward (speechbrain/lobes/models/Xvector.py:95-108) -- each against the oracle's restatement of the same function and
against the value the parent module's fused forward returns for the same bottleneck.

Tolerances: fp32 paths (TDNN on the training kernels, FC classifier) 1e-4 abs. on log-probabilities; the x-vector runs its
TDNN in fp16 on the tensor cores: 1e-2 rel. on the embedding (BASELINE.json's 16-bit bar)."""
import numpy as np
import pytest
import torch

import recipe
from conftest import rel_max
from oracle import anonymizer_ref, xvector_ref

import speech_anonymization_b200 as sab

pytestmark = pytest.mark.gpu
DEV = "cuda"


def cu(x):
    return torch.from_numpy(np.ascontiguousarray(x)).to(DEV)


def _load(cls, args, shapes, seed=8886):
    m = cls(*args)
    w = recipe.make_weights(shapes, seed)
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
    return m.to(DEV), w


def _tw(w, dtype=torch.float64):
    return {k: torch.from_numpy(np.asarray(v)).to(dtype) if np.asarray(v).dtype.kind == "f" else torch.from_numpy(np.asarray(v))
            for k, v in w.items()}


@pytest.mark.parametrize("B,Lp", [(3, 15), (4, 180), (2, 2521)])
def test_tdnn_sex_classifier_eval_forward(B, Lp):
    m, w = _load(sab.ConvAutoencoder, (), recipe.CONVAE_SHAPES)
    clf = m.sex_classifier.eval()
    rng = np.random.Generator(np.random.PCG64(500 + Lp))
    enc = rng.standard_normal((B, 128, Lp)).astype(np.float32)
    noise = recipe.make_noise(11, B, 128)
    with torch.no_grad():
        logp = clf(cu(enc), pool_noise=cu(noise))
        ref = anonymizer_ref.tdnn_sex_classifier(torch.from_numpy(enc).double(), _tw(w), torch.from_numpy(noise).double())
    assert logp.shape == (B, 2) and logp.dtype == torch.float32
    if Lp == 15:        # one pooled frame: the unbiased std is NaN in the reference as well (pooling.py:299)
        assert bool(torch.isnan(ref).all()) and bool(torch.isnan(logp).all())
        return
    assert np.max(np.abs(logp.cpu().numpy() - ref.numpy())) < 1e-4
    with pytest.raises(NotImplementedError):        # eval-mode BatchNorm has no backward here
        clf(cu(enc).requires_grad_(True))
    with pytest.raises(RuntimeError):
        clf(cu(enc[:, :64]))
    with pytest.raises(RuntimeError):
        clf(cu(enc[:, :, :14]))


def test_tdnn_sex_classifier_train_forward_backward():
    """Train mode: batch statistics, gradient w.r.t. the input carries GradReverse's sign flip (ConvAutoEncoder.py:12-28)."""
    m, w = _load(sab.ConvAutoencoder, (), recipe.CONVAE_SHAPES)
    clf = m.sex_classifier.train()
    B, Lp = 6, 90
    rng = np.random.Generator(np.random.PCG64(77))
    enc = rng.standard_normal((B, 128, Lp)).astype(np.float32)
    noise = recipe.make_noise(12, B, 128)
    y = torch.arange(B) % 2
    x = cu(enc).requires_grad_(True)
    logp = clf(x, pool_noise=cu(noise))
    loss = torch.nn.functional.nll_loss(logp, y.to(DEV))
    loss.backward()
    xr = torch.from_numpy(enc).double().requires_grad_(True)
    wr = {k: v.clone().requires_grad_(v.dtype.is_floating_point and "running" not in k) for k, v in _tw(w).items()}
    ref = anonymizer_ref.tdnn_sex_classifier(xr, wr, torch.from_numpy(noise).double(), bn_training=True)
    torch.nn.functional.nll_loss(ref, y).backward()
    assert np.max(np.abs(logp.detach().cpu().numpy() - ref.detach().numpy())) < 1e-4
    g, gr = x.grad.cpu().numpy(), -xr.grad.numpy()          # the oracle's classifier omits GradReverse (identity forward)
    assert rel_max(g, gr) < 2e-3
    gw, gwr = clf.tdnn[3].weight.grad.cpu().numpy(), wr["sex_classifier.tdnn.3.weight"].grad.numpy()
    assert rel_max(gw, gwr) < 2e-3


@pytest.mark.parametrize("B,T", [(3, 51), (2, 1001), (64, 100)])
def test_fc_sex_classifier_forward(B, T):
    m, w = _load(sab.FullyConnectedAutoencoder, (80, B), recipe.FCAE_SHAPES)
    m.eval()
    rng = np.random.Generator(np.random.PCG64(900 + T))
    z = rng.standard_normal((B, T, 20)).astype(np.float32)
    noise = recipe.make_noise(13, B, 40)
    with torch.no_grad():
        logp = m.sex_classifier(cu(z), pool_noise=cu(noise))
        ref = anonymizer_ref.fc_sex_classifier(torch.from_numpy(z).double(), _tw(w), torch.from_numpy(noise).double())
    assert logp.shape == (B, 2)
    assert np.max(np.abs(logp.cpu().numpy() - ref.numpy())) < 1e-4
    with pytest.raises(RuntimeError):
        m.sex_classifier(cu(z[:, :, :19]))
    m.sex_classifier.train()
    with pytest.raises(NotImplementedError):
        with torch.no_grad():
            m.sex_classifier(cu(z))


def test_fc_sex_classifier_matches_fused_parent():
    """The parent's fused forward and encoder -> standalone classifier give the same log-probabilities."""
    m, w = _load(sab.FullyConnectedAutoencoder, (80, 4), recipe.FCAE_SHAPES)
    m.eval()
    B, T = 4, 333
    x = recipe.make_feats(5, B, T)
    noise = recipe.make_noise(14, B, 40)
    with torch.no_grad():
        _, logp_fused = m(cu(x), pool_noise=cu(noise))
        z = m.encoder(cu(x))                              # torch Linear/ReLU: only to make the classifier's input
        logp = m.sex_classifier(z, pool_noise=cu(noise))
    assert float((logp - logp_fused).abs().max()) < 1e-4


def test_xvector_forward():
    e = sab.Xvector(in_channels=80, lin_neurons=128)
    we = recipe.make_weights(recipe.XVECTOR_EMB_SHAPES, 1230)
    e.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in we.items()})
    e = e.to(DEV).eval()
    B, T = 3, 257
    x = recipe.make_feats(44, B, T)
    lens = np.array([1.0, 0.61, 0.8], np.float32)
    noise = recipe.make_noise(44, B, 1500)
    emb = e(cu(x), cu(lens), pool_noise=cu(noise))
    assert emb.shape == (B, 1, 128)
    with torch.no_grad():
        ref = xvector_ref.xvector_embedding(torch.from_numpy(x), torch.from_numpy(lens), {k: torch.from_numpy(np.asarray(v)) for k, v in we.items()},
                                            torch.from_numpy(noise))
    assert rel_max(emb.cpu().numpy(), ref.numpy()) < 1e-2
    # the same module inside XvectorClassifier gives the same embedding
    clf = sab.XvectorClassifier(embedding_model=e).to(DEV).eval()
    emb2 = clf.encode_batch_feat(cu(x), cu(lens), pool_noise=cu(noise))
    assert torch.equal(emb, emb2)
    # a default-constructed Xvector holds the reference's default signature (Xvector.py:45-56) and says what the CUDA path covers
    d = sab.Xvector().to(DEV).eval()
    assert d.blocks[16].w.weight.shape == (512, 3000) and d.blocks[0].conv.weight.shape[1] == 40
    with pytest.raises((NotImplementedError, ValueError)):
        d(cu(recipe.make_feats(1, 1, 50, 40)))
