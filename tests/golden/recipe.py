"""Deterministic, torch-independent inputs and weights for the golden vectors.

Everything is drawn from numpy's PCG64 (stable across numpy versions and
platforms), keyed by (seed, crc32(name)), so the generator script (which runs
the reference, here) and the tests (which run on the GPU box, without the
reference) rebuild bit-identical inputs without storing them.
"""
import zlib
import numpy as np

CONVAE_SHAPES = {
    "encoder.0.weight": (32, 1, 15), "encoder.0.bias": (32,),
    "encoder.2.weight": (64, 32, 5), "encoder.2.bias": (64,),
    "encoder.3.weight": (64,), "encoder.3.bias": (64,),
    "encoder.5.weight": (64, 64, 5), "encoder.5.bias": (64,),
    "encoder.6.weight": (64,), "encoder.6.bias": (64,),
    "encoder.8.weight": (128, 64, 5), "encoder.8.bias": (128,),
    "encoder.9.weight": (128,), "encoder.9.bias": (128,),
    "encoder.11.weight": (128, 128, 5), "encoder.11.bias": (128,),
    "encoder.12.weight": (128,), "encoder.12.bias": (128,),
    "decoder.0.weight": (128, 128, 5), "decoder.0.bias": (128,),
    "decoder.1.weight": (128, 64, 5), "decoder.1.bias": (64,),
    "decoder.2.weight": (64,), "decoder.2.bias": (64,),
    "decoder.4.weight": (64, 64, 5), "decoder.4.bias": (64,),
    "decoder.5.weight": (64, 32, 5), "decoder.5.bias": (32,),
    "decoder.6.weight": (32,), "decoder.6.bias": (32,),
    "decoder.8.weight": (1, 32, 15), "decoder.8.bias": (1,),
    "sex_classifier.tdnn.0.weight": (128, 128, 5), "sex_classifier.tdnn.0.bias": (128,),
    "sex_classifier.tdnn.3.weight": (128, 128, 3), "sex_classifier.tdnn.3.bias": (128,),
    "sex_classifier.tdnn.6.weight": (128, 128, 3), "sex_classifier.tdnn.6.bias": (128,),
    "sex_classifier.classify.0.weight": (128, 256), "sex_classifier.classify.0.bias": (128,),
    "sex_classifier.classify.3.weight": (64, 128), "sex_classifier.classify.3.bias": (64,),
    "sex_classifier.classify.6.weight": (2, 64), "sex_classifier.classify.6.bias": (2,),
}
for _bn, _c in (("tdnn.2", 128), ("tdnn.5", 128), ("tdnn.8", 128), ("norm", 128),
                ("classify.2", 128), ("classify.5", 64)):
    for _s in ("weight", "bias", "running_mean", "running_var"):
        CONVAE_SHAPES["sex_classifier.%s.%s" % (_bn, _s)] = (_c,)
    CONVAE_SHAPES["sex_classifier.%s.num_batches_tracked" % _bn] = ()

CONVRECON_SHAPES = {   # models/EndToEnd.py:40-54
    "encoder.0.weight": (32, 1, 15), "encoder.0.bias": (32,),
    "encoder.1.weight": (32,), "encoder.1.bias": (32,),
    "encoder.3.weight": (64, 32, 5), "encoder.3.bias": (64,),
    "encoder.4.weight": (64,), "encoder.4.bias": (64,),
    "encoder.6.weight": (64, 64, 5), "encoder.6.bias": (64,),
    "encoder.7.weight": (64,), "encoder.7.bias": (64,),
    "encoder.9.weight": (64, 32, 5), "encoder.9.bias": (32,),
    "encoder.10.weight": (32,), "encoder.10.bias": (32,),
    "encoder.12.weight": (1, 32, 15), "encoder.12.bias": (1,),
}

# evaluator_inference.yaml:34-48 (speechbrain/lobes/models/Xvector.py): embedding_model.ckpt / classifier.ckpt keys
XVECTOR_EMB_SHAPES = {"blocks.16.w.weight": (128, 3000), "blocks.16.w.bias": (128,)}
for _i, (_cin, _cout, _k) in enumerate(((80, 512, 5), (512, 512, 3), (512, 512, 3), (512, 512, 1), (512, 1500, 1))):
    XVECTOR_EMB_SHAPES["blocks.%d.conv.weight" % (3 * _i)] = (_cout, _cin, _k)
    XVECTOR_EMB_SHAPES["blocks.%d.conv.bias" % (3 * _i)] = (_cout,)
    for _s in ("weight", "bias", "running_mean", "running_var"):
        XVECTOR_EMB_SHAPES["blocks.%d.norm.%s" % (3 * _i + 2, _s)] = (_cout,)
    XVECTOR_EMB_SHAPES["blocks.%d.norm.num_batches_tracked" % (3 * _i + 2)] = ()
XVECTOR_CLF_SHAPES = {"DNN.block_0.linear.w.weight": (128, 128), "DNN.block_0.linear.w.bias": (128,),
                      "out.w.weight": (2, 128), "out.w.bias": (2,)}
for _p in ("norm.norm", "DNN.block_0.norm.norm"):
    for _s in ("weight", "bias", "running_mean", "running_var"):
        XVECTOR_CLF_SHAPES["%s.%s" % (_p, _s)] = (128,)
    XVECTOR_CLF_SHAPES["%s.num_batches_tracked" % _p] = ()

FCAE_SHAPES = {
    "encoder.0.weight": (60, 80), "encoder.0.bias": (60,),
    "encoder.2.weight": (40, 60), "encoder.2.bias": (40,),
    "encoder.4.weight": (20, 40), "encoder.4.bias": (20,),
    "decoder.0.weight": (40, 20), "decoder.0.bias": (40,),
    "decoder.2.weight": (60, 40), "decoder.2.bias": (60,),
    "decoder.4.weight": (80, 60), "decoder.4.bias": (80,),
    "sex_classifier.initial.0.weight": (40, 20), "sex_classifier.initial.0.bias": (40,),
    "sex_classifier.initial.2.weight": (40, 40), "sex_classifier.initial.2.bias": (40,),
    "sex_classifier.classify.0.weight": (40, 80), "sex_classifier.classify.0.bias": (40,),
    "sex_classifier.classify.3.weight": (40, 40), "sex_classifier.classify.3.bias": (40,),
    "sex_classifier.classify.5.weight": (20, 40), "sex_classifier.classify.5.bias": (20,),
    "sex_classifier.classify.7.weight": (2, 20), "sex_classifier.classify.7.bias": (2,),
}
for _bn, _c in (("norm", 20), ("classify.1", 40), ("classify.6", 20)):
    for _s in ("weight", "bias", "running_mean", "running_var"):
        FCAE_SHAPES["sex_classifier.%s.%s" % (_bn, _s)] = (_c,)
    FCAE_SHAPES["sex_classifier.%s.num_batches_tracked" % _bn] = ()


def _rng(seed, name):
    return np.random.Generator(np.random.PCG64([seed, zlib.crc32(name.encode())]))


def make_weights(shapes, seed):
    """Random but well-conditioned weights for every state_dict key."""
    out = {}
    for name, shape in shapes.items():
        g = _rng(seed, name)
        if name.endswith("num_batches_tracked"):
            out[name] = np.zeros((), dtype=np.int64)
        elif name.endswith("running_mean"):
            out[name] = (0.2 * g.standard_normal(shape)).astype(np.float32)
        elif name.endswith("running_var"):
            out[name] = g.uniform(0.5, 1.5, shape).astype(np.float32)
        elif name.endswith("weight") and len(shape) >= 2:
            bound = 1.0 / np.sqrt(float(np.prod(shape[1:])))
            out[name] = g.uniform(-bound, bound, shape).astype(np.float32)
        elif name.endswith("weight"):
            out[name] = g.uniform(0.5, 1.5, shape).astype(np.float32)
        else:
            out[name] = g.uniform(-0.1, 0.1, shape).astype(np.float32)
    return out


def make_wav(seed, B, n_samples, ragged=False, scale=0.1):
    """Synthetic PCM like SURVEY 8(d): scale*N(0,1); ragged -> zero padded tails,
    relative lengths linspace(0.6, 1.0, B)."""
    g = _rng(seed, "wav")
    wav = (scale * g.standard_normal((B, n_samples))).astype(np.float32)
    lens = np.ones(B, dtype=np.float32)
    if ragged:
        lens = np.linspace(0.6, 1.0, B).astype(np.float32)
        for b in range(B):
            wav[b, int(round(float(lens[b]) * n_samples)):] = 0.0
    return wav, lens


def make_feats(seed, B, T, F=80):
    """Normalised-feature-like input for the anonymizers (zero mean, unit-ish std)."""
    g = _rng(seed, "feats")
    return g.standard_normal((B, T, F)).astype(np.float32)


def make_noise(seed, B, C):
    """A fixed stand-in for StatisticsPooling's Gaussian noise (pooling.py:313-326):
    same range eps*[1, 9] as the reference's rescaled randn."""
    g = _rng(seed, "noise")
    u = g.uniform(0.0, 1.0, (B, C))
    return (1e-5 * (9.0 - 8.0 * u)).astype(np.float32)
