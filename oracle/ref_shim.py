"""TEST INFRASTRUCTURE ONLY -- imports the *unmodified* reference from /root/reference.

This module exists so that (a) ``tests/golden/make_golden.py`` can generate the
committed golden vectors from the reference's own code and (b) the CPU test-suite
can cross-check the oracle restatement (``oracle/*.py``) against the real
reference whenever ``/root/reference`` is mounted (it is NOT on the GPU box, so
nothing under ``-m gpu``, ``smoke()`` or ``bench.py`` may import this file).

The reference cannot be imported the normal way here: ``import speechbrain``
executes ``speechbrain/__init__.py`` -> ``core.py:32`` which needs ``hyperpyyaml``
(not installed, no network).  SURVEY.md section 8(c) describes the work-around used
below: register bare namespace modules whose ``__path__`` points into the
reference tree so that only the leaf modules on the hot path are executed.

Nothing from the reference is copied: the modules are executed where they lie.
"""
import os
import sys
import types
import importlib

REF_ROOT = os.environ.get("SPANON_REFERENCE_ROOT", "/root/reference")
_SB = os.path.join(REF_ROOT, "speechbrain", "speechbrain")


def available() -> bool:
    return os.path.isdir(_SB) and os.path.isdir(os.path.join(REF_ROOT, "models"))


def _ns(name, path):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    m.__path__ = [path]
    m.__package__ = name
    sys.modules[name] = m
    return m


_loaded = {}


def load():
    """Returns a dict of the reference classes on the hot path."""
    if _loaded:
        return _loaded
    if not available():
        raise RuntimeError("reference tree not mounted at %s" % REF_ROOT)
    sys.dont_write_bytecode = True  # the reference tree is read-only
    if "speechbrain" in sys.modules and not getattr(sys.modules["speechbrain"], "_spanon_shim", False):
        raise RuntimeError("a real 'speechbrain' package is already imported")
    sb = _ns("speechbrain", _SB)
    sb._spanon_shim = True
    for sub in ("utils", "lobes", "nnet", "processing", "dataio"):
        setattr(sb, sub, _ns("speechbrain." + sub, os.path.join(_SB, sub)))
    # models/ConvAutoEncoder.py:6 imports EncoderClassifier only for the dead
    # CycleGAN variants; a stub keeps the import side-effect free.
    pre = types.ModuleType("speechbrain.pretrained")

    class _EncoderClassifierStub:
        """models/EndToEnd.py:57-61 loads a pretrained classifier from an absolute path of the authors' machine at
        construction time; the stub makes the class constructible (its classifier is never called by the shim)."""

        @classmethod
        def from_hparams(cls, *a, **k):
            return None

    pre.EncoderClassifier = _EncoderClassifierStub
    sys.modules["speechbrain.pretrained"] = pre
    sb.pretrained = pre

    feats_lobe = importlib.import_module("speechbrain.lobes.features")
    feats_proc = importlib.import_module("speechbrain.processing.features")
    pooling = importlib.import_module("speechbrain.nnet.pooling")

    # the research repo's own top-level packages: models/, utils/
    ref_models = _ns("spanon_ref_models", os.path.join(REF_ROOT, "models"))
    ref_utils = _ns("spanon_ref_utils", os.path.join(REF_ROOT, "utils"))
    convae = importlib.import_module("spanon_ref_models.ConvAutoEncoder")
    fc = importlib.import_module("spanon_ref_models.FullyConnected")
    e2e = importlib.import_module("spanon_ref_models.EndToEnd")
    cos = importlib.import_module("spanon_ref_utils.cosine_similarity_loss")
    cmi = importlib.import_module("spanon_ref_utils.ClusterMI")
    gmi = importlib.import_module("spanon_ref_utils.GroupSamplingMI")
    mil = importlib.import_module("spanon_ref_utils.mi_loss")

    # Xvector.py reaches its layers through attribute access on the package (sb.nnet.containers.Sequential, ...):
    # wire the leaf modules onto the namespace stubs
    for sub in ("containers", "linear", "normalization", "activations", "CNN", "pooling"):
        setattr(sb.nnet, sub, importlib.import_module("speechbrain.nnet." + sub))
    xvec = importlib.import_module("speechbrain.lobes.models.Xvector")

    _loaded.update(
        Xvector=xvec.Xvector,
        XvectorClassifier=xvec.Classifier,
        Fbank=feats_lobe.Fbank,
        STFT=feats_proc.STFT,
        spectral_magnitude=feats_proc.spectral_magnitude,
        Filterbank=feats_proc.Filterbank,
        InputNormalization=feats_proc.InputNormalization,
        StatisticsPooling=pooling.StatisticsPooling,
        ConvAutoencoder=convae.ConvAutoencoder,
        TDNNSexClassifier=convae.TDNNSexClassifier,
        ConvReconstruction=e2e.ConvReconstruction,
        FullyConnectedAutoencoder=fc.FullyConnectedAutoencoder,
        FullyConnSexClassifier=fc.FullyConnSexClassifier,
        CosineSimilarityLoss=cos.CosineSimilarityLoss,
        ClusterMI=cmi.ClusterMI,
        GroupSamplingMI=gmi.GroupSamplingMI,
        MILoss=mil.MILoss,
        pairwise_dists=cmi._pairwise_dists,
        cosine_distance_2d=cmi.cosine_distance_2d,
    )
    return _loaded
