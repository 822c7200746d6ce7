"""ORACLE (test infrastructure, never on the product path).

Functional fp32/fp64 CPU restatement (torch.nn.functional on explicit weight
dicts keyed like the reference's ``state_dict()``) of the two anonymizers:

  ConvAutoencoder.forward          models/ConvAutoEncoder.py:178-200 (layers :141-172)
  TDNNSexClassifier.forward        models/ConvAutoEncoder.py:57-69   (layers :33-55)
  GLU (= x*sigmoid(x))             models/ConvAutoEncoder.py:113-120
  ConvReconstruction.forward       models/EndToEnd.py:65-78           (layers :40-54)
  FullyConnectedAutoencoder        models/FullyConnected.py:127-159
  FullyConnSexClassifier.forward   models/FullyConnected.py:90-104   (layers :68-86)
  StatisticsPooling.forward        speechbrain/speechbrain/nnet/pooling.py:263-312

"parity unpinned" by reference TESTS (the research repo has none for these
classes); pinned instead by executing the reference's own code:
golden vectors in tests/golden/ (made by tests/golden/make_golden.py) and the
live cross-check tests/test_oracle_vs_reference.py.

The conv/linear arithmetic itself is PyTorch's (not vendored in /root/reference,
reference pinned torch 1.10, container has 2.11); it is used here through
torch.nn.functional, i.e. the same library the reference calls.

All BatchNorm layers are evaluated in eval mode (running statistics) unless
``bn_training=True``.  The Gaussian noise StatisticsPooling adds to the mean
(pooling.py:295-298, unseeded) is an explicit argument so both sides of a parity
test can use the same tensor.
"""
import torch
import torch.nn.functional as F


def _t(w, dtype):
    return {k: torch.as_tensor(v).to(dtype) for k, v in w.items()}


def silu_glu(x):
    return x * torch.sigmoid(x)


def instance_norm(x, w, prefix, eps=1e-5):
    """nn.InstanceNorm1d(affine=True), no running stats: biased var over L."""
    return F.instance_norm(x, weight=w[prefix + ".weight"], bias=w[prefix + ".bias"], eps=eps)


def batch_norm(x, w, prefix, training=False, eps=1e-5):
    return F.batch_norm(x, w[prefix + ".running_mean"].clone(), w[prefix + ".running_var"].clone(),
                        w[prefix + ".weight"], w[prefix + ".bias"], training=training, eps=eps)


def statistics_pooling(x, noise=None, eps=1e-5):
    """pooling.py:263-312 with lengths=None.  x [B, L, C] -> [B, 2C]."""
    mean = x.mean(dim=1)
    std = x.std(dim=1)
    if noise is not None:
        mean = mean + noise
    std = std + eps
    return torch.cat((mean, std), dim=1)


def convae_encoder(x, w):
    """models/ConvAutoEncoder.py:141-157.  x [B,1,L] -> [B,128,L/4]."""
    x = silu_glu(F.conv1d(x, w["encoder.0.weight"], w["encoder.0.bias"], stride=1, padding=7))
    x = F.conv1d(x, w["encoder.2.weight"], w["encoder.2.bias"], stride=2, padding=2)
    x = silu_glu(instance_norm(x, w, "encoder.3"))
    x = F.conv1d(x, w["encoder.5.weight"], w["encoder.5.bias"], stride=1, padding=2)
    x = silu_glu(instance_norm(x, w, "encoder.6"))
    x = F.conv1d(x, w["encoder.8.weight"], w["encoder.8.bias"], stride=2, padding=2)
    x = silu_glu(instance_norm(x, w, "encoder.9"))
    x = F.conv1d(x, w["encoder.11.weight"], w["encoder.11.bias"], stride=1, padding=2)
    x = silu_glu(instance_norm(x, w, "encoder.12"))
    return x


def convae_decoder(x, w):
    """models/ConvAutoEncoder.py:160-172.  [B,128,L/4] -> [B,1,L]."""
    x = F.conv1d(x, w["decoder.0.weight"], w["decoder.0.bias"], stride=1, padding=2)
    x = F.conv_transpose1d(x, w["decoder.1.weight"], w["decoder.1.bias"], stride=2, padding=2,
                           output_padding=1)
    x = silu_glu(instance_norm(x, w, "decoder.2"))
    x = F.conv1d(x, w["decoder.4.weight"], w["decoder.4.bias"], stride=1, padding=2)
    x = F.conv_transpose1d(x, w["decoder.5.weight"], w["decoder.5.bias"], stride=2, padding=2,
                           output_padding=1)
    x = silu_glu(instance_norm(x, w, "decoder.6"))
    x = F.conv1d(x, w["decoder.8.weight"], w["decoder.8.bias"], stride=1, padding=7)
    return x


def tdnn_sex_classifier(enc, w, noise=None, bn_training=False, p="sex_classifier."):
    """models/ConvAutoEncoder.py:57-69.  enc [B,128,L4] -> log-probs [B,2].
    GradReverse is the identity in the forward direction (:12-28)."""
    x = batch_norm(enc, w, p + "norm", bn_training)
    for i, dil in ((0, 1), (3, 2), (6, 3)):
        x = F.conv1d(x, w[p + "tdnn.%d.weight" % i], w[p + "tdnn.%d.bias" % i], dilation=dil)
        x = F.relu(x)
        x = batch_norm(x, w, p + "tdnn.%d" % (i + 2), bn_training)
    # :61 -- a memory REINTERPRETATION [B,C,L] -> [B,L,C], not a transpose
    x = x.reshape(x.shape[0], x.shape[2], x.shape[1])
    s = statistics_pooling(x, noise)                                     # [B,256]
    s = F.relu(F.linear(s, w[p + "classify.0.weight"], w[p + "classify.0.bias"]))
    s = batch_norm(s, w, p + "classify.2", bn_training)
    s = F.relu(F.linear(s, w[p + "classify.3.weight"], w[p + "classify.3.bias"]))
    s = batch_norm(s, w, p + "classify.5", bn_training)
    s = F.linear(s, w[p + "classify.6.weight"], w[p + "classify.6.bias"])
    return F.log_softmax(s, 1)


def convae_forward(feats, weights, noise=None, dtype=torch.float32, bn_training=False):
    """models/ConvAutoEncoder.py:178-200.  feats [B,T,F] (T*F % 4 == 0) ->
    (recon [B,T,F], sex log-probs [B,2])."""
    w = _t(weights, dtype)
    feats = torch.as_tensor(feats).to(dtype)
    B, T, Fd = feats.shape
    x = feats.reshape(B, T * Fd).unsqueeze(1)        # :181,188 -- ONE flat 1-D signal
    enc = convae_encoder(x, w)
    if noise is not None:
        noise = torch.as_tensor(noise).to(dtype)
    logp = tdnn_sex_classifier(enc, w, noise, bn_training)
    rec = convae_decoder(enc, w).squeeze(1).reshape(B, T, Fd)
    return rec, logp


def convrecon_forward(feats, weights, dtype=torch.float32):
    """models/EndToEnd.py:65-78 (layers :40-54): the "endtoend" anonymizer without its external pretrained
    classifier.  feats [B,T,F] (T*F even) -> recon [B,T,F]."""
    w = _t(weights, dtype)
    feats = torch.as_tensor(feats).to(dtype)
    B, T, Fd = feats.shape
    x = feats.reshape(B, T * Fd).unsqueeze(1)
    x = F.conv1d(x, w["encoder.0.weight"], w["encoder.0.bias"], stride=1, padding=7)
    x = silu_glu(instance_norm(x, w, "encoder.1"))
    x = F.conv1d(x, w["encoder.3.weight"], w["encoder.3.bias"], stride=2, padding=2)
    x = silu_glu(instance_norm(x, w, "encoder.4"))
    x = F.conv1d(x, w["encoder.6.weight"], w["encoder.6.bias"], stride=1, padding=2)
    x = silu_glu(instance_norm(x, w, "encoder.7"))
    x = F.conv_transpose1d(x, w["encoder.9.weight"], w["encoder.9.bias"], stride=2, padding=2, output_padding=1)
    x = silu_glu(instance_norm(x, w, "encoder.10"))
    x = F.conv1d(x, w["encoder.12.weight"], w["encoder.12.bias"], stride=1, padding=7)
    return x.squeeze(1).reshape(B, T, Fd)


def fc_sex_classifier(z, w, noise=None, bn_training=False, p="sex_classifier."):
    """models/FullyConnected.py:90-104.  z [B,T,20] -> log-probs [B,2]."""
    B, T, C = z.shape
    x = z.reshape(B, C, T)                            # :93 reinterpretation
    x = batch_norm(x, w, p + "norm", bn_training)
    x = x.reshape(B, T, C)                            # :95 reinterpretation back
    x = F.relu(F.linear(x, w[p + "initial.0.weight"], w[p + "initial.0.bias"]))
    x = F.relu(F.linear(x, w[p + "initial.2.weight"], w[p + "initial.2.bias"]))
    s = statistics_pooling(x, noise)                  # [B,80]
    s = F.linear(s, w[p + "classify.0.weight"], w[p + "classify.0.bias"])
    s = F.relu(batch_norm(s, w, p + "classify.1", bn_training))
    s = F.relu(F.linear(s, w[p + "classify.3.weight"], w[p + "classify.3.bias"]))
    s = F.linear(s, w[p + "classify.5.weight"], w[p + "classify.5.bias"])
    s = batch_norm(s, w, p + "classify.6", bn_training)
    s = F.linear(s, w[p + "classify.7.weight"], w[p + "classify.7.bias"])
    return F.log_softmax(s, 1)


def fcae_forward(feats, weights, noise=None, dtype=torch.float32, bn_training=False):
    """models/FullyConnected.py:147-159.  feats [B,T,80] -> (recon [B,T,80], [B,2])."""
    w = _t(weights, dtype)
    x = torch.as_tensor(feats).to(dtype)
    x = F.relu(F.linear(x, w["encoder.0.weight"], w["encoder.0.bias"]))
    x = F.relu(F.linear(x, w["encoder.2.weight"], w["encoder.2.bias"]))
    z = F.linear(x, w["encoder.4.weight"], w["encoder.4.bias"])
    if noise is not None:
        noise = torch.as_tensor(noise).to(dtype)
    logp = fc_sex_classifier(z, w, noise, bn_training)
    x = F.relu(F.linear(z, w["decoder.0.weight"], w["decoder.0.bias"]))
    x = F.relu(F.linear(x, w["decoder.2.weight"], w["decoder.2.bias"]))
    rec = F.linear(x, w["decoder.4.weight"], w["decoder.4.bias"])
    return rec, logp
