"""ORACLE (test infrastructure, never on the product path).

Functional CPU restatement (torch.nn.functional on explicit weight dicts keyed like the reference's ``state_dict()``)
of the external x-vector gender classifier the evaluation configs run on (anonymized) features:

  Xvector.forward                        speechbrain/speechbrain/lobes/models/Xvector.py:95-108  (layers :60-93)
  Conv1d(padding="same", reflect)        speechbrain/speechbrain/nnet/CNN.py:349-352,400-403 (+ get_padding_elem)
  BatchNorm1d on [B,T,C]                 speechbrain/speechbrain/nnet/normalization.py:71-99
  StatisticsPooling.forward (lengths)    speechbrain/speechbrain/nnet/pooling.py:263-312
  Classifier                             speechbrain/speechbrain/lobes/models/Xvector.py:111-172
  EncoderClassifier.classify_batch_feats speechbrain/speechbrain/pretrained/interfaces.py:762-776,825-829

in the configuration of speechbrain_configs/evaluator_inference.yaml:34-48.  The reference has no tests for this
classifier in this configuration ("parity unpinned" by tests); pinned by executing the reference's own classes:
tests/golden/golden_v2.npz (tests/golden/make_golden_v2.py) and tests/test_oracle_vs_reference.py.
"""
import torch
import torch.nn.functional as F

KERNELS = (5, 3, 3, 1, 1)
DILATIONS = (1, 2, 3, 1, 1)


def _bn(x_btc, w, prefix, eps=1e-5):
    """sb BatchNorm1d in eval mode: per-channel affine with the running statistics."""
    s = w[prefix + ".weight"] / torch.sqrt(w[prefix + ".running_var"] + eps)
    return (x_btc - w[prefix + ".running_mean"]) * s + w[prefix + ".bias"]


def xvector_embedding(feats, lens, w, noise=None, eps=1e-5):
    """feats [B,T,80], lens [B] relative (or None) -> x-vectors [B,1,128]."""
    x = feats
    B, T, _ = x.shape
    for i, (k, d) in enumerate(zip(KERNELS, DILATIONS)):
        y = x.transpose(1, 2)                                   # sb layout [B,T,C] -> [B,C,T]
        pad = d * (k - 1) // 2                                  # get_padding_elem: stride 1 => dilation*(k-1)/2 each side
        if pad > 0:
            y = F.pad(y, (pad, pad), mode="reflect")
        y = F.conv1d(y, w["blocks.%d.conv.weight" % (3 * i)], w["blocks.%d.conv.bias" % (3 * i)], dilation=d)
        x = _bn(F.leaky_relu(y.transpose(1, 2), 0.01), w, "blocks.%d.norm" % (3 * i + 2))
    means, stds = [], []
    for b in range(B):
        n = T if lens is None else int(torch.round(lens[b] * T))
        means.append(x[b, :n].mean(dim=0))
        stds.append(x[b, :n].std(dim=0))
    mean, std = torch.stack(means), torch.stack(stds)
    if noise is not None:
        mean = mean + noise
    pooled = torch.cat((mean, std + eps), dim=1).unsqueeze(1)   # [B,1,3000]
    return F.linear(pooled, w["blocks.16.w.weight"], w["blocks.16.w.bias"])


def xvector_classifier(emb, w):
    """Classifier: act -> norm -> DNN.block_0 (linear, act, norm) -> out -> log-softmax.  emb [B,1,128] -> [B,1,2]."""
    x = _bn(F.leaky_relu(emb, 0.01), w, "norm.norm")
    x = F.linear(x, w["DNN.block_0.linear.w.weight"], w["DNN.block_0.linear.w.bias"])
    x = _bn(F.leaky_relu(x, 0.01), w, "DNN.block_0.norm.norm")
    x = F.linear(x, w["out.w.weight"], w["out.w.bias"])
    return F.log_softmax(x, dim=-1)


def classify_batch_feats(feats, lens, w_emb, w_clf, noise=None, dtype=torch.float32):
    """-> (embeddings [B,1,128], log-probs [B,2])."""
    we = {k: torch.as_tensor(v).to(dtype) for k, v in w_emb.items() if not k.endswith("num_batches_tracked")}
    wc = {k: torch.as_tensor(v).to(dtype) for k, v in w_clf.items() if not k.endswith("num_batches_tracked")}
    feats = torch.as_tensor(feats).to(dtype)
    lens = None if lens is None else torch.as_tensor(lens).to(torch.float32)
    noise = None if noise is None else torch.as_tensor(noise).to(dtype)
    emb = xvector_embedding(feats, lens, we, noise)
    return emb, xvector_classifier(emb, wc).squeeze(1)
