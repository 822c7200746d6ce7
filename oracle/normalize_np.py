"""ORACLE (test infrastructure, never on the product path).

numpy restatement of the reference's (fork-modified) InputNormalization with
norm_type in {"global", "batch", "sentence"} and of the x36 time padding:

  InputNormalization.forward        speechbrain/speechbrain/processing/features.py:998-1121
  InputNormalization._compute_current_stats                                   :1123-1148
  time padding to a multiple of 36  speechbrain_convae_train.py:62-63 (dup :86-87)

Pinned by the reference's own known-answer test
(speechbrain/tests/unittests/test_features.py:100-119: [1,2,3,0,0,0], len 0.5 ->
[-1,0,1,-2,-2,-2]) and by golden vectors generated from the reference.
"""
import numpy as np


def valid_frames(rel_len, T):
    """features.py:1021  torch.round(lengths[i] * T).int()  -- fp32 product, round
    half to even (torch.round), then truncation to int32."""
    prod = np.float32(rel_len) * np.float32(T)
    return int(np.rint(prod))


def current_stats(x):
    """features.py:1123-1148: mean, UNBIASED std over the valid frames (dim 0),
    std floored at eps=1e-10.  A single valid frame gives NaN std like torch."""
    x = np.asarray(x)
    mean = x.mean(axis=0)
    with np.errstate(invalid="ignore", divide="ignore"):
        std = x.std(axis=0, ddof=1) if x.shape[0] > 1 else np.full(x.shape[1:], np.nan)
    # torch.max(nan, eps) propagates NaN; np.maximum does too
    std = np.maximum(std, 1e-10)
    return mean, std


class InputNormalizationOracle:
    """State machine equivalent to the reference module (features.py:973-996)."""

    def __init__(self, mean_norm=True, std_norm=True, norm_type="global",
                 avg_factor=None, update_until_epoch=3):
        assert norm_type in ("global", "batch", "sentence")
        self.mean_norm, self.std_norm = mean_norm, std_norm
        self.norm_type = norm_type
        self.avg_factor = avg_factor
        self.update_until_epoch = update_until_epoch
        self.glob_mean = None       # reference: torch.tensor([0]) sentinel
        self.glob_std = None
        self.count = 0
        self.training = True

    def __call__(self, x, lengths, epoch=0):
        x = np.array(x, dtype=np.float64, copy=True)
        B, T = x.shape[0], x.shape[1]
        means, stds = [], []
        for b in range(B):
            n = valid_frames(lengths[b], T)
            m, s = current_stats(x[b, :n])
            if not self.mean_norm:
                m = np.zeros(1)
            if not self.std_norm:
                s = np.ones(1)
            means.append(m)
            stds.append(s)
            if self.norm_type == "sentence":                 # features.py:1031-1033
                x[b] = (x[b] - m) / s
        if self.norm_type in ("batch", "global"):
            cur_mean = np.mean(np.stack(means), axis=0)      # features.py:1082-1083
            cur_std = np.mean(np.stack(stds), axis=0)
            if self.norm_type == "batch":
                x = (x - cur_mean) / cur_std
            else:
                if self.training:                            # features.py:1090-1112
                    if self.count == 0:
                        self.glob_mean, self.glob_std = cur_mean, cur_std
                    elif epoch < self.update_until_epoch:
                        w = 1.0 / (self.count + 1) if self.avg_factor is None else self.avg_factor
                        self.glob_mean = (1 - w) * self.glob_mean + w * cur_mean
                        self.glob_std = (1 - w) * self.glob_std + w * cur_std
                    self.count += 1
                if self.glob_mean is None:                   # features.py:1114-1117
                    self.glob_mean = cur_mean
                if self.glob_std is None:
                    self.glob_std = cur_std
                x = (x - self.glob_mean) / self.glob_std     # features.py:1119
        return x


def padded_frames(T, multiple=36):
    """speechbrain_convae_train.py:62  T -> next multiple of 36 (no-op if already)."""
    return T if T % multiple == 0 else T + (multiple - T % multiple)


def pad_time(feats, multiple=36):
    """speechbrain_convae_train.py:62-63  zero-pad [B,T,F] along T (after normalisation)."""
    B, T, F = feats.shape
    Tp = padded_frames(T, multiple)
    out = np.zeros((B, Tp, F), dtype=feats.dtype)
    out[:, :T] = feats
    return out
