"""Host-side mirror of the reference's loss utilities, backed by libspanon's reduction kernels.

  CosineSimilarityLoss     utils/cosine_similarity_loss.py:5-56
  cosine_distance_2d       utils/ClusterMI.py:9-10
  ClusterMI                utils/ClusterMI.py:65-121
  GroupSampler / GroupSamplingMI   utils/GroupSamplingMI.py:14-61
  MILoss                   utils/mi_loss.py:8-16
  SimilarityMetricsStats   utils/utility_similarity_aggregator.py:4-52
"""
import ctypes as C

import numpy as np
import torch
import torch.nn as nn
from torch.nn.modules.loss import _Loss

from . import _lib


class _CosineLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x1, x2, eps):
        B, T, D = x1.shape
        loss = torch.empty((1,), dtype=torch.float32, device=x1.device)
        with torch.cuda.device(x1.device):
            _lib.check(_lib.lib().spanon_cosine_loss(_lib.ptr(x1), _lib.ptr(x2), B, T, D, eps, _lib.ptr(loss), None,
                                                     _lib.stream_ptr(x1.device)), "cosine_loss")
        ctx.save_for_backward(x1, x2)
        ctx.eps = eps
        return loss[0]

    @staticmethod
    def backward(ctx, grad_out):
        x1, x2 = ctx.saved_tensors
        B, T, D = x1.shape
        g1 = torch.empty_like(x1) if ctx.needs_input_grad[0] else None
        g2 = torch.empty_like(x2) if ctx.needs_input_grad[1] else None
        go = grad_out.reshape(1).to(torch.float32).contiguous()
        with torch.cuda.device(x1.device):
            _lib.check(_lib.lib().spanon_cosine_loss_bwd(_lib.ptr(x1), _lib.ptr(x2), _lib.ptr(go), B, T, D, ctx.eps,
                                                         _lib.ptr(g1), _lib.ptr(g2), _lib.stream_ptr(x1.device)),
                       "cosine_loss_bwd")
        return g1, g2, None


class CosineSimilarityLoss(_Loss):
    """utils/cosine_similarity_loss.py:45-56: sum(1 - cos(x1, x2, dim=2, eps=1e-6)) / x1.shape[1]."""
    __constants__ = ['margin', 'reduction']
    eps = 1e-6

    def __init__(self, size_average=None, reduce=None, reduction: str = 'mean') -> None:
        super().__init__(size_average, reduce, reduction)

    def forward(self, input1, input2):
        _lib.require_cuda(input1, "input1")
        _lib.require_cuda(input2, "input2")
        if input1.dim() != 3 or input1.shape != input2.shape:
            raise RuntimeError("CosineSimilarityLoss expects two [batch, time, dim] tensors of equal shape "
                               "(CosineSimilarity(dim=2)); got %s and %s" % (tuple(input1.shape), tuple(input2.shape)))
        return _CosineLossFn.apply(input1.contiguous(), input2.contiguous(), self.eps)


def cosine_distance_2d(x, y):
    """utils/ClusterMI.py:9-10 -- the only metric the MI kernel implements (used as a marker)."""
    return 1 - torch.nn.functional.cosine_similarity(x, y, dim=1)


def knn_mi(X, y, idx=None, n_classes=2, k=3, return_dists=False):
    """One launch for all index sets: MI [n_sets] (bits), neighbour counts [n_sets, N], distances."""
    _lib.require_cuda(X, "X")
    if X.dim() != 2:
        raise ValueError("X must be [N, D]")
    X = X.contiguous()
    y = torch.as_tensor(y).to(device=X.device, dtype=torch.int32).contiguous()
    if idx is None:
        n_sets, N, idx_t = 1, X.shape[0], None
    else:
        idx_t = torch.as_tensor(idx).to(device=X.device, dtype=torch.int32).contiguous()
        n_sets, N = idx_t.shape
    mi = torch.empty((n_sets,), dtype=torch.float32, device=X.device)
    m_i = torch.empty((n_sets, N), dtype=torch.int32, device=X.device)
    dists = torch.empty((n_sets, N, N), dtype=torch.float32, device=X.device) if return_dists else None
    with torch.cuda.device(X.device):
        _lib.check(_lib.lib().spanon_knn_mi_stats(_lib.ptr(X), _lib.ptr(y), _lib.ptr(idx_t), n_sets, N, X.shape[1],
                                                  n_classes, k, _lib.ptr(dists), _lib.ptr(m_i), _lib.ptr(mi),
                                                  _lib.stream_ptr(X.device)), "knn_mi_stats")
    return mi, m_i, dists


class ClusterMI(nn.Module):
    """utils/ClusterMI.py:65-121 (Ross 2014 kNN estimator of MI between a discrete and a continuous set)."""

    def __init__(self, n_classes=2, k=3, dist_metric=cosine_distance_2d):
        super().__init__()
        assert n_classes >= 2, "Number of classes. Needs to be larger than or equal to two - " + str(n_classes) + " given."
        if dist_metric is not cosine_distance_2d:
            raise NotImplementedError("only cosine_distance_2d (the reference default) has a kernel")
        self.k = k
        self.k_digamma = torch.digamma(torch.tensor(k).float())
        self.n_classes = n_classes
        self.distance = dist_metric

    def forward(self, X, y):
        return self._mutual_information(X, y)

    def _mutual_information(self, X, y):
        mi, _, _ = knn_mi(X, y, None, self.n_classes, self.k)
        return mi[0]


class GroupSampler(object):
    """utils/GroupSamplingMI.py:14-40: same numpy RNG call sequence (one np.random.choice per group per
    draw, groups in sorted-key order like pandas groupby), without pandas."""

    def __init__(self, groups, sample_size):
        self.groups = groups
        self.sample_size = sample_size
        self.create_groups()

    def sample_groups(self):
        samples = []
        for g in self.group_samples.keys():
            samples_ = np.random.choice(self.samples_per_group[g], self.sample_size, replace=False)
            samples.extend(self.group_samples[g][samples_])
        return samples

    def create_groups(self):
        keys = [g.item() if hasattr(g, "item") else g for g in list(self.groups)]
        order = sorted(set(keys))
        self.group_samples = {key: np.array([i for i, g in enumerate(keys) if g == key]) for key in order}
        self.samples_per_group = {key: len(value) for key, value in self.group_samples.items()}


class GroupSamplingMI(ClusterMI):
    """utils/GroupSamplingMI.py:42-61; the n_iterations index sets are evaluated by ONE kernel launch."""

    def __init__(self, n_samples=1, n_classes=2, n_iterations=100, k=3, dist_metric=cosine_distance_2d):
        super().__init__(n_classes, k, dist_metric)
        self.n_samples = n_samples
        self.n_iterations = n_iterations

    def forward(self, X, y, groups):
        group_sampler = GroupSampler(groups, self.n_samples)
        idx = np.stack([np.asarray(group_sampler.sample_groups(), dtype=np.int32)
                        for _ in range(self.n_iterations)])
        mi_t, _, _ = knn_mi(X, y, idx, self.n_classes, self.k)
        mi_cpu = mi_t.cpu()
        mi = list(mi_cpu.unbind(0))
        return mi, torch.mean(mi_cpu), torch.std(mi_cpu)


class MILoss(nn.Module):
    """utils/mi_loss.py:8-16 (returns the list of per-iteration MI values, like the reference)."""

    def forward(self, X, y, batch, batch_size, n_classes=2, samples_set_per_batch=1):
        group_sampling_MI = GroupSamplingMI(n_samples=batch_size // samples_set_per_batch, n_classes=n_classes)
        mi_loss, mi_mean, mi_stdd = group_sampling_MI(X, y, batch)
        return mi_loss


class SimilarityMetricsStats(object):
    """utils/utility_similarity_aggregator.py:4-52 (running mean of cosine scores; plain bookkeeping)."""

    def __init__(self):
        self.clear()

    def clear(self):
        self.scores = []
        self.summary = {}
        self.value = 0
        self.denom = 0

    def append(self, scores):
        self.scores.extend(scores.detach())
        self.value += torch.sum(scores.detach())
        self.denom += scores.shape[0]

    def peek(self):
        return self.value / (1.0 * self.denom)

    def summarize(self):
        if isinstance(self.scores, list):
            self.scores = torch.stack(self.scores)
        self.summary["average"] = torch.sum(self.scores) / self.scores.shape[0]
        return self.summary["average"]
