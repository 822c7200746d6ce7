"""ORACLE (test infrastructure, never on the product path).

numpy float64 restatement of the reference's loss statistics:

  CosineSimilarityLoss.forward   utils/cosine_similarity_loss.py:45-56
  cosine_distance_2d             utils/ClusterMI.py:9-10
  _pairwise_dists                utils/ClusterMI.py:12-63
  ClusterMI._mutual_information  utils/ClusterMI.py:88-121
  GroupSampler / GroupSamplingMI utils/GroupSamplingMI.py:14-61

torch.nn.CosineSimilarity / F.cosine_similarity (PyTorch, not vendored) are
restated from their documented definition in torch >= 1.12:
    cos = <x1,x2> / (max(||x1||, eps) * max(||x2||, eps))
(torch 1.10, which the reference's logs pin, used max(||x1||*||x2||, eps); the
two agree unless a norm is below eps).  "parity unpinned" by reference tests
(there are none for utils/); pinned by golden vectors produced by executing the
reference code (tests/golden/make_golden.py).
"""
import numpy as np
from scipy.special import digamma


def cosine_similarity(x1, x2, axis, eps):
    x1 = np.asarray(x1, dtype=np.float64)
    x2 = np.asarray(x2, dtype=np.float64)
    dot = (x1 * x2).sum(axis=axis)
    n1 = np.maximum(np.sqrt((x1 * x1).sum(axis=axis)), eps)
    n2 = np.maximum(np.sqrt((x2 * x2).sum(axis=axis)), eps)
    return dot / (n1 * n2)


def cosine_similarity_loss(x1, x2, eps=1e-6):
    """utils/cosine_similarity_loss.py:53-56: sum(1 - cos(dim=2)) / shape[1]."""
    sim = cosine_similarity(x1, x2, axis=2, eps=eps)
    loss = 1.0 - sim
    return loss.sum() / loss.shape[1]


def pairwise_cosine_dists(X):
    """utils/ClusterMI.py:12-63 with dist_fn = 1 - cosine_similarity(dim=1), eps=1e-8.
    The roll construction just fills the symmetric matrix; diagonal stays 0."""
    X = np.asarray(X, dtype=np.float64)
    N = X.shape[0]
    D = np.zeros((N, N))
    for i in range(N):
        for j in range(N):
            if i != j:
                D[i, j] = 1.0 - cosine_similarity(X[i], X[j], axis=0, eps=1e-8)
    return D


def knn_mi_counts(D, y, k=3):
    """utils/ClusterMI.py:104-115 -> integer m_i per sample (bit-exact target)."""
    y = np.asarray(y)
    N = D.shape[0]
    same = (y[:, None] == y[None, :])
    d_same = np.where(same, D, 10e6)
    # topk(k+1, smallest): the diagonal 0 is one of them; anchor = (k+1)-th smallest
    anchor = np.sort(d_same, axis=1)[:, k]
    m = (D <= anchor[:, None]).sum(axis=1) - 1
    return m, anchor


def cluster_mi(X, y, n_classes=2, k=3, dists=None):
    """utils/ClusterMI.py:88-121 -> MI in bits."""
    y = np.asarray(y)
    N = len(y)
    D = pairwise_cosine_dists(X) if dists is None else dists
    N_x = np.array([(y == i).sum() for i in range(n_classes)], dtype=np.float64)
    avg_N_x = ((N_x / N) * digamma(N_x)).sum()
    m, _ = knn_mi_counts(D, y, k)
    avg_m = digamma(m.astype(np.float64)).mean()
    mi = digamma(float(N)) - avg_N_x + digamma(float(k)) - avg_m
    return mi / np.log(2.0)
