"""Minimal stand-in for faiss-cpu 1.7.4 (TEST INFRASTRUCTURE, see oracle/__init__.py).

The reference imports ``faiss`` at module top (search/li/LearnedIndex.py:6,
search/li/clustering/faiss_kmeans.py:5) and calls exactly:

* ``faiss.knn(xq, xb, k, metric=faiss.METRIC_INNER_PRODUCT)``
  (search/li/LearnedIndex.py:360-365) -- hot path;
* ``faiss.Kmeans(d, k, verbose=False, seed=2023).train(x)`` and
  ``.index.search(x, 1)`` (search/li/clustering/faiss_kmeans.py:14-22) -- build only.

faiss is a pinned third-party dependency (requirements.txt:8, ``faiss-cpu==1.7.4``)
that is neither vendored under /root/reference nor installable here (no network).
This module restates the published behaviour of those entry points:

``knn``: both inputs are cast to contiguous float32; similarities are the exact
f32 inner products ``xq @ xb.T`` (BLAS sgemm, here through torch so the CPU
baseline is not sandbagged); per query the ``k`` largest are returned sorted in
descending order as ``(D float32[nq,k], I int64[nq,k])``; when ``nb < k`` the
trailing slots hold ``I = -1`` and ``D = -FLT_MAX`` (the neutral element of
faiss's min-heap for inner product).  ``assert d == d2`` as faiss does.
"""
import numpy as np
import torch

METRIC_INNER_PRODUCT = 0
METRIC_L2 = 1

_FLT_MAX = np.finfo(np.float32).max


def knn(xq, xb, k, metric=METRIC_L2):
    xq = np.ascontiguousarray(xq, dtype=np.float32)
    xb = np.ascontiguousarray(xb, dtype=np.float32)
    nq, d = xq.shape
    nb, d2 = xb.shape
    assert d == d2
    tq = torch.from_numpy(xq)
    tb = torch.from_numpy(xb)
    if metric == METRIC_INNER_PRODUCT:
        score = torch.mm(tq, tb.t())
        largest = True
        pad = -_FLT_MAX
    elif metric == METRIC_L2:
        score = torch.cdist(tq, tb, p=2.0) ** 2
        largest = False
        pad = _FLT_MAX
    else:
        raise NotImplementedError(f"metric {metric}")
    kk = min(k, nb)
    D = np.full((nq, k), pad, dtype=np.float32)
    I = np.full((nq, k), -1, dtype=np.int64)
    if kk > 0:
        d_top, i_top = torch.topk(score, kk, dim=1, largest=largest, sorted=True)
        D[:, :kk] = d_top.numpy()
        I[:, :kk] = i_top.numpy()
    return D, I


class _FlatIndex:
    def __init__(self, centroids):
        self._c = centroids

    def search(self, x, k):
        return knn(x, self._c, k, metric=METRIC_L2)


class Kmeans:
    """Plain Lloyd k-means; only what faiss_kmeans.py:14-22 touches."""

    def __init__(self, d, k, niter=20, verbose=False, seed=1234, **_):
        self.d, self.k, self.niter, self.seed = d, k, niter, seed
        self.centroids = None
        self.index = None

    def train(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        rng = np.random.RandomState(self.seed)
        c = x[rng.choice(x.shape[0], self.k, replace=False)].copy()
        tx = torch.from_numpy(x)
        for _ in range(self.niter):
            tc = torch.from_numpy(c)
            lab = torch.cdist(tx, tc).argmin(dim=1)
            for j in range(self.k):
                m = lab == j
                if m.any():
                    c[j] = tx[m].mean(dim=0).numpy()
        self.centroids = c
        self.index = _FlatIndex(c)
