"""Pandas-free CPU restatement of the LMI query path (TEST INFRASTRUCTURE).

See ``oracle/__init__.py`` for who may import this.  Every function cites the
reference file:line it restates (paths relative to /root/reference).  The
arithmetic goes through the same library calls the reference makes on CPU
(torch ``linear``/``relu``/``softmax``/``topk`` for the MLP; an exact f32
``Q @ X.T`` + per-row top-k for ``faiss.knn``), so on the same machine the
outputs are bit-identical to the reference's except where ``torch.topk`` /
``np.argsort`` are free to order exact ties.

Pinned against the reference executed in-process by
``tests/test_oracle_vs_reference.py`` (live, build container) and
``tests/golden/*.npz`` (committed; produced by ``oracle/make_golden.py``).

What the restatement changes (layout only, never arithmetic): the reference
re-groups the whole DataFrame and copies every bucket out of it once per round
(search/li/LearnedIndex.py:350,357).  ``BucketMajor`` performs that grouping
once, up front, so the oracle scales to 10M rows in RAM.
"""
from __future__ import annotations

import time
from collections import defaultdict
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

EMPTY_VALUE = -1  # search/li/PriorityQueue.py:4
PER_BUCKET_K = 10  # `k` is not forwarded to _search_single_bucket (LearnedIndex.py:111-117, :334)
FLT_MAX = np.finfo(np.float32).max

Layers = List[Tuple[np.ndarray, np.ndarray]]  # [(W f32[out,in], b f32[out]), ...]


# --------------------------------------------------------------------------- #
# model.py
# --------------------------------------------------------------------------- #
def to_f32_tensor(x) -> torch.Tensor:
    """``data_X_to_torch`` -- search/li/model.py:102-105."""
    return torch.from_numpy(np.array(x).astype(np.float32))


def mlp_logits(layers: Layers, x: torch.Tensor) -> torch.Tensor:
    """``Model.forward`` over ``Sequential(Linear, ReLU, ..., Linear)`` --
    search/li/model.py:29-79 (layer zoo), :97-99 (forward)."""
    h = x
    with torch.no_grad():
        for i, (W, b) in enumerate(layers):
            h = torch.nn.functional.linear(
                h, torch.from_numpy(np.asarray(W, np.float32)), torch.from_numpy(np.asarray(b, np.float32))
            )
            if i + 1 < len(layers):
                h = torch.relu(h)
    return h


def predict_proba(layers: Layers, x: torch.Tensor) -> Tuple[np.ndarray, np.ndarray]:
    """``NeuralNetwork.predict_proba`` -- search/li/model.py:226-241: softmax over
    the class axis, then ``topk`` over ALL classes (a full descending sort)."""
    out = mlp_logits(layers, x)
    prob = torch.softmax(out, dim=0 if out.dim() == 1 else 1)
    probs, classes = prob.topk(prob.shape[1])
    return probs.numpy(), classes.numpy()


def predict(layers: Layers, x: torch.Tensor) -> np.ndarray:
    """``NeuralNetwork.predict`` -- search/li/model.py:213-224 (argmax of logits)."""
    return torch.max(mlp_logits(layers, x), 1)[1].numpy()


# --------------------------------------------------------------------------- #
# utils.py
# --------------------------------------------------------------------------- #
def filter_path_idxs(paths: np.ndarray, path: Sequence[int]) -> np.ndarray:
    """search/li/utils.py:61-65 -- rows of ``paths`` equal to ``path`` on every level."""
    return np.flatnonzero((paths == np.asarray(path)).all(axis=1))


# --------------------------------------------------------------------------- #
# PriorityQueue.py
# --------------------------------------------------------------------------- #
class BatchedQueue:
    """``PriorityQueue`` -- search/li/PriorityQueue.py:7-94.  One dense queue per
    query; the best entry sits at the END of the occupied prefix."""

    def __init__(self, n_queries: int, capacity: int, n_levels: int):
        # :18-34
        self.prob = np.full((n_queries, capacity), EMPTY_VALUE, dtype=np.float32)
        self.path = np.full((n_queries, capacity, n_levels), EMPTY_VALUE, dtype=np.int32)
        self.size = np.zeros(n_queries, dtype=np.int32)
        self.dirty = np.zeros(n_queries, dtype=bool)
        self.n_levels = n_levels

    def push(self, rows: np.ndarray, path: np.ndarray, prob: np.ndarray) -> None:
        # :36-50
        at = self.size[rows]
        self.prob[rows, at] = prob
        self.path[rows, at, :] = path
        self.dirty[rows] = True
        self.size[rows] += 1

    def pop(self, rows: np.ndarray) -> np.ndarray:
        # :52-56
        self.size[rows] -= 1
        return self.path[rows, self.size[rows], :]

    def resort(self) -> None:
        # :58-94 -- ascending argsort (numpy default kind, not stable) per distinct length
        for length in np.unique(self.size):
            if length in (0, 1):
                continue
            rows = np.flatnonzero((self.size == length) & self.dirty)
            perm = self.prob[rows, :length].argsort()
            self.prob[rows, :length] = np.take_along_axis(self.prob[rows, :length], perm, axis=1)
            for lvl in range(self.n_levels):
                self.path[rows, :length, lvl] = np.take_along_axis(
                    self.path[rows, :length, lvl], perm, axis=1
                )
            self.dirty[rows] = False


# --------------------------------------------------------------------------- #
# LearnedIndex.py
# --------------------------------------------------------------------------- #
class OracleIndex:
    """``LearnedIndex`` state -- search/li/LearnedIndex.py:23-39, as raw weights."""

    def __init__(
        self,
        root_layers: Layers,
        internal: Optional[Dict[Tuple[int, ...], Layers]] = None,
        bucket_paths: Optional[List[Tuple[int, ...]]] = None,
    ):
        self.root_layers = root_layers
        self.internal = dict(internal or {})
        if bucket_paths is None:
            bucket_paths = [(i,) for i in range(root_layers[-1][0].shape[0])]
        self.bucket_paths = [tuple(p) for p in bucket_paths]


def precompute_bucket_order(
    index: OracleIndex, queries_navigation: np.ndarray, n_buckets: int, n_categories: Sequence[int]
) -> Tuple[np.ndarray, float]:
    """``_precompute_bucket_order`` -- search/li/LearnedIndex.py:163-252 together
    with ``_visit_internal_nodes`` :254-301 and ``_visit_buckets`` :303-325."""
    nq = queries_navigation.shape[0]
    n_levels = len(n_categories)
    t_inf = 0.0

    t0 = time.time()
    root_prob, root_cls = predict_proba(index.root_layers, to_f32_tensor(queries_navigation))
    t_inf += time.time() - t0

    order = np.full((nq, n_buckets, n_levels), EMPTY_VALUE, dtype=np.int32)
    if n_levels == 1:
        # :209-214
        order[:, :n_buckets, 0] = root_cls[:, :n_buckets]
        return order, t_inf

    # :216-227 -- seed the queue with level-1 nodes, least probable first
    queue = BatchedQueue(nq, int(np.prod(n_categories)), n_levels)
    everyone = np.arange(nq)
    for j in reversed(range(n_categories[0])):
        p = np.full((nq, n_levels), EMPTY_VALUE, dtype=np.int32)
        p[:, 0] = root_cls[:, j]
        queue.push(everyone, p, root_prob[:, j])

    filled = np.zeros(nq, dtype=np.int32)
    while not np.all(filled == n_buckets):  # :234
        active = np.flatnonzero(filled < n_buckets)
        visiting = queue.pop(active)

        # _visit_internal_nodes :269-299 -- children are pushed with their RAW
        # (not parent-multiplied) probabilities, most probable child first
        for path, layers in index.internal.items():
            hit = filter_path_idxs(visiting, path)
            rows = active[hit]
            if rows.shape[0] == 0:
                continue
            t0 = time.time()
            prob, cls = predict_proba(layers, to_f32_tensor(queries_navigation[rows]))
            t_inf += time.time() - t0
            depth = len(path) - list(path).count(EMPTY_VALUE)
            for j in range(cls.shape[1]):
                child = np.full((rows.shape[0], n_levels), EMPTY_VALUE, dtype=np.int32)
                child[:] = np.array(path)
                child[:, depth] = cls[:, j]
                queue.push(rows, child, prob[:, j])

        # _visit_buckets :315-325
        for path in index.bucket_paths:
            hit = filter_path_idxs(visiting, path)
            rows = active[hit]
            if rows.shape[0] == 0:
                continue
            order[rows, filled[rows], :] = np.array(path)
            filled[rows] += 1

        queue.resort()  # :250
    return order, t_inf


def knn_inner_product(xq: np.ndarray, xb: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
    """``faiss.knn(xq, xb, k, metric=METRIC_INNER_PRODUCT)`` as called at
    search/li/LearnedIndex.py:360-365 (faiss-cpu 1.7.4, requirements.txt:8; not
    vendored).  Same definition as ``oracle/faiss_shim/faiss.py``."""
    xq = np.ascontiguousarray(xq, dtype=np.float32)
    xb = np.ascontiguousarray(xb, dtype=np.float32)
    assert xq.shape[1] == xb.shape[1]
    nq, nb = xq.shape[0], xb.shape[0]
    D = np.full((nq, k), -FLT_MAX, dtype=np.float32)
    I = np.full((nq, k), -1, dtype=np.int64)
    kk = min(k, nb)
    if kk > 0:
        s = torch.mm(torch.from_numpy(xq), torch.from_numpy(xb).t())
        d_top, i_top = torch.topk(s, kk, dim=1, largest=True, sorted=True)
        D[:, :kk] = d_top.numpy()
        I[:, :kk] = i_top.numpy()
    return D, I


class BucketMajor:
    """The grouping ``data_navigation.groupby(category columns)`` produces
    (search/li/LearnedIndex.py:350): groups in ascending key order, rows of a group
    in their original order.  Built once instead of once per round."""

    def __init__(self, data_search: np.ndarray, data_prediction: np.ndarray, ids: Optional[np.ndarray] = None):
        pred = np.asarray(data_prediction)
        if pred.ndim == 1:
            pred = pred[:, None]
        n, n_levels = pred.shape
        assert data_search.shape[0] == n
        if ids is None:
            ids = np.arange(1, n + 1)  # search/search.py:190-191: 1-based DataFrame index
        # lexicographic key order, stable within a key
        perm = np.lexsort(tuple(pred[:, l] for l in reversed(range(n_levels))))
        sorted_pred = pred[perm]
        if n:
            change = np.any(sorted_pred[1:] != sorted_pred[:-1], axis=1)
            starts = np.concatenate(([0], np.flatnonzero(change) + 1))
        else:
            starts = np.zeros(0, dtype=np.int64)
        self.paths = [tuple(int(v) for v in sorted_pred[s]) for s in starts]
        self.offsets = np.concatenate((starts, [n])).astype(np.int64)
        self.vectors = np.ascontiguousarray(np.asarray(data_search)[perm])
        self.ids = np.asarray(ids)[perm]
        self.n_levels = n_levels

    @classmethod
    def from_grouped(cls, vectors: np.ndarray, offsets: np.ndarray, ids: np.ndarray, paths=None) -> "BucketMajor":
        """Adopt rows that are ALREADY grouped by bucket (ascending bucket id, original
        order inside a bucket) -- the same grouping ``__init__`` derives -- without the
        extra full copy.  ``offsets`` int64[C+1]; buckets without rows are not listed as
        groups, exactly as ``groupby`` omits absent keys (:350)."""
        self = cls.__new__(cls)
        offsets = np.asarray(offsets, dtype=np.int64)
        sizes = np.diff(offsets)
        present = np.flatnonzero(sizes > 0)
        all_paths = paths if paths is not None else [(int(b),) for b in range(len(sizes))]
        self.paths = [tuple(all_paths[b]) for b in present]
        self.offsets = np.concatenate((offsets[present], [offsets[-1]])).astype(np.int64)
        self.vectors = vectors
        self.ids = np.asarray(ids)
        self.n_levels = len(self.paths[0]) if self.paths else 1
        return self


def search_round(
    layout: BucketMajor, queries_search: np.ndarray, bucket_path: np.ndarray, k: int = PER_BUCKET_K
) -> Tuple[np.ndarray, np.ndarray, float]:
    """``_search_single_bucket`` -- search/li/LearnedIndex.py:327-373: every query is
    scanned against the ONE bucket ``bucket_path`` names for it."""
    nq = queries_search.shape[0]
    nns = np.zeros((nq, k), dtype=np.uint32)  # :340
    dists = np.full((nq, k), np.inf, dtype=float)  # :341
    t_seq = 0.0
    for g, path in enumerate(layout.paths):
        lo, hi = layout.offsets[g], layout.offsets[g + 1]
        who = filter_path_idxs(bucket_path, path)  # :353
        if hi - lo == 0 or who.shape[0] == 0:
            continue
        t0 = time.time()
        sim, idx = knn_inner_product(queries_search[who], layout.vectors[lo:hi], k)  # :360-365
        t_seq += time.time() - t0
        nns[who] = layout.ids[lo:hi][idx]  # :370 (idx == -1 wraps to the bucket's last id)
        dists[who] = 1 - sim  # :368, :371 (f32 arithmetic, widened on store)
    return dists, nns, t_seq


def merge_rounds(dists_a, nns_a, dists_b, nns_b, k: int):
    """Cross-round merge -- search/li/LearnedIndex.py:125-146: concatenate, stable
    argsort by distance, keep the first ``k`` columns."""
    d = np.hstack((dists_a, dists_b))
    i = np.hstack((nns_a, nns_b))
    keep = d.argsort(kind="stable", axis=1)[:, :k]
    return np.take_along_axis(d, keep, axis=1), np.take_along_axis(i, keep, axis=1)


def search(
    index: OracleIndex,
    queries_navigation: np.ndarray,
    data_search: Optional[np.ndarray],
    queries_search: np.ndarray,
    data_prediction: Optional[np.ndarray],
    n_categories: Sequence[int],
    n_buckets: int = 1,
    k: int = 10,
    ids: Optional[np.ndarray] = None,
    layout: Optional[BucketMajor] = None,
):
    """``LearnedIndex.search`` -- search/li/LearnedIndex.py:41-161.

    Returns ``(dists f64[nq,k'], nns u32[nq,k'], measured_time, bucket_order)``;
    ``k' = 10`` when ``n_buckets == 1`` (``k`` is not forwarded, :111-117).
    ``data_navigation`` is not an argument: the reference only uses it to carry
    the ``data_prediction`` labels and the row ids (:101-104, :350).
    """
    times: Dict[str, float] = defaultdict(float)
    t_start = time.time()
    order, times["inference"] = precompute_bucket_order(index, queries_navigation, n_buckets, n_categories)
    if layout is None:
        layout = BucketMajor(data_search, data_prediction, ids)
        t_start = time.time() - times["inference"]  # layout build is not part of the reference's search
    out_d = out_i = None
    for r in range(n_buckets):
        t0 = time.time()
        d, i, t_seq = search_round(layout, queries_search, order[:, r, :])
        times["search_within_buckets"] += time.time() - t0
        times["seq_search"] += t_seq
        times["sort"] += 0.0  # never incremented by the reference (:348, :373)
        if out_d is None:
            out_d, out_i = d, i
        else:
            out_d, out_i = merge_rounds(out_d, out_i, d, i, k)
            assert out_d.shape == out_i.shape == (queries_search.shape[0], k)  # :142-146
    times["search"] = time.time() - t_start
    return out_d, out_i, dict(times), order


def merge_ranks(dists_per_rank: Sequence[np.ndarray], nns_per_rank: Sequence[np.ndarray], k: int):
    """Cross-GPU merge (no reference counterpart; the reference is single-process): the
    same stable rule as ``merge_rounds`` applied over ranks in rank order -- the k smallest
    per query, ties keeping the lower rank, then the lower position."""
    d = np.hstack(list(dists_per_rank))
    i = np.hstack(list(nns_per_rank))
    keep = d.argsort(kind="stable", axis=1)[:, :k]
    return np.take_along_axis(d, keep, axis=1), np.take_along_axis(i, keep, axis=1)


# --------------------------------------------------------------------------- #
# Ground truth + recall (Baseline.py, eval/eval.py)
# --------------------------------------------------------------------------- #
def brute_force(queries: np.ndarray, data: np.ndarray, k: int = 10, ids: Optional[np.ndarray] = None, block: int = 1024):
    """Exact k-NN by ``1 - <q, x>`` over ALL rows, ids 1-based -- the ground truth
    ``Baseline.search`` defines (search/li/Baseline.py:11-16; cosine == inner
    product on the L2-normalised vectors the driver feeds, search/search.py:142-144)."""
    n = data.shape[0]
    if ids is None:
        ids = np.arange(1, n + 1)
    tq = torch.from_numpy(np.ascontiguousarray(queries, dtype=np.float32))
    tx = torch.from_numpy(np.ascontiguousarray(data, dtype=np.float32))
    out_d = np.empty((queries.shape[0], k), dtype=np.float32)
    out_i = np.empty((queries.shape[0], k), dtype=np.int64)
    for s in range(0, queries.shape[0], block):
        sim = torch.mm(tq[s : s + block], tx.t())
        d_top, i_top = torch.topk(sim, k, dim=1)
        out_d[s : s + block] = 1 - d_top.numpy()
        out_i[s : s + block] = ids[i_top.numpy()]
    return out_d, out_i


def get_recall(I: np.ndarray, gt: np.ndarray, k: int) -> float:
    """``get_recall`` -- eval/eval.py:37-45 (set-intersection recall@k)."""
    assert k <= I.shape[1]
    assert len(I) == len(gt)
    hits = 0
    for a, b in zip(I, gt):
        hits += len(set(a[:k].tolist()) & set(b[:k].tolist()))
    return hits / (len(I) * k)
