"""Multi-level navigation (``n_levels > 1``) -- the branch of
``LearnedIndex._precompute_bucket_order`` at search/li/LearnedIndex.py:216-252 together with
``_visit_internal_nodes`` (:254-301), ``_visit_buckets`` (:303-325) and ``PriorityQueue``
(search/li/PriorityQueue.py:7-94).

The reference runs a batched best-first search: per query a sorted array of (path,
probability) entries; each iteration pops the most probable entry, runs the popped node's
model on the queries that popped it (one ``predict_proba`` per internal model per iteration)
and pushes every child with the child's RAW probability.  A node's probability is a pure
function of (model, query), so the GPU form is

1. dense grouped inference: every internal model is evaluated on the whole query batch into one
   table ``node_prob f32[nq, n_nodes]`` with a column per tree node (breadth-first numbering) --
   all models of a tree level per launch (``lmi_linear_tc`` on the stacked first layers,
   ``lmi_linear_tc_grouped`` after, ``lmi_softmax_segments`` in place), or one launch set per
   model (``NeuralNetwork.logits`` -> ``lmi_softmax_probs``) when a level is not uniform;
2. ``lmi_navigate_tree``: one warp per query replays the best-first search over that table
   with the frontier in shared memory and writes the flat bucket ids in visiting order.

Flat bucket id of a full path ``(c_0, ..., c_{L-1})`` = mixed radix over ``n_categories``
(``c_0`` most significant) -- the key order ``groupby(category_L1..L)`` iterates in (:350).
"""
from __future__ import annotations

import os
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch

from . import _native
from .model import NeuralNetwork

EMPTY_VALUE = -1  # search/li/PriorityQueue.py:4


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def flat_bucket_ids(paths: np.ndarray, n_categories: Sequence[int]) -> np.ndarray:
    """int64[m] flat ids of the full paths ``paths`` (int[m, L]); rows with an entry outside
    ``[0, n_categories[l])`` (e.g. ``EMPTY_VALUE``) get ``prod(n_categories)`` -- a bucket no
    query can probe."""
    paths = np.asarray(paths, dtype=np.int64)
    cats = np.asarray(list(n_categories), dtype=np.int64)
    total = int(np.prod(cats))
    strides = np.concatenate((np.cumprod(cats[::-1])[::-1][1:], [1])).astype(np.int64)
    valid = ((paths >= 0) & (paths < cats)).all(axis=1)
    flat = (paths * strides).sum(axis=1)
    return np.where(valid, flat, total)


def unflatten_bucket_ids(flat: np.ndarray, n_categories: Sequence[int]) -> np.ndarray:
    """Inverse of ``flat_bucket_ids``: int32[..., L]; negative ids give ``EMPTY_VALUE`` rows."""
    flat = np.asarray(flat, dtype=np.int64)
    cats = np.asarray(list(n_categories), dtype=np.int64)
    strides = np.concatenate((np.cumprod(cats[::-1])[::-1][1:], [1])).astype(np.int64)
    out = (flat[..., None] // strides) % cats
    out = np.where(flat[..., None] < 0, EMPTY_VALUE, out)
    return out.astype(np.int32)


class NavigationTree:
    """Static structure of a multi-level index + the device tables ``lmi_navigate_tree`` reads."""

    def __init__(self, root_model: NeuralNetwork, internal_models: Dict[Tuple[int, ...], NeuralNetwork],
                 bucket_paths: List[Tuple[int, ...]], n_categories: Sequence[int]):
        self.n_categories = [int(c) for c in n_categories]
        L = self.n_levels = len(self.n_categories)
        assert L >= 2
        self.device = root_model.device
        self.root_model = root_model
        self.grouped = os.environ.get("LMI_NAV_GROUPED", "1") != "0"  # 0: one launch set per model (verification)
        # :222 seeds the queue with pred_l1_paths[:, :n_categories[0]]
        assert root_model.n_output_neurons == self.n_categories[0], \
            "root model outputs != n_categories[0]"
        # breadth-first numbering: base[t] = first node id of depth t (1-based depth)
        width = np.cumprod(self.n_categories)  # nodes at depth t = width[t-1]
        self.base = np.concatenate(([0], np.cumsum(width)))  # base[t-1] -> first id of depth t
        self.n_nodes = int(self.base[L])
        self.n_roots = self.n_categories[0]
        first_child = np.full(self.n_nodes, -1, dtype=np.int32)
        n_children = np.zeros(self.n_nodes, dtype=np.int32)
        node_bucket = np.full(self.n_nodes, -1, dtype=np.int32)

        self.models: List[Tuple[int, NeuralNetwork]] = []  # (first child column, model)
        by_depth: Dict[int, Dict[int, NeuralNetwork]] = {}  # depth -> {flat prefix: model}
        for path, model in internal_models.items():
            path = tuple(int(v) for v in path)
            depth = len(path) - path.count(EMPTY_VALUE)  # :283
            assert len(path) == L and 0 < depth < L and all(v == EMPTY_VALUE for v in path[depth:]), \
                f"internal model key {path} is not a right-padded path"
            if any(not (0 <= path[l] < self.n_categories[l]) for l in range(depth)):
                continue  # a node nobody can reach
            v = self._node(path[:depth])
            assert model.n_output_neurons <= self.n_categories[depth], \
                f"model at {path} has more classes than n_categories[{depth}]"
            assert model.input_dim == root_model.input_dim
            fc = int(self.base[depth]) + self._flat(path[:depth]) * self.n_categories[depth]
            first_child[v] = fc
            n_children[v] = model.n_output_neurons
            self.models.append((fc, model))
            by_depth.setdefault(depth, {})[self._flat(path[:depth])] = model
        for path in bucket_paths:
            path = tuple(int(v) for v in path)
            assert len(path) == L, f"bucket path {path} does not have {L} levels"
            if any(not (0 <= path[l] < self.n_categories[l]) for l in range(L)):
                continue  # never popped: children are only ever pushed with classes < n_categories
            node_bucket[self._node(path)] = self._flat(path)
        self.first_child_host, self.n_children_host, self.node_bucket_host = first_child, n_children, node_bucket
        self.first_child = torch.from_numpy(first_child).to(self.device)
        self.n_children = torch.from_numpy(n_children).to(self.device)
        self.node_bucket = torch.from_numpy(node_bucket).to(self.device)
        self.n_reachable_buckets = self._count_reachable()
        # one softmax segment per model: the root's [0, C1) and every internal model's child columns
        seg_start = [0] + [fc for fc, _ in self.models]
        seg_len = [self.n_roots] + [m.n_output_neurons for _, m in self.models]
        self.seg_start = torch.tensor(seg_start, dtype=torch.int32, device=self.device)
        self.seg_len = torch.tensor(seg_len, dtype=torch.int32, device=self.device)
        self._groups = self._stack_levels(by_depth)
        self._root_stack = None  # the root's layers in the stacked-level format (built on first use)

    def _flat(self, prefix: Sequence[int]) -> int:
        f = 0
        for l, c in enumerate(prefix):
            f = f * self.n_categories[l] + int(c)
        return f

    def _node(self, prefix: Sequence[int]) -> int:
        return int(self.base[len(prefix) - 1]) + self._flat(prefix)

    def _count_reachable(self) -> int:
        stack = list(range(self.n_roots))
        found = 0
        while stack:
            v = stack.pop()
            if self.node_bucket_host[v] >= 0:
                found += 1
            fc, nc = int(self.first_child_host[v]), int(self.n_children_host[v])
            stack.extend(range(fc, fc + nc))
        return found

    # ------------------------------------------------------------------ grouped inference
    def _stack_levels(self, by_depth):
        """Per depth: the models of ALL nodes of that depth stacked for the grouped tensor-core
        launches (``lmi_linear_tc`` on the stacked first layer, ``lmi_linear_tc_grouped`` after),
        node order == column order of the table.  ``None`` when the level is not uniform enough
        (different layer shapes) or the tensor-core path does not apply; the per-model loop of
        ``node_probabilities`` covers that."""
        if not self.root_model._tc_ok():
            return None
        levels = []
        for depth in sorted(by_depth):
            models = by_depth[depth]
            G = int(np.prod(self.n_categories[:depth]))
            shapes = {tuple(tuple(W.shape) for W, _ in m.layers[:-1]) + (m.layers[-1][0].shape[1],)
                      for m in models.values()}
            if len(shapes) != 1 or not all(m._tc_ok() for m in models.values()):
                return None
            proto = next(iter(models.values()))
            n_layers = len(proto.layers)
            c_pad = self.n_categories[depth]
            stacked = []
            for i in range(n_layers):
                out_i = c_pad if i + 1 == n_layers else proto.layers[i][0].shape[0]
                in_i = proto.layers[i][0].shape[1]
                if in_i % 4:
                    return None
                W = torch.zeros((G, out_i, in_i), dtype=torch.float32, device=self.device)
                b = torch.zeros((G, out_i), dtype=torch.float32, device=self.device)
                for g, m in models.items():
                    Wm, bm = m.layers[i]
                    W[g, : Wm.shape[0]] = Wm
                    b[g, : bm.shape[0]] = bm
                stacked.append(dict(W=W.contiguous(), b=b.contiguous(), parts=None, in_dim=in_i, out_dim=out_i))
            levels.append(dict(G=G, col0=int(self.base[depth]), layers=stacked))
        return levels

    def _forward_into_table(self, x_parts, d_in, G, layers, table, col0, split):
        """Runs ``G`` stacked models (shared input ``x_parts``) and writes the last layer's logits
        into ``table[:, col0 : col0 + G*out]``."""
        L = _native.lib()
        nq = table.shape[0]
        ptr = lambda t: t.data_ptr() if t is not None else None
        parts, ldx = x_parts, d_in
        for i, lyr in enumerate(layers):
            if lyr["parts"] is None:
                lyr["parts"] = split(lyr["W"], False)
            wp = lyr["parts"]
            last = i + 1 == len(layers)
            width = G * lyr["out_dim"]
            if last:
                y, nxt, ldy, ycol = table, (None, None, None), table.shape[1], col0
            else:
                y, ldy, ycol = None, width, 0
                nxt = tuple(torch.empty((nq, width), dtype=torch.float32, device=self.device) for _ in range(3))
            if i == 0:  # shared input: one plain launch over the stacked output features
                groups, in_dim, out_dim = 1, d_in, width
            else:
                groups, in_dim, out_dim = G, lyr["in_dim"], lyr["out_dim"]
            _native.check(L.lmi_linear_tc_grouped(ptr(parts[0]), ptr(parts[1]), ptr(parts[2]), ldx, ptr(wp[0]),
                                                  ptr(wp[1]), ptr(wp[2]), lyr["b"].data_ptr(), nq, in_dim, out_dim,
                                                  groups, int(not last), ptr(y), ptr(nxt[0]), ptr(nxt[1]),
                                                  ptr(nxt[2]), ldy, ycol, _stream_ptr()), "lmi_linear_tc_grouped")
            parts, ldx = nxt, width

    def _node_probabilities_grouped(self, qn: torch.Tensor) -> torch.Tensor:
        L = _native.lib()
        nq = qn.shape[0]
        table = torch.empty((nq, self.n_nodes), dtype=torch.float32, device=self.device)
        root = self.root_model
        parts = root.split_input(qn)
        if self._root_stack is None:
            self._root_stack = [dict(W=W, b=b, parts=None, in_dim=W.shape[1], out_dim=W.shape[0])
                                for W, b in root.layers]
        self._forward_into_table(parts, root.input_dim, 1, self._root_stack, table, 0, root._split)
        for lvl in self._groups:
            self._forward_into_table(parts, root.input_dim, lvl["G"], lvl["layers"], table, lvl["col0"], root._split)
        _native.check(L.lmi_softmax_segments(table.data_ptr(), nq, self.n_nodes, self.seg_start.data_ptr(),
                                             self.seg_len.data_ptr(), int(self.seg_start.shape[0]), _stream_ptr()),
                      "lmi_softmax_segments")
        return table

    # ----------------------------------------------------------------------------------
    def node_probabilities(self, qn: torch.Tensor) -> torch.Tensor:
        """f32[nq, n_nodes]: root probabilities in columns [0, C1), every internal model's
        in the columns of its children (columns of unreachable nodes stay unwritten)."""
        L = _native.lib()
        nq = qn.shape[0]
        table = torch.empty((nq, self.n_nodes), dtype=torch.float32, device=self.device)
        if nq == 0:
            return table
        if self._groups is not None and self.grouped:
            return self._node_probabilities_grouped(qn)
        parts = self.root_model.split_input(qn)  # one hi/mid/lo split shared by every model
        for col0, model in [(0, self.root_model)] + self.models:
            logits = model.logits(qn, parts=parts)
            _native.check(L.lmi_softmax_probs(logits.data_ptr(), nq, logits.shape[1], table.data_ptr(),
                                              self.n_nodes, col0, _stream_ptr()), "lmi_softmax_probs")
        return table

    def bucket_order(self, qn: torch.Tensor, n_buckets: int) -> Tuple[torch.Tensor, torch.Tensor]:
        """(flat bucket ids int32[nq, n_buckets] in visiting order, node_prob table)."""
        if n_buckets > self.n_reachable_buckets:
            # the reference pops from an empty queue here (PriorityQueue.py:52-56) and never terminates
            raise ValueError(f"n_buckets={n_buckets} exceeds the {self.n_reachable_buckets} reachable buckets")
        L = _native.lib()
        nq = qn.shape[0]
        table = self.node_probabilities(qn)
        order = torch.empty((nq, n_buckets), dtype=torch.int32, device=self.device)
        _native.check(L.lmi_navigate_tree(table.data_ptr(), nq, self.n_nodes, self.first_child.data_ptr(),
                                          self.n_children.data_ptr(), self.node_bucket.data_ptr(), self.n_roots,
                                          n_buckets, order.data_ptr(), _stream_ptr()), "lmi_navigate_tree")
        return order, table
