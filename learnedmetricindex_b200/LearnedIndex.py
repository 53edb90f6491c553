"""Drop-in for ``li.LearnedIndex.LearnedIndex`` (search/li/LearnedIndex.py) whose
``search()`` runs on a B200 through the C ABI of ``include/lmi_b200.h``.

Same constructor (:23-28), same ``search`` signature, return dtypes and id semantics
(:41-51, :161): ``(dists f64[nq,k], nns u32[nq,k], measured_time)``.  The reference's
``B x n_buckets_in_index`` Python double loop (:107 x :350) becomes four launches:

    MLP + softmax + top-B        lmi_linear / lmi_softmax_topn     (model.py:226-241)
    query-to-bucket inversion    lmi_invert                        (utils.py:61-65)
    bucket scan + fused top-10   lmi_scan_topk (tcgen05 / TMA)     (LearnedIndex.py:350-365)
    merge across probes          lmi_merge_topk                    (LearnedIndex.py:125-146, :368-371)

Per round every query is scanned against exactly one bucket and buckets are disjoint,
so after ``n_buckets`` rounds the reference's result is the top-k by ``1 - <q,x>`` over
the union of each query's per-bucket top-10 lists; that is what the fused pass computes.
"""
from __future__ import annotations

import os
import time
from collections import defaultdict
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from . import _native
from .layout import DeviceBucketLayout
from .model import NeuralNetwork, to_device_matrix
from .navigation import NavigationTree, flat_bucket_ids, unflatten_bucket_ids

EMPTY_VALUE = -1  # search/li/PriorityQueue.py:4
PER_BUCKET_K = 10  # `k` is not forwarded to _search_single_bucket (LearnedIndex.py:111-117, :334)
MAX_RANGES = 32  # row ranges a bucket is cut into at most
LISTS_PER_ITEM = 2  # LMI_LISTS_PER_ITEM: partial top-10 lists a scan work item writes per query
SCAN_ROWS_PER_TILE = 64
DEFAULT_RANGE_ROWS = 24576  # upper bound on the rows of a scan work item (see plan_scan; sustained 10M sweep: 8192 +5 %, 16384 0, 24576 -1.9 %, 32768 -1.0 %)
GRAPH_MAX_NQ = int(os.environ.get("LMI_GRAPH_MAX_NQ", 4096))  # `search` replays a CUDA graph up to this batch size
GRAPH_CACHE = 8  # captured (shape, budget) combinations kept per index


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def plan_scan(nq: int, n_probe: int, n_buckets_in_index: int, n_rows: int, max_bucket: int,
              n_sms: int = 148, items_per_sm: int = 8) -> Tuple[int, int]:
    """Rows per scan work item (multiple of 64) and the resulting ranges per bucket.

    Work items are (bucket, row range, tile of 128 probing queries).  Aim at
    ``items_per_sm`` items per SM so the static round-robin over persistent CTAs balances,
    keep ranges long so the per-item top-10 warm-up and query gather amortise, and never
    exceed ``MAX_RANGES`` lists per (query, probe)."""
    pairs = max(1, nq * n_probe)
    touched = min(n_buckets_in_index, pairs)
    tiles = max(touched, -(-pairs // 128))  # sum_b ceil(m_b / 128), roughly
    mean_bucket = max(1, n_rows // max(1, n_buckets_in_index))
    total_rows = tiles * mean_bucket
    want = max(1, n_sms * items_per_sm)
    length = -(-total_rows // want)
    # short ranges keep the dynamic schedule balanced and the query tiles of a range within an L2
    # window of each other; long enough that the per-item query gather and top-10 warm-up amortise
    length = min(length, int(os.environ.get("LMI_RANGE_ROWS", DEFAULT_RANGE_ROWS)))
    length = max(length, 1024, -(-max_bucket // MAX_RANGES))
    length = min(length, max(max_bucket, SCAN_ROWS_PER_TILE))
    length = -(-length // SCAN_ROWS_PER_TILE) * SCAN_ROWS_PER_TILE
    rmax = max(1, -(-max_bucket // length))
    assert rmax <= MAX_RANGES
    return length, rmax


class LearnedIndex:
    def __init__(self, root_model, internal_models=None, bucket_paths=None, scan_impl: int = 0):
        self.root_model = NeuralNetwork.from_reference(root_model)
        """The root model of the index."""
        self.internal_models = {tuple(p): NeuralNetwork.from_reference(m) for p, m in (internal_models or {}).items()}
        """Path -> internal model (multi-level indexes)."""
        self.bucket_paths = [tuple(p) for p in (bucket_paths or [])]
        """List of paths to the buckets."""
        self.scan_impl = scan_impl  # 0: tcgen05/TMA kernel; 1: CUDA-core verification kernel (tests)
        self.use_graphs = os.environ.get("LMI_CUDA_GRAPH", "1") != "0"
        self.device = self.root_model.device
        self._layout: Optional[DeviceBucketLayout] = None
        self._layout_key = None
        self._layout_refs = None
        self.last_bucket_order: Optional[torch.Tensor] = None
        self.last_probs: Optional[torch.Tensor] = None
        self._tree: Optional[NavigationTree] = None
        self._graphs: Dict[tuple, dict] = {}
        self._n_sms = torch.cuda.get_device_properties(self.device).multi_processor_count

    @classmethod
    def from_reference(cls, li, **kw) -> "LearnedIndex":
        """From a (possibly unpickled) reference ``LearnedIndex`` (:23-39)."""
        return cls(li.root_model, getattr(li, "internal_models", {}), getattr(li, "bucket_paths", []), **kw)

    # ------------------------------------------------------------------ pickling ----
    def __getstate__(self):
        """``save_as_pickle(filename, li)`` (search/li/utils.py:14-29, search/search.py:226-235) keeps the
        models and bucket paths; the device layout, like ``data_prediction`` in the reference
        (search.py:228), is not part of the pickle and is rebuilt by the first ``search``."""
        return {"root_model": self.root_model, "internal_models": self.internal_models,
                "bucket_paths": self.bucket_paths, "scan_impl": self.scan_impl}

    def __setstate__(self, state):
        self.__init__(state["root_model"], state["internal_models"], state["bucket_paths"],
                      scan_impl=state.get("scan_impl", 0))

    # ------------------------------------------------------------------ layout ------
    def attach_layout(self, layout: DeviceBucketLayout) -> None:
        """Use a device layout built elsewhere (e.g. data generated on the GPU)."""
        self._layout = layout
        self._layout_key = "attached"
        self._layout_refs = None
        self._graphs.clear()  # captured graphs hold pointers into the previous layout

    def prepare(self, data_search, data_prediction, n_categories, data_navigation=None) -> DeviceBucketLayout:
        """Build (once) the bucket-major device layout for ``data_search`` / ``data_prediction``."""
        key = (id(data_search), id(data_prediction), tuple(n_categories))
        if self._layout is not None and self._layout_key == key:
            return self._layout
        pred = np.asarray(data_prediction)
        if pred.ndim == 1:
            pred = pred[:, None]
        assert pred.shape[1] == len(n_categories), "data_prediction columns != levels"
        if len(n_categories) == 1:
            labels, n_layout_buckets = pred[:, 0], int(n_categories[0])
        else:
            # groupby(category_L1..L) key order == mixed-radix flat id (:343-350); rows whose path is
            # incomplete (EMPTY_VALUE) land in one trailing bucket no query can probe
            labels = flat_bucket_ids(pred, n_categories)
            n_layout_buckets = int(np.prod(n_categories))
            if (labels == n_layout_buckets).any():
                n_layout_buckets += 1
        ids = None
        if data_navigation is not None and hasattr(data_navigation, "index"):
            ids = np.asarray(data_navigation.index)
            if hasattr(data_search, "index") and not data_search.index.equals(data_navigation.index):
                data_search = data_search.loc[data_navigation.index]  # label lookup, as :357 does
        self._layout = DeviceBucketLayout.from_host(data_search, labels, n_layout_buckets, ids=ids,
                                                    device=self.device)
        self._layout_key = key
        self._layout_refs = (data_search, data_prediction)  # keep ids from being recycled
        self._graphs.clear()
        return self._layout

    # ------------------------------------------------------------------ navigation --
    def navigation_tree(self, n_categories: List[int]) -> NavigationTree:
        if self._tree is None or self._tree.n_categories != [int(c) for c in n_categories]:
            self._tree = NavigationTree(self.root_model, self.internal_models, self.bucket_paths, n_categories)
        return self._tree

    def _bucket_order_device(self, qn: torch.Tensor, n_buckets: int, n_categories: List[int]):
        """(bucket ids int32[nq, n_buckets] -- flat ids for a multi-level index --, probabilities)."""
        if len(n_categories) != 1:
            return self.navigation_tree(n_categories).bucket_order(qn, n_buckets)
        C = self.root_model.n_output_neurons
        if n_buckets > C:
            raise ValueError(f"n_buckets={n_buckets} exceeds the {C} buckets of the index")
        return self.root_model.topn_device(qn, n_buckets)

    def _precompute_bucket_order(self, queries_navigation, n_buckets: int, n_categories: List[int]):
        """search/li/LearnedIndex.py:163-214: int32[nq, n_buckets, n_levels] and inference seconds."""
        qn = to_device_matrix(queries_navigation, self.device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        order, _ = self._bucket_order_device(qn, n_buckets, list(n_categories))
        e1.record()
        e1.synchronize()
        order = order.cpu().numpy()
        if len(n_categories) == 1:
            return order.reshape(qn.shape[0], n_buckets, 1), e0.elapsed_time(e1) / 1e3
        return unflatten_bucket_ids(order, n_categories), e0.elapsed_time(e1) / 1e3

    # ------------------------------------------------------------------ search ------
    def search_device(self, qn: torch.Tensor, qs: torch.Tensor, n_categories: List[int], n_buckets: int, k: int,
                      events=None, out=None, bucket_order: Optional[torch.Tensor] = None):
        """Everything between the H2D and D2H copies.  ``qn`` / ``qs``: device matrices
        (fp16 or f32).  Returns device tensors (dists f64[nq,k'], nns int32-bits[nq,k']);
        ``out=(dists, nns)`` supplies them (e.g. views into an all-gather send buffer)."""
        L = _native.lib()
        layout = self._layout
        assert layout is not None, "call prepare()/attach_layout() first"
        dev = self.device
        nq = qs.shape[0]
        B = int(n_buckets)
        C = layout.n_buckets
        k_out = PER_BUCKET_K if B == 1 else int(k)
        assert 0 < k_out <= PER_BUCKET_K * B, "k exceeds what n_buckets per-bucket top-10 lists can supply"
        assert qs.shape[1] == layout.d_data, "query / data dimensionality mismatch"  # faiss: d == d2

        def mark(i):
            if events is not None:
                events[i].record()

        if bucket_order is None:
            mark(0)
            order, probs = self._bucket_order_device(qn, B, n_categories)
        else:  # computed elsewhere (query-sharded navigation of the multi-GPU path); events[0] is the caller's
            order, probs = bucket_order, None
            assert order.shape == (nq, B) and order.dtype == torch.int32 and order.is_contiguous()
        self.last_bucket_order, self.last_probs = order, probs
        mark(1)

        if out is not None:
            dists, nns = out
            assert dists.shape == (nq, k_out) and dists.dtype == torch.float64 and dists.is_contiguous()
            assert nns.shape == (nq, k_out) and nns.dtype == torch.int32 and nns.is_contiguous()
        else:
            dists = torch.empty((nq, k_out), dtype=torch.float64, device=dev)
            nns = torch.empty((nq, k_out), dtype=torch.int32, device=dev)
        if nq == 0:
            for i in (2, 3, 4):
                mark(i)
            return dists, nns

        if layout.d != layout.d_data:  # zero-pad K to the scan kernel's multiple of 64
            q16 = torch.empty((nq, layout.d), dtype=torch.float16, device=dev)
            _native.check(L.lmi_scatter_rows(qs.data_ptr(), int(qs.dtype == torch.float16), None, nq, layout.d_data,
                                             q16.data_ptr(), layout.d, _stream_ptr()), "lmi_scatter_rows")
        elif qs.dtype == torch.float16:
            q16 = qs
        else:
            q16 = torch.empty(qs.shape, dtype=torch.float16, device=dev)
            _native.check(L.lmi_cast_f16(qs.data_ptr(), q16.data_ptr(), qs.numel(), _stream_ptr()), "lmi_cast_f16")

        range_len, rmax = plan_scan(nq, B, C, layout.n, layout.max_bucket, self._n_sms)
        P = nq * B
        max_items = int(L.lmi_max_items(nq, B, C, rmax))
        lmax = rmax * LISTS_PER_ITEM
        counts = torch.empty(C, dtype=torch.int32, device=dev)
        bucket_lists = torch.empty(C, dtype=torch.int32, device=dev)
        offsets = torch.empty(C + 1, dtype=torch.int32, device=dev)
        pair_list = torch.empty(P, dtype=torch.int32, device=dev)
        slot_of_pair = torch.empty(P, dtype=torch.int32, device=dev)
        items = torch.empty(max_items * 8, dtype=torch.int32, device=dev)
        n_items = torch.zeros(2, dtype=torch.int32, device=dev)
        work = torch.empty(int(L.lmi_invert_workspace_ints(nq, B, C)), dtype=torch.int32, device=dev)
        _native.check(L.lmi_invert(order.data_ptr(), nq, B, C, layout.bucket_off.data_ptr(), range_len,
                                   counts.data_ptr(), offsets.data_ptr(), bucket_lists.data_ptr(), pair_list.data_ptr(),
                                   slot_of_pair.data_ptr(), items.data_ptr(), n_items.data_ptr(), max_items,
                                   work.data_ptr(), _stream_ptr()), "lmi_invert")
        mark(2)

        part_sim = torch.empty(P * lmax * PER_BUCKET_K, dtype=torch.float32, device=dev)
        part_row = torch.empty(P * lmax * PER_BUCKET_K, dtype=torch.int32, device=dev)
        if layout.n > 0:
            _native.check(L.lmi_scan_topk(layout.vectors.data_ptr(), layout.n, layout.d, q16.data_ptr(), nq, B,
                                          pair_list.data_ptr(), items.data_ptr(), n_items.data_ptr(), max_items,
                                          lmax, part_sim.data_ptr(), part_row.data_ptr(), self.scan_impl,
                                          _stream_ptr()), "lmi_scan_topk")
        mark(3)
        _native.check(L.lmi_merge_topk(part_sim.data_ptr(), part_row.data_ptr(), slot_of_pair.data_ptr(),
                                       order.data_ptr(), layout.bucket_off.data_ptr(), layout.ids.data_ptr(),
                                       bucket_lists.data_ptr(), nq, B, C, lmax, k_out, dists.data_ptr(),
                                       nns.data_ptr(), _stream_ptr()),
                      "lmi_merge_topk")
        mark(4)
        self.last_plan = dict(range_len=range_len, rmax=rmax, max_items=max_items, n_items=n_items,
                              counts=counts)
        return dists, nns

    # ------------------------------------------------------------------ CUDA graph --
    def search_graphed(self, qn, qs, n_categories: List[int], n_buckets: int, k: int):
        """``search_device`` as ONE CUDA-graph replay: the ~15 launches of a search (split, Linear
        layers, softmax/top-n or tree navigation, inversion, scan, merge) are captured once per
        (batch shape, budget) with static input/output buffers, so a small batch costs one graph
        launch instead of ~15 host-side launches.  ``qn`` / ``qs`` may be host (numpy / pinned
        torch) or device matrices; they are copied into the static inputs.  Returns
        ``(dists, nns, stage_ms)`` -- the static output tensors, valid until the next replay of the
        same shape, and the per-stage milliseconds of the eager run made at capture time."""
        shared = qs is qn
        def as_tensor(x):
            if isinstance(x, torch.Tensor):
                return x
            a = np.asarray(x)
            if a.dtype not in (np.float16, np.float32):
                a = a.astype(np.float32)  # model.py:104
            return torch.from_numpy(np.ascontiguousarray(a))
        tn = as_tensor(qn)
        ts = tn if shared else as_tensor(qs)
        key = (tuple(tn.shape), tn.dtype, tuple(ts.shape), ts.dtype, shared, tuple(n_categories), int(n_buckets),
               int(k), id(self._layout), self.scan_impl)
        entry = self._graphs.get(key)
        if entry is None:
            if len(self._graphs) >= GRAPH_CACHE:
                self._graphs.pop(next(iter(self._graphs)))
            sn = torch.empty(tn.shape, dtype=tn.dtype, device=self.device)
            ss = sn if shared else torch.empty(ts.shape, dtype=ts.dtype, device=self.device)
            sn.copy_(tn, non_blocking=True)
            if not shared:
                ss.copy_(ts, non_blocking=True)
            # eager run first: fills every lazy cache (weight splits, tensor-map driver entry point, the
            # navigation tree) outside the capture and yields the per-stage timings
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
            self.search_device(sn, ss, n_categories, n_buckets, k, events=ev)
            torch.cuda.current_stream().synchronize()
            stage_ms = [ev[j].elapsed_time(ev[j + 1]) for j in range(4)]
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                d, i = self.search_device(sn, ss, n_categories, n_buckets, k)
            entry = dict(graph=graph, qn=sn, qs=ss, dists=d, nns=i, stage_ms=stage_ms, order=self.last_bucket_order)
            self._graphs[key] = entry
        else:
            entry["qn"].copy_(tn, non_blocking=True)
            if not shared:
                entry["qs"].copy_(ts, non_blocking=True)
        entry["graph"].replay()
        self.last_bucket_order = entry["order"]
        return entry["dists"], entry["nns"], entry["stage_ms"]

    def search(
        self,
        data_navigation,
        queries_navigation,
        data_search,
        queries_search,
        data_prediction,
        n_categories: List[int],
        n_buckets: int = 1,
        k: int = 10,
    ) -> Tuple[np.ndarray, np.ndarray, Dict[str, float]]:
        """Searches for ``k`` nearest neighbours of each query in its ``n_buckets`` most
        probable buckets -- search/li/LearnedIndex.py:41-161.

        Inputs are the reference's (host DataFrames / ndarrays); ``data_*`` may be ``None``
        after ``prepare()`` / ``attach_layout()``.  Unlike the reference, ``data_navigation``
        is not mutated (:101-104, :153-157 add and drop ``category_L*`` columns; net effect
        on the frame is identical)."""
        measured_time: Dict[str, float] = defaultdict(float)
        s = time.time()
        n_categories = list(n_categories)
        if data_search is not None and data_prediction is not None:
            self.prepare(data_search, data_prediction, n_categories, data_navigation)
        assert self._layout is not None, "no index data: pass data_search/data_prediction or attach_layout()"

        if 0 < len(queries_search) <= GRAPH_MAX_NQ and self.use_graphs:
            # latency-bound batch sizes: one graph replay; the stage split reported is the one measured
            # when this shape was captured, "search" is this call's wall time
            dists_d, nns_d, st = self.search_graphed(
                queries_navigation, queries_navigation if queries_search is queries_navigation else queries_search,
                n_categories, n_buckets, k)
            dists = dists_d.cpu().numpy()
            nns = nns_d.cpu().numpy().view(np.uint32)
            stage = [t / 1e3 for t in st]
        else:
            qn = to_device_matrix(queries_navigation, self.device)
            qs = qn if queries_search is queries_navigation else to_device_matrix(queries_search, self.device)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
            dists_d, nns_d = self.search_device(qn, qs, n_categories, n_buckets, k, events=ev)
            dists = dists_d.cpu().numpy()
            nns = nns_d.cpu().numpy().view(np.uint32)
            stage = [ev[j].elapsed_time(ev[j + 1]) / 1e3 for j in range(4)]

        measured_time["inference"] = stage[0]
        measured_time["search_within_buckets"] = stage[1] + stage[2] + stage[3]
        measured_time["seq_search"] = stage[2]
        measured_time["sort"] = stage[3]
        measured_time["search"] = time.time() - s
        return dists, nns, measured_time
