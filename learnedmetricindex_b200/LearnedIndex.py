"""Drop-in for ``li.LearnedIndex.LearnedIndex`` (search/li/LearnedIndex.py) whose
``search()`` runs on a B200 through the C ABI of ``include/lmi_b200.h``.

Same constructor (:23-28), same ``search`` signature, return dtypes and id semantics
(:41-51, :161): ``(dists f64[nq,k], nns u32[nq,k], measured_time)``.  The reference's
``B x n_buckets_in_index`` Python double loop (:107 x :350) becomes four stages, six launches, replayed
as ONE CUDA graph per search:

    MLP + softmax + top-B        lmi_mlp_topn, one launch (tcgen05 tf32x3; per-layer lmi_linear_tc +
                                 lmi_softmax_topn for shapes it does not take)  (model.py:226-241)
    query-to-bucket inversion    lmi_invert, three launches        (utils.py:61-65)
    bucket scan + fused top-10   lmi_scan_topk (tcgen05 / TMA)     (LearnedIndex.py:350-365)
    merge across probes          lmi_merge_topk                    (LearnedIndex.py:125-146, :368-371)

Per round every query is scanned against exactly one bucket and buckets are disjoint,
so after ``n_buckets`` rounds the reference's result is the top-k by ``1 - <q,x>`` over
the union of each query's per-bucket top-10 lists; that is what the fused pass computes.
"""
from __future__ import annotations

import os
import time
import warnings
from collections import OrderedDict, defaultdict
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from . import _native
from .layout import DeviceBucketLayout
from .model import NeuralNetwork, to_device_matrix
from .navigation import NavigationTree, flat_bucket_ids, unflatten_bucket_ids

EMPTY_VALUE = -1  # search/li/PriorityQueue.py:4
PER_BUCKET_K = 10  # `k` is not forwarded to _search_single_bucket (LearnedIndex.py:111-117, :334)
MAX_RANGES = 32  # row ranges a bucket is cut into at most (lmi_merge_topk: <= 256 lists, <= 32 in its fast path)
LISTS_PER_ITEM = 2  # LMI_LISTS_PER_ITEM: partial top-10 lists a scan work item writes per query
SCAN_ROWS_PER_TILE = 64
SCAN_MAX_D = 768  # widest search representation the tcgen05 scan kernels take (lmi_scan.cu: SCAN_MAX_D)
DEFAULT_RANGE_ROWS = 24576  # upper bound on the rows of a scan work item (see plan_scan; sustained 10M sweep: 8192 +5 %, 16384 0, 24576 -1.9 %, 32768 -1.0 %)
RANGE_DIV_FULL_TILES, RANGE_DIV_SPARSE_TILES, RANGE_DIV_FEW_QUERIES = 1.5, 3.0, 6.0  # plan_scan: long items per SM
DEFAULT_SCAN_IMPL = 0
GRAPH_CACHE = 8  # captured (shape, budget) combinations kept per index (least recently used goes first)
WS_CACHE = 8  # batch shapes whose workspace is kept
WS_ARRAYS = 12  # LMI_WS_ARRAYS
(WS_COUNTS, WS_OFFSETS, WS_BUCKET_LISTS, WS_PAIR_LIST, WS_SLOT_OF_PAIR, WS_ITEMS, WS_N_ITEMS, WS_INVERT_SCRATCH,
 WS_PART_SIM, WS_PART_ROW, WS_Q16, WS_GATE) = range(WS_ARRAYS)


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def _round_up(v: int, m: int) -> int:
    return -(-int(v) // m) * m


def plan_scan(nq: int, n_probe: int, n_buckets_in_index: int, n_rows: int, max_bucket: int,
              n_sms: int = 148, n_nonempty: Optional[int] = None, rows_per_tile: int = SCAN_ROWS_PER_TILE) -> Dict[str, int]:
    """How ``lmi_invert`` cuts the probed buckets into scan work items.

    Work items are (bucket, row range, tile of 128 probing queries), drawn in list order by the scan's
    persistent CTAs.  Two lengths (rows per item): ``range_len`` while there is plenty of work -- long, so
    the per-item query gather and top-10 warm-up amortise -- and a shorter ``tail_range_len`` for the last
    ``tail_permille`` of the launch's work, so the CTAs run dry within a short item of each other (with one
    length the ragged end of the launch is about half a long item per SM).  ``n_nonempty``: buckets that hold
    rows in THIS layout (a rank of a sharded index owns only some); the estimate of the per-SM work uses it.
    Returns ``dict(range_len, tail_range_len, tail_permille, rmax)``; ``rmax`` = most ranges of any bucket."""
    C = max(1, n_buckets_in_index)
    local = C if n_nonempty is None else max(1, min(C, n_nonempty))
    pairs = max(1.0, nq * n_probe * local / C)  # probes that land in a bucket held here
    tiles = max(min(local, pairs), pairs / 128.0)  # sum_b ceil(m_b / 128), roughly
    mean_bucket = max(1.0, n_rows / local)
    rows_per_sm = tiles * mean_bucket / max(1, n_sms)  # rows (x 128 queries) one CTA scans in the launch
    floor_rows = max(1024, -(-max_bucket // MAX_RANGES))
    env = os.environ.get
    # long items per SM the main cut aims for.  Measured on a B200 (tools/shard_plan_lab.py,
    # profiles/r03_shard_plan_lab_*.jsonl): where the query tiles are full (>= 128 probing queries per bucket:
    # the tensor-bound regime) few, long items win -- a 1/8 bucket shard of the 10M index scans in 0.70 ms with
    # 1.5 items per SM against 0.87-1.06 ms with 6 (every item re-gathers its 128 queries into tensor memory and
    # re-warms its top-10, and the two-level cut below evens out the end of the launch anyway); batches that
    # leave the tiles mostly empty are HBM bound and want more items in flight (3: -3 % at 100-1000 queries);
    # a handful of queries keeps the finest cut so that every SM streams.
    per_bucket = pairs / local
    div = float(env("LMI_RANGE_DIV", 0)) or (RANGE_DIV_FULL_TILES if per_bucket >= 128 else
                                             RANGE_DIV_SPARSE_TILES if per_bucket >= 2 else RANGE_DIV_FEW_QUERIES)
    main = int(env("LMI_RANGE_ROWS", 0)) or int(min(DEFAULT_RANGE_ROWS, max(rows_per_sm / div, 2048.0)))
    main = max(main, floor_rows)
    main = _round_up(min(main, max(max_bucket, rows_per_tile)), rows_per_tile)
    tail = int(env("LMI_TAIL_ROWS", 0)) or main // 3
    # keep every bucket within 16 ranges (= 32 partial lists per probe: the merge's one-list-per-lane path)
    tail = _round_up(max(tail, floor_rows, -(-max_bucket // 16)), rows_per_tile)
    permille = env("LMI_TAIL_PERMILLE")
    if permille is not None:
        permille = int(permille)
    else:
        # the short items should cover about the last 1.25 long items of every SM
        permille = int(min(400.0, max(50.0, 1250.0 * main / max(rows_per_sm, 1.0))))
    if tail >= main:
        tail, permille = 0, 0
    rmax = max(1, -(-max_bucket // (tail or main)))
    assert rmax <= MAX_RANGES
    return dict(range_len=main, tail_range_len=tail, tail_permille=permille, rmax=rmax)


def results_to_host(owner, dists_d: torch.Tensor, nns_d: torch.Tensor):
    """Device results -> fresh host arrays ``(dists f64, nns u32)`` (LearnedIndex.py:161): both D2H copies go into
    pinned staging buffers kept per result shape on ``owner`` and are enqueued back to back, one synchronisation."""
    cache = owner.__dict__.setdefault("_host_out", {})
    key = (tuple(dists_d.shape), tuple(nns_d.shape))
    buf = cache.get(key)
    if buf is None:
        if len(cache) >= 8:
            cache.pop(next(iter(cache)))
        buf = (torch.empty(dists_d.shape, dtype=torch.float64).pin_memory(),
               torch.empty(nns_d.shape, dtype=torch.int32).pin_memory())
        cache[key] = buf
    buf[0].copy_(dists_d, non_blocking=True)
    buf[1].copy_(nns_d, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return buf[0].numpy().copy(), buf[1].numpy().view(np.uint32).copy()


class _BatchWorkspace:
    """Every transient array of one batch shape, in ONE allocation carved by ``lmi_workspace_layout``
    (SURVEY.md section 8b): a search allocates nothing but its two small result tensors."""

    def __init__(self, lib, layout: DeviceBucketLayout, nq: int, B: int, n_sms: int, q_needs_copy: bool,
                 rows_per_tile: int = SCAN_ROWS_PER_TILE):
        import ctypes

        self.nq, self.B = nq, B
        C = layout.n_buckets
        self.plan = plan_scan(nq, B, C, layout.n, layout.max_bucket, n_sms, layout.n_nonempty, rows_per_tile)
        self.rmax = self.plan["rmax"]
        self.lmax = self.rmax * LISTS_PER_ITEM
        self.max_items = int(lib.lmi_max_items(nq, B, C, self.rmax))
        offs = (ctypes.c_int64 * (WS_ARRAYS + 1))()
        total = int(lib.lmi_workspace_layout(nq, B, C, self.rmax, layout.d if q_needs_copy else 0, offs))
        assert total > 0
        self.buf = torch.empty(total, dtype=torch.uint8, device=layout.device)
        base = self.buf.data_ptr()
        self.ptr = [base + int(offs[i]) for i in range(WS_ARRAYS)]
        self.counts = self.buf[int(offs[WS_COUNTS]): int(offs[WS_COUNTS]) + 4 * C].view(torch.int32)
        self.n_items = self.buf[int(offs[WS_N_ITEMS]): int(offs[WS_N_ITEMS]) + 8].view(torch.int32)
        self.q16 = None
        if q_needs_copy:
            o = int(offs[WS_Q16])
            self.q16 = self.buf[o: o + 2 * nq * layout.d].view(torch.float16).view(nq, layout.d)
        self.inexact = torch.zeros(1, dtype=torch.int64, device=layout.device)  # queries changed by the fp16 cast


class LearnedIndex:
    def __init__(self, root_model, internal_models=None, bucket_paths=None, scan_impl: Optional[int] = None,
                 allow_fp16_rounding: Optional[bool] = None):
        self.root_model = NeuralNetwork.from_reference(root_model)
        """The root model of the index."""
        self.internal_models = {tuple(p): NeuralNetwork.from_reference(m) for p, m in (internal_models or {}).items()}
        """Path -> internal model (multi-level indexes)."""
        self.bucket_paths = [tuple(p) for p in (bucket_paths or [])]
        """List of paths to the buckets."""
        # lmi_scan_topk impl: 0 = tcgen05/TMA kernel, one CTA per work item, 64-row score tiles; 3 = the same with
        # 128-row score tiles (one accumulator); 4 = 128-row tiles, two accumulators, the queries split between
        # tensor memory and shared memory; 2 = 64-row tiles over clusters of two CTAs that share bucket rows by
        # TMA multicast; 1 = CUDA-core verification kernel (tests)
        self.scan_impl = int(os.environ.get("LMI_SCAN_IMPL", DEFAULT_SCAN_IMPL)) if scan_impl is None else int(scan_impl)
        # The scan computes f32 arithmetic on fp16 values -- exactly the reference's arithmetic when the data is
        # fp16 on disk (SISAP `emb`).  f32 data that fp16 cannot represent is ROUNDED (2^-11 relative per
        # component): None = warn once, True = accept silently, False = refuse (ValueError).
        self.allow_fp16_rounding = allow_fp16_rounding
        self.use_graphs = os.environ.get("LMI_CUDA_GRAPH", "1") != "0"
        self.device = self.root_model.device
        self._layout: Optional[DeviceBucketLayout] = None
        self._layout_key = None
        self._layout_refs = None
        self.last_bucket_order: Optional[torch.Tensor] = None
        self.last_probs: Optional[torch.Tensor] = None
        self.last_plan: Optional[dict] = None
        self._tree: Optional[NavigationTree] = None
        self._graphs: "OrderedDict[tuple, dict]" = OrderedDict()
        self._workspaces: "OrderedDict[tuple, _BatchWorkspace]" = OrderedDict()
        self._warned_rounding = False
        self._n_sms = torch.cuda.get_device_properties(self.device).multi_processor_count
        self.scan_flags = int(os.environ.get("LMI_SCAN_FLAGS", "0"))  # lmi_scan_topk tuning flags (results identical)
        # the ranges of a bucket share their running 10th-best through global memory (lmi_scan_topk's `gate`)
        self.share_gate = os.environ.get("LMI_SCAN_SHARE_GATE", "1") != "0"

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
                "bucket_paths": self.bucket_paths, "scan_impl": self.scan_impl,
                "allow_fp16_rounding": self.allow_fp16_rounding}

    def __setstate__(self, state):
        self.__init__(state["root_model"], state["internal_models"], state["bucket_paths"],
                      scan_impl=state.get("scan_impl", 0), allow_fp16_rounding=state.get("allow_fp16_rounding"))

    # ------------------------------------------------------------------ layout ------
    def _forget_layout_state(self):
        self._graphs.clear()  # captured graphs hold pointers into the previous layout
        self._workspaces.clear()

    def _check_rounding(self, inexact: int, what: str):
        """fp16 storage changed ``inexact`` values of ``what``: refuse, warn once, or accept."""
        if inexact <= 0 or self.allow_fp16_rounding is True:
            return
        msg = (f"{inexact} values of {what} are not representable in fp16 and were rounded (relative error <= 2^-11 "
               f"per component; distances of unit vectors move by up to ~5e-4, typically ~1e-5): the scan computes "
               f"f32 arithmetic on fp16 values, which equals the reference's arithmetic only for data that is fp16 "
               f"on disk (SISAP `emb`).  Pass allow_fp16_rounding=True to accept, False to refuse.")
        if self.allow_fp16_rounding is False:
            raise ValueError(msg)
        if not self._warned_rounding:
            warnings.warn(msg, UserWarning, stacklevel=3)
            self._warned_rounding = True

    def attach_layout(self, layout: DeviceBucketLayout) -> None:
        """Use a device layout built elsewhere (e.g. data generated on the GPU)."""
        self._check_rounding(layout.inexact_values, "data_search")
        self._layout = layout
        self._layout_key = "attached"
        self._layout_refs = None
        self._forget_layout_state()

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
        layout = DeviceBucketLayout.from_host(data_search, labels, n_layout_buckets, ids=ids, device=self.device)
        self._check_rounding(layout.inexact_values, "data_search")
        self._layout = layout
        self._layout_key = key
        self._layout_refs = (data_search, data_prediction)  # keep ids from being recycled
        self._forget_layout_state()
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
    def _workspace(self, nq: int, B: int, q_needs_copy: bool) -> _BatchWorkspace:
        key = (nq, B, bool(q_needs_copy), id(self._layout))
        ws = self._workspaces.get(key)
        if ws is None:
            if len(self._workspaces) >= WS_CACHE:
                self._workspaces.popitem(last=False)
            ws = _BatchWorkspace(_native.lib(), self._layout, nq, B, self._n_sms, q_needs_copy,
                                 128 if (self.scan_impl & 0xff) in (3, 4) else SCAN_ROWS_PER_TILE)
            self._workspaces[key] = ws
        else:
            self._workspaces.move_to_end(key)
        return ws

    def search_device(self, qn: torch.Tensor, qs: torch.Tensor, n_categories: List[int], n_buckets: int, k: int,
                      events=None, out=None, bucket_order: Optional[torch.Tensor] = None):
        """Everything between the H2D and D2H copies.  ``qn`` / ``qs``: device matrices
        (fp16 or f32).  Returns device tensors (dists f64[nq,k'], nns int32-bits[nq,k']);
        ``out=(dists, nns)`` supplies them (e.g. views into an all-gather send buffer).
        All transient arrays live in a per-shape workspace that is reused by the next search of the
        same shape on this index: one stream at a time per index (the reference is single-threaded too)."""
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

        probe_out = None
        if out is not None:
            # (dists, nns) or (dists, nns, probe ranks int16[nq,k]) -- the third is what a cross-GPU merge needs to
            # order equal distances like the unsharded search (lmi_merge_topk's probe_rank)
            dists, nns = out[0], out[1]
            if len(out) > 2:
                probe_out = out[2]
                assert probe_out.shape == (nq, k_out) and probe_out.element_size() == 2 and probe_out.is_contiguous()
            assert dists.shape == (nq, k_out) and dists.dtype == torch.float64 and dists.is_contiguous()
            assert nns.shape == (nq, k_out) and nns.dtype == torch.int32 and nns.is_contiguous()
        else:
            dists = torch.empty((nq, k_out), dtype=torch.float64, device=dev)
            nns = torch.empty((nq, k_out), dtype=torch.int32, device=dev)
        if nq == 0:
            for i in (2, 3, 4):
                mark(i)
            return dists, nns

        needs_copy = layout.d != layout.d_data or qs.dtype != torch.float16
        ws = self._workspace(nq, B, needs_copy)
        st = _stream_ptr()
        if layout.d != layout.d_data:  # zero-pad K to the scan kernel's multiple of 64
            q16 = ws.q16
            _native.check(L.lmi_scatter_rows(qs.data_ptr(), int(qs.dtype == torch.float16), None, nq, layout.d_data,
                                             q16.data_ptr(), layout.d, ws.inexact.data_ptr(), st), "lmi_scatter_rows")
        elif qs.dtype == torch.float16:
            q16 = qs
        else:
            q16 = ws.q16
            _native.check(L.lmi_cast_f16(qs.data_ptr(), q16.data_ptr(), qs.numel(), ws.inexact.data_ptr(), st),
                          "lmi_cast_f16")
        self._last_ws = ws

        plan, P = ws.plan, ws.ptr
        _native.check(L.lmi_invert(order.data_ptr(), nq, B, C, layout.bucket_off.data_ptr(), plan["range_len"],
                                   plan["tail_range_len"], plan["tail_permille"], P[WS_COUNTS], P[WS_OFFSETS],
                                   P[WS_BUCKET_LISTS], P[WS_PAIR_LIST], P[WS_SLOT_OF_PAIR], P[WS_ITEMS], P[WS_N_ITEMS],
                                   ws.max_items, P[WS_INVERT_SCRATCH], st), "lmi_invert")
        mark(2)
        if layout.n > 0:
            impl = self.scan_impl
            if layout.d > SCAN_MAX_D and (impl & 0xff) != 1:
                # the tcgen05 kernels keep a work item's queries in tensor memory (<= 768 dimensions); wider data
                # still searches correctly, on the CUDA-core kernel -- far from the tensor roofline, so say so once
                if not getattr(self, "_warned_wide", False):
                    warnings.warn(f"search dimensionality {layout.d_data} > {SCAN_MAX_D}: the scan runs on the "
                                  f"CUDA-core kernel (lmi_scan_topk impl 1), not on the tensor cores", UserWarning,
                                  stacklevel=3)
                    self._warned_wide = True
                impl = 1
            _native.check(L.lmi_scan_topk(layout.vectors.data_ptr(), layout.n, layout.d, q16.data_ptr(), nq, B,
                                          P[WS_PAIR_LIST], P[WS_ITEMS], P[WS_N_ITEMS], ws.max_items, ws.lmax,
                                          P[WS_PART_SIM], P[WS_PART_ROW], impl | (self.scan_flags << 8),
                                          P[WS_GATE] if self.share_gate else None, st),
                          "lmi_scan_topk")
        mark(3)
        _native.check(L.lmi_merge_topk(P[WS_PART_SIM], P[WS_PART_ROW], P[WS_SLOT_OF_PAIR], order.data_ptr(),
                                       layout.bucket_off.data_ptr(), layout.ids.data_ptr(), P[WS_BUCKET_LISTS], nq, B,
                                       C, ws.lmax, k_out, dists.data_ptr(), nns.data_ptr(),
                                       probe_out.data_ptr() if probe_out is not None else None, st), "lmi_merge_topk")
        mark(4)
        self.last_plan = dict(plan, max_items=ws.max_items, n_items=ws.n_items, counts=ws.counts)
        return dists, nns

    @property
    def launches_per_search(self) -> int:
        """Kernels of this library one 1-level search launches (navigation + inversion + scan + merge)."""
        return self.root_model.launches_per_topn() + 3 + 1 + 1

    def _surface_query_rounding(self):
        """After a host-visible sync: did the fp16 copy of the last f32 query batch change any value?"""
        ws = getattr(self, "_last_ws", None)
        if ws is None or ws.q16 is None or self.allow_fp16_rounding is True or self._warned_rounding:
            return
        n = int(ws.inexact.item())
        if n:
            ws.inexact.zero_()
            self._check_rounding(n, "queries_search")

    # ------------------------------------------------------------------ CUDA graph --
    def search_graphed(self, qn, qs, n_categories: List[int], n_buckets: int, k: int):
        """``search_device`` as ONE CUDA-graph replay, for any batch size: the launches of a search
        (navigation, inversion, scan, merge) are captured once per (batch shape, budget) with static
        input/output buffers -- the only shape-dependent data, the work-item list and its length, is
        produced on the device -- so a search costs one graph launch on the host.  ``qn`` / ``qs`` may be
        host (numpy / pinned torch) or device matrices; they are copied into the static inputs.
        Stage timings come from event-record nodes inside the graph, i.e. they are THIS replay's.
        Returns ``(dists, nns, events)`` -- the static output tensors, valid until the next replay of
        the same shape, and the five stage events (read them after synchronising)."""
        shared = qs is qn

        def as_tensor(x):
            if isinstance(x, torch.Tensor):
                t = x
            else:
                a = np.asarray(x)
                if a.dtype not in (np.float16, np.float32):
                    a = a.astype(np.float32)  # model.py:104
                t = torch.from_numpy(np.ascontiguousarray(a))
            return t if t.dtype in (torch.float16, torch.float32) else t.float()  # as to_device_matrix does

        tn = as_tensor(qn)
        ts = tn if shared else as_tensor(qs)
        # the key holds nothing but shapes, dtypes and the budget: inputs -- host or device -- are always copied
        # into the entry's static device buffers (a device-to-device copy of the batch costs microseconds next
        # to the scan), so a caller that hands over a freshly allocated device tensor per call still replays
        # the one captured graph instead of re-capturing on every new address
        key = (tuple(tn.shape), tn.dtype, tuple(ts.shape), ts.dtype, shared, tuple(n_categories), int(n_buckets),
               int(k), id(self._layout), self.scan_impl)
        entry = self._graphs.get(key)
        if entry is None:
            if len(self._graphs) >= GRAPH_CACHE:
                self._graphs.popitem(last=False)  # least recently used
            sn = torch.empty(tn.shape, dtype=tn.dtype, device=self.device)
            ss = sn if shared else torch.empty(ts.shape, dtype=ts.dtype, device=self.device)
            sn.copy_(tn, non_blocking=True)
            if not shared:
                ss.copy_(ts, non_blocking=True)
            # eager run first: fills every lazy cache (weight splits, workspaces, the tensor-map driver entry
            # point, the navigation tree) outside the capture
            d, i = self.search_device(sn, ss, n_categories, n_buckets, k)
            torch.cuda.current_stream().synchronize()
            ev = [torch.cuda.Event(enable_timing=True, external=True) for _ in range(5)]
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                d, i = self.search_device(sn, ss, n_categories, n_buckets, k, events=ev, out=(d, i))
            # the graph holds raw pointers into the batch workspace and the navigation workspace: keep both alive
            # with the entry, whatever the workspace caches evict
            entry = dict(graph=graph, qn=sn, qs=ss, dists=d, nns=i, events=ev, order=self.last_bucket_order,
                         plan=self.last_plan, keep=(getattr(self, "_last_ws", None),
                                                    dict(self.root_model._fused_ws)))
            self._graphs[key] = entry
        else:
            self._graphs.move_to_end(key)
            if entry["qn"] is not tn:  # (a caller may pass the static buffer itself: nothing to copy then)
                entry["qn"].copy_(tn, non_blocking=True)
            if not shared and entry["qs"] is not ts:
                entry["qs"].copy_(ts, non_blocking=True)
        entry["graph"].replay()
        self.last_bucket_order, self.last_plan = entry["order"], entry["plan"]
        return entry["dists"], entry["nns"], entry["events"]

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

        if len(queries_search) > 0 and self.use_graphs:
            # one graph replay per search; the stage split is read from event nodes inside the graph
            dists_d, nns_d, ev = self.search_graphed(
                queries_navigation, queries_navigation if queries_search is queries_navigation else queries_search,
                n_categories, n_buckets, k)
        else:
            qn = to_device_matrix(queries_navigation, self.device)
            qs = qn if queries_search is queries_navigation else to_device_matrix(queries_search, self.device)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
            dists_d, nns_d = self.search_device(qn, qs, n_categories, n_buckets, k, events=ev)
        dists, nns = results_to_host(self, dists_d, nns_d)
        stage = [ev[j].elapsed_time(ev[j + 1]) / 1e3 for j in range(4)]
        self._surface_query_rounding()

        measured_time["inference"] = stage[0]
        measured_time["search_within_buckets"] = stage[1] + stage[2] + stage[3]
        measured_time["seq_search"] = stage[2]
        measured_time["sort"] = stage[3]
        measured_time["search"] = time.time() - s
        return dists, nns, measured_time
