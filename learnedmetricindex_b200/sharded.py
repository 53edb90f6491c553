"""Bucket-sharded search over the GPUs of one box (one process per GPU).

The reference is single-process; this is the multi-GPU form BASELINE.json's
north_star asks for.  Buckets are independent units, so whole buckets are assigned to
ranks; MLP weights and the query batch are replicated; every rank runs the single-GPU
path (``LearnedIndex.search_device``) over the buckets it owns -- a probed bucket that
lives elsewhere is an empty bucket locally and yields the reference's "no candidate"
filler (id 0, +inf; LearnedIndex.py:340-341) -- and the only exchange step moves the per-rank
``[nq, k]`` results (``nq*k*14`` bytes per rank with the probe ranks; 1.4 MB at 10k x 10).

Two exchange paths, bit-identical results:
  * ``exchange="peer"`` (the default where symmetric memory works): every rank writes its packed block
    ``[dists f64 | nns u32 | probe u16]`` into symmetric memory, a device-side barrier orders the writes,
    and ``lmi_merge_ranks_peer`` reads the peers' blocks over NVLink INSIDE the merge kernel -- no
    collective-library call on the data path.  Navigation is query-sharded the same way (each rank infers
    1/world of the batch, ``lmi_gather_peer`` collects the bucket order), and ``search_graphed`` replays
    the whole sharded search as one CUDA graph.
  * ``exchange="nccl"`` (fallback, agreed on by all ranks together): ONE all-gather of the packed blocks
    followed by ``lmi_merge_ranks`` on every rank.

The local merge kernel writes straight into this rank's send block, so there is no staging copy.

Host-side logic (partitioning, buffer packing, gather, merge order) is backend-agnostic:
``tests/test_sharded_gloo.py`` runs it with world_size 2 over gloo on CPU, injecting a
numpy merge in place of the CUDA kernel.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist

SHARDED_GRAPH_CACHE = 8  # captured (shape, budget) entries kept per sharded index, least recently used goes first


def partition_buckets(sizes: Sequence[int], world: int, weights: Optional[Sequence[float]] = None,
                      speeds: Optional[Sequence[float]] = None) -> List[List[int]]:
    """Longest-processing-time greedy assignment of whole buckets to ``world`` ranks.

    ``sizes[b]`` = rows of bucket b; ``weights[b]`` (optional) = expected probes of bucket b,
    so the balanced quantity is scan work ``s_b * w_b`` (HBM bytes when all weights are 1).
    ``speeds[r]`` (optional) = relative scan throughput of rank r -- the GPUs of one box differ by several
    percent under their power caps -- a bucket then goes to the rank that would FINISH it earliest.
    Deterministic: ties go to the lower bucket id / lower rank, so every rank derives the
    same map without communicating."""
    sizes = np.asarray(sizes, dtype=np.float64)
    cost = sizes if weights is None else sizes * np.asarray(weights, dtype=np.float64)
    speed = np.ones(world) if speeds is None else np.asarray(speeds, dtype=np.float64)
    assert speed.shape == (world,) and (speed > 0).all()
    order = sorted(range(len(sizes)), key=lambda b: (-cost[b], b))
    load = [0.0] * world
    mine: List[List[int]] = [[] for _ in range(world)]
    for b in order:
        r = min(range(world), key=lambda w: ((load[w] + float(cost[b])) / speed[w] if speeds is not None else load[w], w))
        mine[r].append(b)
        load[r] += float(cost[b])
    return [sorted(m) for m in mine]


def slice_bounds(sizes: Sequence[int], world: int, shares: Optional[Sequence[float]] = None):
    """Row slice ``[lo_b, hi_b)`` of every bucket for each rank: ``(lo[world, C], hi[world, C])``.  Equal slices by
    default; ``shares[r]`` (relative scan throughput of rank r) makes rank r's slice that fraction of every bucket.
    The slices are contiguous and in rank order, which the cross-GPU merge's tie order relies on."""
    sizes = np.asarray(sizes, dtype=np.int64)
    if shares is None:
        cuts = [sizes * r // world for r in range(world + 1)]
    else:
        sh = np.asarray(shares, dtype=np.float64)
        assert sh.shape == (world,) and (sh > 0).all()
        cum = np.concatenate(([0.0], np.cumsum(sh) / sh.sum()))
        cuts = [np.floor(sizes * c + 1e-9).astype(np.int64) for c in cum]
        cuts[-1] = sizes.copy()
    lo = np.stack(cuts[:-1])
    hi = np.stack(cuts[1:])
    return lo, np.maximum(hi, lo)


def owner_of_bucket(parts: List[List[int]], n_buckets: int) -> np.ndarray:
    own = np.full(n_buckets, -1, dtype=np.int64)
    for r, bs in enumerate(parts):
        own[bs] = r
    assert (own >= 0).all()
    return own


class GatherBuffers:
    """Packed all-gather buffers: per rank ``[nq*k f64][nq*k u32][nq*k u16]`` padded to 16 bytes -- distances,
    ids, and the probe rank of every entry (``lmi_merge_topk``'s ``probe_rank``), which the rank merge uses to
    order equal distances exactly as the unsharded search does (zeros = absent: ties keep the lower rank)."""

    def __init__(self, nq: int, k: int, world: int, rank: int, device):
        self.nq, self.k, self.world, self.rank = nq, k, world, rank
        self.bytes_d = nq * k * 8
        self.bytes_i = nq * k * 4
        self.bytes_p = nq * k * 2
        self.off_p = self.bytes_d + self.bytes_i
        self.per_rank = -(-(self.off_p + self.bytes_p) // 16) * 16
        self.send = torch.zeros(self.per_rank, dtype=torch.uint8, device=device)
        self.recv = torch.zeros(self.per_rank * world, dtype=torch.uint8, device=device)

    def local_views(self) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        d = self.send[: self.bytes_d].view(torch.float64).view(self.nq, self.k)
        i = self.send[self.bytes_d : self.bytes_d + self.bytes_i].view(torch.int32).view(self.nq, self.k)
        p = self.send[self.off_p : self.off_p + self.bytes_p].view(torch.int16).view(self.nq, self.k)
        return d, i, p

    def rank_views(self, w: int) -> Tuple[torch.Tensor, torch.Tensor]:
        base = w * self.per_rank
        d = self.recv[base : base + self.bytes_d].view(torch.float64).view(self.nq, self.k)
        i = self.recv[base + self.bytes_d : base + self.bytes_d + self.bytes_i].view(torch.int32).view(self.nq, self.k)
        return d, i

    @property
    def stride_d(self) -> int:  # elements between consecutive ranks' dist blocks
        return self.per_rank // 8

    @property
    def stride_i(self) -> int:
        return self.per_rank // 4

    @property
    def stride_p(self) -> int:
        return self.per_rank // 2


class PeerBuffers:
    """The exchange step over PEER MEMORY instead of a collective: every rank's packed result block
    ``[nq*k f64 | nq*k u32]`` lives in a symmetric-memory buffer (``torch.distributed.
    _symmetric_memory``: the same allocation mapped into every process of the box over NVLink), the
    local merge kernel writes straight into it, a cross-rank barrier orders the step, and
    ``lmi_merge_ranks_peer`` reads all peers' blocks through NVSwitch inside the merge kernel --
    all-gather and k-way merge in one launch.  Two blocks alternate so a rank may start the next
    search while slower peers still read its previous results (one barrier per search)."""

    def __init__(self, nq: int, k: int, world: int, rank: int, device, group=None, n_blocks: int = 2):
        import torch.distributed._symmetric_memory as symm_mem

        self.nq, self.k, self.world, self.rank = nq, k, world, rank
        self.bytes_d = nq * k * 8
        self.bytes_i = nq * k * 4
        self.bytes_p = nq * k * 2  # probe rank of every entry: see GatherBuffers
        self.off_p = self.bytes_d + self.bytes_i
        self.per_rank = -(-(self.off_p + self.bytes_p) // 16) * 16
        self.blocks, self.handles = [], []
        for _ in range(n_blocks):  # a captured graph owns ONE block (fixed pointers); two graphs alternate instead
            t = symm_mem.empty(self.per_rank, dtype=torch.uint8, device=device)
            t.zero_()
            self.handles.append(symm_mem.rendezvous(t, group if group is not None else dist.group.WORLD))
            self.blocks.append(t)
        self.turn = 0

    def local_views(self) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        blk = self.blocks[self.turn]
        d = blk[: self.bytes_d].view(torch.float64).view(self.nq, self.k)
        i = blk[self.bytes_d : self.bytes_d + self.bytes_i].view(torch.int32).view(self.nq, self.k)
        p = blk[self.off_p : self.off_p + self.bytes_p].view(torch.int16).view(self.nq, self.k)
        return d, i, p

    def exchange_and_merge(self, out=None):
        from . import _native

        hdl = self.handles[self.turn]
        dev = self.blocks[0].device
        if out is not None:
            out_d, out_i = out
        else:
            out_d = torch.empty((self.nq, self.k), dtype=torch.float64, device=dev)
            out_i = torch.empty((self.nq, self.k), dtype=torch.int32, device=dev)
        hdl.barrier(channel=0)  # every rank's block of this turn is written (device-side, on this stream)
        _native.check(_native.lib().lmi_merge_ranks_peer(hdl.buffer_ptrs_dev, self.bytes_d, self.off_p, self.world,
                                                         self.nq, self.k, out_d.data_ptr(), out_i.data_ptr(),
                                                         torch.cuda.current_stream().cuda_stream),
                      "lmi_merge_ranks_peer")
        self.turn = (self.turn + 1) % len(self.blocks)
        return out_d, out_i


class PeerGather:
    """All-gather of one fixed-size block per rank by loads from peer memory (``lmi_gather_peer``):
    the rank writes its block into a symmetric-memory buffer, a device-side barrier orders the
    step, one kernel copies every peer's block over NVLink.  Two blocks alternate (see
    ``PeerBuffers``)."""

    def __init__(self, bytes_per_rank: int, world: int, device, group=None, n_blocks: int = 2):
        import torch.distributed._symmetric_memory as symm_mem

        self.bytes = -(-max(bytes_per_rank, 16) // 16) * 16
        self.world = world
        self.blocks, self.handles = [], []
        self.recv = None  # n_blocks == 1 (graph capture): ONE receive buffer with a fixed address
        for _ in range(n_blocks):
            t = symm_mem.empty(self.bytes, dtype=torch.uint8, device=device)
            t.zero_()
            self.handles.append(symm_mem.rendezvous(t, group if group is not None else dist.group.WORLD))
            self.blocks.append(t)
        self.turn = 0

    def send_block(self) -> torch.Tensor:
        return self.blocks[self.turn]

    def gather(self) -> torch.Tensor:
        """uint8[world * bytes]: every rank's block, in rank order (a fresh tensor)."""
        from . import _native

        hdl = self.handles[self.turn]
        if len(self.blocks) == 1:
            if self.recv is None:
                self.recv = torch.empty(self.world * self.bytes, dtype=torch.uint8, device=self.blocks[0].device)
            recv = self.recv
        else:
            recv = torch.empty(self.world * self.bytes, dtype=torch.uint8, device=self.blocks[0].device)
        hdl.barrier(channel=0)
        _native.check(_native.lib().lmi_gather_peer(hdl.buffer_ptrs_dev, self.bytes, self.world, recv.data_ptr(),
                                                    torch.cuda.current_stream().cuda_stream), "lmi_gather_peer")
        self.turn = (self.turn + 1) % len(self.blocks)
        return recv


def merge_ranks_device(buf: GatherBuffers, out_d: torch.Tensor, out_i: torch.Tensor) -> None:
    """``lmi_merge_ranks`` over the gathered buffer (CUDA)."""
    from . import _native

    L = _native.lib()
    base = buf.recv.data_ptr()
    _native.check(L.lmi_merge_ranks(base, base + buf.bytes_d, base + buf.off_p, buf.stride_d, buf.stride_i,
                                    buf.stride_p, buf.world, buf.nq, buf.k, out_d.data_ptr(), out_i.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream),
                  "lmi_merge_ranks")


def gather_and_merge(buf: GatherBuffers, group=None,
                     merge_fn: Callable[[GatherBuffers, torch.Tensor, torch.Tensor], None] = merge_ranks_device):
    """The exchange step: all-gather the packed per-rank results, then k-way merge."""
    dist.all_gather_into_tensor(buf.recv, buf.send, group=group)
    out_d = torch.empty((buf.nq, buf.k), dtype=torch.float64, device=buf.send.device)
    out_i = torch.empty((buf.nq, buf.k), dtype=torch.int32, device=buf.send.device)
    merge_fn(buf, out_d, out_i)
    return out_d, out_i


def query_slice(nq: int, world: int, rank: int) -> Tuple[int, int, int]:
    """(start, stop, padded slice length) of the queries whose bucket order ``rank`` computes."""
    per = -(-nq // world) if nq else 0
    start = min(nq, rank * per)
    return start, min(nq, start + per), per


def gather_bucket_order(order_local: torch.Tensor, nq: int, n_probe: int, group=None) -> torch.Tensor:
    """The navigation step sharded by queries: every rank runs the MLP + top-B on its slice only
    and the int32 ``[slice, B]`` results are all-gathered (``nq*B*4`` bytes in total; 160 KB at
    10k x 4) -- every rank needs the whole bucket order to find the probes of its own buckets."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    start, stop, per = query_slice(nq, world, rank)
    send = torch.full((per, n_probe), -1, dtype=torch.int32, device=order_local.device)
    if stop > start:
        send[: stop - start] = order_local
    recv = torch.empty((world * per, n_probe), dtype=torch.int32, device=order_local.device)
    dist.all_gather_into_tensor(recv, send, group=group)
    return recv[:nq].contiguous()


class ShardedLearnedIndex:
    """``LearnedIndex`` whose buckets are spread over the ranks of ``group``.

    ``local`` is this rank's ``LearnedIndex`` with a layout holding only the buckets it
    owns (``DeviceBucketLayout.restrict_to_buckets`` or a layout built from the local rows).
    """

    def __init__(self, local, group=None, shard_navigation: bool = True, exchange: Optional[str] = None):
        import os

        self.local = local
        self.group = group
        self.shard_navigation = shard_navigation  # MLP + top-B on nq/world queries per rank, then all-gather
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        # "peer": results are exchanged by reading peer memory inside the merge kernel (PeerBuffers);
        # "nccl": all-gather + lmi_merge_ranks (GatherBuffers; also what the gloo tests run)
        self.exchange = exchange or os.environ.get("LMI_EXCHANGE", "peer")
        self._bufs = {}
        self._graphs = OrderedDict()
        if self.exchange == "peer" and self.world > 1 and torch.device(local.device).type == "cuda":
            self.exchange = "peer" if self._agree_on_peer_memory(local.device, group) else "nccl"

    def _agree_on_peer_memory(self, device, group) -> bool:
        """Symmetric memory needs P2P mappings between all GPUs of the group.  ``rendezvous`` is itself a
        collective, so the ranks agree in TWO steps: first every rank tests what it can test alone (the
        module imports, an allocation succeeds, every peer GPU is P2P-reachable) and the flags are
        MIN-reduced -- only if all say yes does anybody enter ``rendezvous`` -- then the outcome of the
        rendezvous is MIN-reduced as well.  A rank that cannot take part therefore never leaves the
        others waiting inside a collective; all of them fall back to the NCCL exchange together."""
        def agree(ok: bool) -> bool:
            flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
            return bool(flag.item())

        t = None
        try:
            import torch.distributed._symmetric_memory as symm_mem

            me = torch.device(device).index
            me = torch.cuda.current_device() if me is None else me
            ok = all(torch.cuda.can_device_access_peer(me, o) for o in range(torch.cuda.device_count()) if o != me) \
                if torch.cuda.device_count() >= self.world else True  # ranks on other devices than the visible ones
            t = symm_mem.empty(16, dtype=torch.uint8, device=device) if ok else None
        except Exception:  # noqa: BLE001 -- any failure means "use the collective library"
            ok = False
        if not agree(ok):
            return False
        try:
            symm_mem.rendezvous(t, group if group is not None else dist.group.WORLD)
            ok = True
        except Exception:  # noqa: BLE001
            ok = False
        return agree(ok)

    def _buffers(self, nq: int, k: int, device):
        key = (nq, k)
        if key not in self._bufs:
            if self.exchange == "peer" and torch.device(device).type == "cuda" and self.world > 1:
                self._bufs[key] = PeerBuffers(nq, k, self.world, self.rank, device, self.group)
            else:
                self._bufs[key] = GatherBuffers(nq, k, self.world, self.rank, device)
        return self._bufs[key]

    def search_device(self, qn, qs, n_categories, n_buckets: int, k: int, events=None):
        from .LearnedIndex import PER_BUCKET_K

        nq = qs.shape[0]
        k_out = PER_BUCKET_K if int(n_buckets) == 1 else int(k)
        buf = self._buffers(nq, k_out, qs.device)
        order = None
        if self.shard_navigation and self.world > 1 and nq >= self.world:
            if events is not None:
                events[0].record()
            start, stop, _ = query_slice(nq, self.world, self.rank)
            local_order, _ = self.local._bucket_order_device(qn[start:stop], int(n_buckets), list(n_categories))
            if isinstance(buf, PeerBuffers):
                order = self._gather_rows_peer(local_order, nq)
            else:
                order = gather_bucket_order(local_order, nq, int(n_buckets), self.group)
        self.local.search_device(qn, qs, n_categories, n_buckets, k, events=events, out=buf.local_views(),
                                 bucket_order=order)
        if isinstance(buf, PeerBuffers):
            return buf.exchange_and_merge()
        return gather_and_merge(buf, self.group)

    # ------------------------------------------------------------------ CUDA graph ----------------
    def search_graphed(self, qn, qs, n_categories, n_buckets: int, k: int):
        """The whole sharded search as ONE graph replay per call (peer-memory exchange only): query upload
        gather -> navigation of this rank's query slice -> bucket-order gather -> inversion / scan / merge of
        the local buckets -> device-side barrier -> rank merge reading the peers' blocks.  Host inputs: only
        this rank's 1/world slice crosses PCIe (copied into the graph's symmetric send block before the
        replay), the rest arrives over NVLink inside the graph.  Every (shape, budget) owns TWO captured
        instances that alternate, each with its own symmetric blocks: an instance's blocks are rewritten
        two searches later, when every peer has passed the barrier in between (same argument as the
        double-buffered eager path).  Returns ``(dists, nns, events)`` (static tensors of the instance)."""
        from .LearnedIndex import PER_BUCKET_K
        from .model import to_device_matrix

        assert self.exchange == "peer" and self.world > 1
        shared = qs is qn
        nq = len(qs)
        dev = self.local.device
        host_in = not (isinstance(qn, torch.Tensor) and qn.is_cuda)
        dt = lambda x: (x.dtype if isinstance(x, torch.Tensor) else
                        (torch.float16 if np.asarray(x).dtype == np.float16 else torch.float32))
        if not host_in:
            assert qn.is_contiguous() and qs.is_contiguous()
        # The key must be the same on every rank: creating an instance is a collective (symmetric-memory
        # rendezvous, an eager run with device barriers), so a rank that hit its cache while another missed
        # would leave the job hanging.  It therefore holds shapes, dtypes and the budget only -- never a device
        # address, which is per-rank allocator state; device inputs are copied into a static buffer instead.
        key = (nq, tuple(np.shape(qn)), dt(qn), tuple(np.shape(qs)), dt(qs), shared, host_in, tuple(n_categories),
               int(n_buckets), int(k))
        entry = self._graphs.get(key)
        if entry is None:
            while len(self._graphs) >= SHARDED_GRAPH_CACHE:
                # least recently used first (every rank sees the same call sequence, so all evict the same
                # entry); the graphs go before the workspaces and symmetric blocks they point into
                _, old = self._graphs.popitem(last=False)
                for old_inst in old["instances"]:
                    if old_inst is not None:
                        old_inst.pop("graph", None)
                        old_inst.clear()
            entry = dict(instances=[None, None], turn=0)
            self._graphs[key] = entry
        else:
            self._graphs.move_to_end(key)
        t = entry["turn"]
        entry["turn"] ^= 1
        inst = entry["instances"][t]
        start, stop, per = query_slice(nq, self.world, self.rank)
        k_out = PER_BUCKET_K if int(n_buckets) == 1 else int(k)

        def upload(inst):
            """Stage this call's queries where the graph reads them."""
            for name, src in (("n", qn),) + ((("s", qs),) if not shared else ()):
                if host_in:  # H2D of the 1/world slice only, straight into the symmetric send block
                    part = src[start:stop]
                    if not isinstance(part, torch.Tensor):
                        part = torch.from_numpy(np.ascontiguousarray(part))
                    dst = inst["send_" + name]
                    dst[: part.shape[0]].copy_(part.to(dst.dtype), non_blocking=True)
                else:  # device inputs: one device-to-device copy into the instance's static buffer
                    inst["full_" + name].copy_(src, non_blocking=True)

        if inst is None:
            inst = {}
            dtypes = {"n": dt(qn), "s": dt(qs)}
            shapes = {"n": tuple(np.shape(qn)), "s": tuple(np.shape(qs))}
            for name in ("n",) + (("s",) if not shared else ()):
                if host_in:
                    row_bytes = shapes[name][1] * torch.empty(0, dtype=dtypes[name]).element_size()
                    pg = PeerGather(per * row_bytes, self.world, dev, self.group, n_blocks=1)
                    inst["pg_" + name] = pg
                    inst["send_" + name] = pg.send_block()[: per * row_bytes].view(dtypes[name]).view(per, shapes[name][1])
                else:
                    inst["full_" + name] = torch.empty(shapes[name], dtype=dtypes[name], device=dev)
            inst["pg_order"] = PeerGather(per * int(n_buckets) * 4, self.world, dev, self.group, n_blocks=1)
            inst["res"] = PeerBuffers(nq, k_out, self.world, self.rank, dev, self.group, n_blocks=1)
            inst["out"] = (torch.empty((nq, k_out), dtype=torch.float64, device=dev),
                           torch.empty((nq, k_out), dtype=torch.int32, device=dev))

            def body(events):
                full = {}
                for name in ("n",) + (("s",) if not shared else ()):
                    if host_in:
                        pg = inst["pg_" + name]
                        rb = inst["send_" + name].shape[1] * inst["send_" + name].element_size()
                        raw = pg.gather().view(self.world, pg.bytes)[:, : per * rb]
                        full[name] = raw.reshape(-1).view(dtypes[name]).view(self.world * per, shapes[name][1])[:nq]
                    else:
                        full[name] = inst["full_" + name]
                fn = full["n"]
                fs = fn if shared else full["s"]
                if events is not None:
                    events[0].record()
                local_order, _ = self.local._bucket_order_device(fn[start:stop], int(n_buckets), list(n_categories))
                pg = inst["pg_order"]
                blk = pg.send_block()[: per * int(n_buckets) * 4].view(torch.int32).view(per, int(n_buckets))
                blk[: local_order.shape[0]].copy_(local_order)
                order = pg.gather().view(self.world, pg.bytes)[:, : per * int(n_buckets) * 4].reshape(-1) \
                    .view(torch.int32).view(self.world * per, int(n_buckets))[:nq]
                self.local.search_device(fn, fs, n_categories, n_buckets, k, events=events,
                                         out=inst["res"].local_views(), bucket_order=order.contiguous())
                inst["res"].exchange_and_merge(out=inst["out"])

            upload(inst)
            body(None)  # eager run: every rank executes the same barriers, lazy caches fill outside the capture
            torch.cuda.current_stream().synchronize()
            # the graph holds raw pointers into the local index's batch workspace and navigation workspace; both
            # caches evict on their own schedule, so the instance keeps them alive for as long as its graph
            inst["keep"] = (getattr(self.local, "_last_ws", None), dict(self.local.root_model._fused_ws))
            inst["events"] = [torch.cuda.Event(enable_timing=True, external=True) for _ in range(5)]
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                body(inst["events"])
            inst["graph"] = graph
            entry["instances"][t] = inst
        upload(inst)
        inst["graph"].replay()
        return inst["out"][0], inst["out"][1], inst["events"]

    def _gather_rows_peer(self, part: torch.Tensor, n_rows: int) -> torch.Tensor:
        """All-gather of row blocks over peer memory: ``part`` holds this rank's rows
        ``query_slice(n_rows, world, rank)`` of a matrix; returns all ``n_rows`` rows."""
        _, _, per = query_slice(n_rows, self.world, self.rank)
        row_bytes = part.shape[1] * part.element_size()
        key = ("rows", per, part.shape[1], part.dtype)
        if key not in self._bufs:
            self._bufs[key] = PeerGather(per * row_bytes, self.world, part.device, self.group)
        pg = self._bufs[key]
        blk = pg.send_block()[: per * row_bytes].view(part.dtype).view(per, part.shape[1])
        blk[: part.shape[0]] = part
        full = pg.gather().view(self.world, pg.bytes)[:, : per * row_bytes].reshape(-1)
        return full.view(part.dtype).view(self.world * per, part.shape[1])[:n_rows].contiguous()

    def _upload_sharded(self, host_matrix, dev) -> torch.Tensor:
        """Host queries -> device on every rank: each rank copies only its 1/world slice over PCIe and
        the slices are all-gathered over NVLink (every rank is handed the same host batch)."""
        from .model import to_device_matrix

        nq = len(host_matrix)
        if self.world == 1 or nq < self.world:
            return to_device_matrix(host_matrix, dev)
        start, stop, per = query_slice(nq, self.world, self.rank)
        part = to_device_matrix(host_matrix[start:stop], dev)
        if self.exchange == "peer" and torch.device(dev).type == "cuda":
            return self._gather_rows_peer(part, nq)
        send = torch.zeros((per, part.shape[1]), dtype=part.dtype, device=dev)
        send[: stop - start] = part
        recv = torch.empty((self.world * per, part.shape[1]), dtype=part.dtype, device=dev)
        dist.all_gather_into_tensor(recv, send, group=self.group)
        return recv[:nq]

    def search(self, data_navigation, queries_navigation, data_search, queries_search, data_prediction,
               n_categories, n_buckets: int = 1, k: int = 10):
        """Same signature as ``LearnedIndex.search``; the layout must already be attached
        (each rank holds only its shard, so the host frames are not used)."""
        import time
        from collections import defaultdict

        from .model import to_device_matrix

        t0 = time.time()
        dev = self.local.device
        times = defaultdict(float)
        if (self.exchange == "peer" and self.world > 1 and torch.device(dev).type == "cuda" and self.local.use_graphs
                and self.shard_navigation and len(queries_search) >= self.world):
            d, i, ev = self.search_graphed(queries_navigation,
                                           queries_navigation if queries_search is queries_navigation else queries_search,
                                           list(n_categories), n_buckets, k)
            from .LearnedIndex import results_to_host

            dists, nns = results_to_host(self, d, i)
            stage = [ev[j].elapsed_time(ev[j + 1]) / 1e3 for j in range(4)]
            times.update(inference=stage[0], search_within_buckets=stage[1] + stage[2] + stage[3],
                         seq_search=stage[2], sort=stage[3])
        else:
            qn = self._upload_sharded(queries_navigation, dev)
            qs = qn if queries_search is queries_navigation else self._upload_sharded(queries_search, dev)
            d, i = self.search_device(qn, qs, list(n_categories), n_buckets, k)
            dists = d.cpu().numpy()
            nns = i.cpu().numpy().view(np.uint32)
        times["search"] = time.time() - t0
        return dists, nns, times
