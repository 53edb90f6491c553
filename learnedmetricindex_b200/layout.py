"""Bucket-major device layout of the index.

The reference has no bucket storage format: ``data_prediction`` (i64[n, L]) plus the
DataFrame index *are* the layout, re-derived per round by a groupby and per-bucket
``.loc`` copies (search/li/LearnedIndex.py:101-104, :350, :357).  Here the rows are
permuted once into

* ``vectors``     fp16[n, d]   rows grouped by bucket (ascending bucket id), original
                               order kept inside a bucket (what groupby yields);
* ``bucket_off``  int64[C+1]   row range of bucket b;
* ``ids``         uint32[n]    DataFrame label of each row (1-based by the driver's
                               convention, search/search.py:190-191), stored as int32 bits;

all resident in HBM (10M x 768 fp16 = 15.4 GB of the B200's 180 GB).
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np
import torch

from . import _native


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


SCAN_K_MULTIPLE = 64  # lmi_scan_topk streams K in 128-byte (64 x fp16) swizzled boxes


def padded_dim(d: int) -> int:
    return -(-int(d) // SCAN_K_MULTIPLE) * SCAN_K_MULTIPLE


class DeviceBucketLayout:
    def __init__(self, vectors: torch.Tensor, bucket_off: torch.Tensor, ids: torch.Tensor, bucket_paths=None,
                 d_data: Optional[int] = None, inexact_values: int = 0):
        assert vectors.is_cuda and vectors.dtype == torch.float16 and vectors.is_contiguous()
        assert bucket_off.dtype == torch.int64 and ids.dtype == torch.int32
        self.vectors = vectors
        self.bucket_off = bucket_off.to(vectors.device)
        self.bucket_off_host = bucket_off.cpu().numpy()
        self.ids = ids
        self.n, self.d = vectors.shape  # d: row pitch (padded to a multiple of 64 with zero columns)
        self.d_data = int(d_data) if d_data is not None else self.d  # the data's own dimensionality
        self.n_buckets = self.bucket_off_host.shape[0] - 1
        sizes = np.diff(self.bucket_off_host)
        self.max_bucket = int(sizes.max()) if sizes.size else 0
        self.n_nonempty = int((sizes > 0).sum())  # buckets holding rows HERE (a shard owns only some)
        # values the f32 -> fp16 conversion changed while this layout was built (0 for fp16 sources)
        self.inexact_values = int(inexact_values)
        self.bucket_paths = bucket_paths  # multi-level: bucket id -> path tuple
        self.device = vectors.device

    # ------------------------------------------------------------------------------
    @classmethod
    def from_device(cls, x: torch.Tensor, labels: torch.Tensor, n_buckets: int,
                    ids: Optional[torch.Tensor] = None, chunk: int = 1 << 20, bucket_paths=None,
                    free_input: bool = False) -> "DeviceBucketLayout":
        """``x`` (fp16/f32 [n,d]) and bucket ids ``labels`` (int64[n]) already on the GPU.  A row whose
        label is negative has no bucket and is dropped (the reference keeps such rows under an
        ``EMPTY_VALUE`` path that no query can reach, LearnedIndexBuilder.py:67-76)."""
        L = _native.lib()
        dev = x.device
        n, d = x.shape
        labels = labels.to(dev, torch.int64).reshape(-1)
        assert labels.shape[0] == n
        assert int(labels.max().item()) < n_buckets if n else True, "bucket label out of range"
        keep = labels >= 0
        counts = torch.bincount(labels[keep], minlength=n_buckets)[:n_buckets]
        off = torch.zeros(n_buckets + 1, dtype=torch.int64, device=dev)
        off[1:] = torch.cumsum(counts, 0)
        n_kept = int(off[-1].item())
        # stable sort by bucket -> destination row of every source row (dropped rows sort last: dest -1)
        order = torch.sort(torch.where(keep, labels, torch.full_like(labels, n_buckets)), stable=True).indices
        dest = torch.empty(n, dtype=torch.int64, device=dev)
        dest[order] = torch.arange(n, dtype=torch.int64, device=dev)
        dest = torch.where(keep, dest, torch.full_like(dest, -1))
        if ids is None:
            ids = torch.arange(1, n + 1, dtype=torch.int64, device=dev)
        ids_sorted = ids.to(dev, torch.int64)[order][:n_kept]
        ids32 = (ids_sorted & 0xFFFFFFFF).to(torch.int64)
        ids32 = torch.where(ids32 >= 2 ** 31, ids32 - 2 ** 32, ids32).to(torch.int32).contiguous()
        del order
        ld = padded_dim(d)
        out = torch.empty((max(n_kept, 1), ld), dtype=torch.float16, device=dev)
        inexact = torch.zeros(1, dtype=torch.int64, device=dev)
        x = x.contiguous()
        for s in range(0, n, chunk):
            m = min(chunk, n - s)
            _native.check(L.lmi_scatter_rows(x[s : s + m].data_ptr(), int(x.dtype == torch.float16),
                                             dest[s : s + m].data_ptr(), m, d, out.data_ptr(), ld,
                                             inexact.data_ptr(), _stream_ptr()),
                          "lmi_scatter_rows")
        torch.cuda.current_stream().synchronize()
        return cls(out[:n_kept] if n_kept else out[:0], off, ids32, bucket_paths, d_data=d,
                   inexact_values=int(inexact.item()))

    @classmethod
    def from_host(cls, data_search, labels: np.ndarray, n_buckets: int, ids: Optional[np.ndarray] = None,
                  device: Optional[torch.device] = None, chunk_rows: int = 1 << 18, bucket_paths=None
                  ) -> "DeviceBucketLayout":
        """Streams a host matrix (ndarray or DataFrame, fp16/f32) to the GPU in chunks and
        permutes it there; the host never holds a second copy."""
        L = _native.lib()
        dev = device or torch.device("cuda", torch.cuda.current_device())
        if hasattr(data_search, "to_numpy"):
            if ids is None:
                ids = np.asarray(data_search.index)
            mat = data_search.to_numpy(copy=False) if hasattr(data_search, "to_numpy") else np.asarray(data_search)
        else:
            mat = np.asarray(data_search)
        n, d = mat.shape
        labels = np.asarray(labels).reshape(-1).astype(np.int64)
        assert labels.shape[0] == n
        if (labels < 0).any() or (labels >= n_buckets).any():
            raise ValueError("bucket label out of range")
        if ids is None:
            ids = np.arange(1, n + 1)
        ids = np.asarray(ids).astype(np.int64)
        order = np.argsort(labels, kind="stable")
        counts = np.bincount(labels, minlength=n_buckets)
        off = np.zeros(n_buckets + 1, dtype=np.int64)
        off[1:] = np.cumsum(counts)
        dest = np.empty(n, dtype=np.int64)
        dest[order] = np.arange(n, dtype=np.int64)
        ids32 = torch.from_numpy((ids[order] & 0xFFFFFFFF).astype(np.uint32).view(np.int32).copy()).to(dev)
        ld = padded_dim(d)
        out = torch.empty((max(n, 1), ld), dtype=torch.float16, device=dev)
        inexact = torch.zeros(1, dtype=torch.int64, device=dev)
        is_f16 = mat.dtype == np.float16
        if mat.dtype not in (np.float16, np.float32):
            mat = mat.astype(np.float32)
        for s in range(0, n, chunk_rows):
            m = min(chunk_rows, n - s)
            blk = torch.from_numpy(np.ascontiguousarray(mat[s : s + m])).to(dev)
            dst = torch.from_numpy(dest[s : s + m]).to(dev)
            _native.check(L.lmi_scatter_rows(blk.data_ptr(), int(is_f16), dst.data_ptr(), m, d, out.data_ptr(),
                                             ld, inexact.data_ptr(), _stream_ptr()), "lmi_scatter_rows")
            torch.cuda.current_stream().synchronize()
        return cls(out[:n], torch.from_numpy(off), ids32, bucket_paths, d_data=d, inexact_values=int(inexact.item()))

    # ------------------------------------------------------------------------------
    def restrict_to_buckets(self, mine: Sequence[int]) -> "DeviceBucketLayout":
        """Shard: keep only the rows of buckets ``mine`` (others become empty) -- the
        per-GPU layout when buckets are sharded across ranks."""
        mine = sorted(int(b) for b in mine)
        off = self.bucket_off_host
        sizes = np.zeros(self.n_buckets, dtype=np.int64)
        sizes[mine] = np.diff(off)[mine]
        new_off = np.zeros(self.n_buckets + 1, dtype=np.int64)
        new_off[1:] = np.cumsum(sizes)
        rows = [torch.arange(off[b], off[b + 1], device=self.device) for b in mine if off[b + 1] > off[b]]
        sel = torch.cat(rows) if rows else torch.zeros(0, dtype=torch.int64, device=self.device)
        vec = self.vectors[sel].contiguous() if sel.numel() else self.vectors[:0]
        if vec.shape[0] == 0:
            vec = torch.zeros((1, self.d), dtype=torch.float16, device=self.device)[:0]
        return DeviceBucketLayout(vec, torch.from_numpy(new_off), self.ids[sel].contiguous(), self.bucket_paths,
                                  d_data=self.d_data, inexact_values=self.inexact_values)

    def restrict_to_slice(self, rank: int, world: int, shares=None) -> "DeviceBucketLayout":
        """Shard: keep rows ``[s_b*rank//world, s_b*(rank+1)//world)`` of EVERY bucket -- every rank
        then scans the same buckets for the same probes with 1/world of the rows (perfectly balanced).
        ``shares[r]``: relative scan throughput of rank r; its slice is then that fraction of every bucket.
        Valid for ``k <= 10``: an object ranked beyond 10 inside its bucket cannot reach the overall
        top 10, so merging the ranks' lists gives the unsharded result (filler ids of undersized
        buckets excepted)."""
        from .sharded import slice_bounds

        off = self.bucket_off_host
        sizes = np.diff(off)
        lo_all, hi_all = slice_bounds(sizes, world, shares)
        lo = off[:-1] + lo_all[rank]
        hi = off[:-1] + hi_all[rank]
        new_off = np.zeros(self.n_buckets + 1, dtype=np.int64)
        new_off[1:] = np.cumsum(hi - lo)
        rows = [torch.arange(a, b, device=self.device) for a, b in zip(lo.tolist(), hi.tolist()) if b > a]
        sel = torch.cat(rows) if rows else torch.zeros(0, dtype=torch.int64, device=self.device)
        vec = self.vectors[sel].contiguous() if sel.numel() else \
            torch.zeros((1, self.d), dtype=torch.float16, device=self.device)[:0]
        return DeviceBucketLayout(vec, torch.from_numpy(new_off), self.ids[sel].contiguous(), self.bucket_paths,
                                  d_data=self.d_data, inexact_values=self.inexact_values)
