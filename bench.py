#!/usr/bin/env python
"""Benchmark of the LMI query path (``LearnedIndex.search``) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload b|a|c|ml]

A *step* is one pass of the hot path over one batch: MLP -> top-B buckets -> inversion ->
bucket scan with fused top-10 -> merge (-> all-gather + rank merge when N > 1) for all
``nq`` queries.  Metric: QPS (= nq / step time) at recall@10 >= 0.90.

Workloads (BASELINE.json ``configs``):
  b (default)  10M x 768 fp16, 120 buckets, MLP-4 (768->512->120), 10k queries, k=10, B=4  -- headline
  a            100k x 768, otherwise as b (the reference's CPU-runnable size)
  c            100M x 768 (153.6 GB: fits one B200; the sharded configuration for N = 2/4/8)
  ml           1M x 768, 2-level index [10, 10] with a 96-d PCA navigation representation, 10 of 100
               buckets (single GPU; see run_multilevel)

``--gpus N`` (launched under torchrun, one rank per GPU) shards the SAME workload over the N ranks
("strong" scaling): whole buckets per rank (LPT on rows^2), or -- when a 1/N slice of a bucket still
holds several scan ranges (``--shard``, automatic for 100M) -- an equal row slice of every bucket.
Every rank scans its rows for the whole query batch; the exchange step reads the peers' [nq,k]
lists out of peer memory inside the merge kernel (``LMI_EXCHANGE=nccl``: all-gather + merge).

Printed JSON (one line, rank 0): ``value`` = device-timed QPS with inputs resident in HBM;
``e2e`` = QPS through ``LearnedIndex.search`` with HOST queries (pinned) -> H2D -> kernels ->
D2H of (dists, nns); ``roofline`` for the scan kernel (CUDA events around that launch,
algorithmic flops/bytes as in DESIGN.md); ``cpu_baseline`` = the CPU oracle
(oracle/lmi_oracle.py, a restatement of the reference's search) timed on this box's host
cores on a bounded query sample.  ``--impl reference`` times only that CPU path.
Extra keys: ``kernel_ms`` (per-stage CUDA-event means; ``scan_per_rank`` when N > 1),
``host_enqueue_ms_per_step``, ``smallest_budget_with_recall_0.90`` (the metric's other reading: QPS
at the smallest probed-bucket budget that still reaches recall 0.90; ``value`` stays on the README's
budget of 4 buckets).

Synthetic data: there is no network for the SISAP LAION files, so vectors are drawn from a
768-d mixture (learnedmetricindex_b200/synthetic.py), L2-normalised, fp16; the index is a
stand-in for LearnedIndexBuilder.build (k-means labels -> trained MLP-4 -> bucket = argmax).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# torchrun exports OMP_NUM_THREADS=1 to every rank; the reference arm is a CPU measurement on ALL host
# cores (rank 0 alone works there), so that default must not leak into it
if "--impl" in sys.argv and "reference" in sys.argv and os.environ.get("OMP_NUM_THREADS") == "1":
    del os.environ["OMP_NUM_THREADS"]

import numpy as np
import torch
import torch.distributed as dist

WORKLOADS = {
    "a": dict(n=100_000, nq=10_000, buckets=120, name="100k x 768 fp16, 120 buckets, MLP-4, 10k queries, k=10"),
    "b": dict(n=10_000_000, nq=10_000, buckets=120, name="10M x 768 fp16, 120 buckets, MLP-4, 10k queries, k=10"),
    "c": dict(n=100_000_000, nq=10_000, buckets=120, name="100M x 768 fp16, 120 buckets, MLP-4, 10k queries, k=10"),
    # 2-level index (the reference CLI's default tree shape): see run_multilevel; --probe defaults to 10 of 100
    "ml": dict(n=1_000_000, nq=10_000, buckets=100, name="1M x 768 fp16, n_categories [10, 10], 10k queries, k=10"),
}
D = 768
K_NN = 10
CHUNK = 1 << 20
TRAIN_ROWS = 200_000


def log(*a):
    if int(os.environ.get("RANK", "0")) == 0:
        print("[bench]", *a, file=sys.stderr, flush=True)


# --------------------------------------------------------------------------------------
# clocks (sampled DURING the timed regions)
# --------------------------------------------------------------------------------------
_REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
            0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x10: "sync_boost"}


class ClockSampler:
    def __init__(self, device_index: int, period_s: float = 0.01):
        self.samples = []  # (t, sm_mhz, reasons_mask, power_w)
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        self._h = None
        self.period = period_s
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nv = pynvml
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            if not uuid.startswith("GPU-"):
                uuid = "GPU-" + uuid
            try:
                self._h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if hasattr(uuid, "encode") else uuid)
            except Exception:
                self._h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception as e:  # no NVML: report that instead of inventing clocks
            self.error = repr(e)

    def _loop(self):
        nv, h = self._nv, self._h
        while not self._stop.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                except Exception:
                    pw = float("nan")
                self.samples.append((time.time(), float(mhz), int(mask), pw))
            except Exception:
                pass
            self._stop.wait(self.period)

    def start(self):
        if self._h is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2)

    def summary(self, windows):
        """Median SM clock and the union of throttle reasons over samples inside ``windows``."""
        if self._h is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "NVML unavailable"}
        sel = [s for s in self.samples if any(a <= s[0] <= b for a, b in windows)]
        if not sel:
            sel = self.samples
        mhz = sorted(s[1] for s in sel)
        mask = 0
        for s in sel:
            mask |= s[2]
        reasons = [name for bit, name in _REASONS.items() if mask & bit]
        pw = [s[3] for s in sel if s[3] == s[3]]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(sel), "power_w_max": max(pw) if pw else None}


# --------------------------------------------------------------------------------------
# workload construction (untimed setup; torch ops + this repo's layout kernels)
# --------------------------------------------------------------------------------------
def chunk_plan(n):
    return [(s, min(CHUNK, n - s)) for s in range(0, n, CHUNK)]


def make_chunk(ci, m, spec, device):
    from learnedmetricindex_b200 import synthetic as syn

    return syn.make_vectors(m, spec, device=device, stream=1000 + ci, chunk=CHUNK)


STANDIN_ASSET = os.path.join(ROOT, "bench_assets", "standin_mlp4_768_120.npz")


def load_standin_asset(spec, C, device):
    """The committed stand-in weights (tools/make_standin_asset.py: trained once, on the CPU, from a draw of the
    default mixture) when the workload is the default one -- so every run, box, --gpus N and arm searches the
    SAME index; None otherwise."""
    try:
        z = np.load(STANDIN_ASSET)
    except Exception:
        return None
    c, d, _, _, centres, latent = (int(v) for v in z["meta"])
    if (c, d, centres, latent) != (C, D, spec.n_centres, spec.latent_dim) or abs(float(z["noise"]) - spec.noise) > 1e-12:
        return None
    return [(torch.from_numpy(z["W1"]).to(device), torch.from_numpy(z["b1"]).to(device)),
            (torch.from_numpy(z["W2"]).to(device), torch.from_numpy(z["b2"]).to(device))]


def train_index(spec, C, device, rank, world, steps):
    """Stand-in for LearnedIndexBuilder.build: the committed weights for the default workload, else rank 0 trains
    on a fresh 200k-row draw and the weights are broadcast."""
    from learnedmetricindex_b200 import synthetic as syn

    layers = load_standin_asset(spec, C, device)
    if layers is not None:
        log(f"index stand-in: committed weights {os.path.relpath(STANDIN_ASSET, ROOT)}")
        return layers
    shapes = syn.mlp_layer_shapes("MLP-4", D, C)
    if rank == 0:
        xt = syn.make_vectors(TRAIN_ROWS, spec, device=device, stream=2).float()
        g = torch.Generator(device=device).manual_seed(syn.SEED)
        cent = syn._kmeans_centroids(xt, C, 10, g)
        labels = (xt @ cent.t()).argmax(dim=1)
        layers = syn.train_standin_mlp(xt, labels, C, "MLP-4", steps=steps)
        del xt
    else:
        layers = [(torch.empty(o, i, device=device), torch.empty(o, device=device)) for o, i in shapes]
    if world > 1:
        for W, b in layers:
            dist.broadcast(W, 0)
            dist.broadcast(b, 0)
    return layers


def stable_rank_in_group(labels, C):
    """For each row: its position among the rows of the same label (original order)."""
    order = torch.sort(labels, stable=True).indices
    counts = torch.bincount(labels, minlength=C)
    starts = torch.cumsum(counts, 0) - counts
    pos = torch.empty_like(order)
    pos[order] = torch.arange(labels.numel(), device=labels.device)
    return pos - starts[labels], counts


def build_workload(args, device, rank, world, use_repo_kernels=True):
    """Returns dict with the device layout of THIS rank's buckets, MLP layers, queries."""
    from learnedmetricindex_b200 import synthetic as syn

    n, nq, C = args.n, args.nq, args.buckets
    spec = syn.MixtureSpec(n_centres=args.centres, noise=args.noise, latent_dim=args.latent)
    t0 = time.time()
    layers = train_index(spec, C, device, rank, world, args.train_steps)
    log(f"index stand-in trained in {time.time() - t0:.1f}s")

    if use_repo_kernels:
        from learnedmetricindex_b200 import NeuralNetwork

        net = NeuralNetwork(layers, device=device)
        predict = lambda x: net.predict_device(x)
    else:
        predict = lambda x: syn.torch_mlp_argmax(layers, x)

    # pass 1: bucket of every row (data_prediction = argmax MLP(x), LearnedIndexBuilder.py:76)
    t0 = time.time()
    plan = chunk_plan(n)
    preds = []
    sizes = torch.zeros(C, dtype=torch.int64, device=device)
    for ci, (s, m) in enumerate(plan):
        x = make_chunk(ci, m, spec, device)
        p = predict(x).to(torch.int64)
        preds.append(p.to(torch.int32))
        sizes += torch.bincount(p, minlength=C)
        del x
    sizes_h = sizes.cpu().numpy()
    log(f"pass 1 (bucket assignment of {n} rows) {time.time() - t0:.1f}s; bucket sizes min/mean/max = "
        f"{sizes_h.min()}/{int(sizes_h.mean())}/{sizes_h.max()}")

    from learnedmetricindex_b200.sharded import owner_of_bucket, partition_buckets

    # Balance the ranks' SCAN WORK, not their bytes: a bucket costs rows x (probing queries rounded up to
    # the scan's 128-query tiles).  The expected probes per bucket come from a calibration batch drawn
    # from the data distribution (its own random stream, NOT the timed queries), routed by the index.
    weights = sizes_h.astype(np.float64)  # fallback: queries follow the data, so work goes like size^2
    probe = int(getattr(args, "probe", 0) or 0)
    if world > 1 and use_repo_kernels and probe > 0 and getattr(args, "balance", "size2") == "calibrated":
        qcal = syn.make_vectors(nq, spec, device=device, stream=2)
        order_cal, _ = net.topn_device(qcal, min(probe, C))
        m_b = torch.bincount(order_cal.reshape(-1).to(torch.int64), minlength=C).cpu().numpy()
        weights = 128.0 * np.ceil(np.maximum(m_b, 1) / 128.0)
        del qcal, order_cal
    parts = partition_buckets(sizes_h, world, weights=weights)
    # this rank's slice [lo, hi) of every bucket's rows (in bucket order): the whole bucket or nothing when
    # whole buckets are assigned to ranks; an equal 1/world slice of EVERY bucket with --shard rows (every
    # rank then has the same buckets, probes and tile structure: perfectly balanced scan work)
    if getattr(args, "shard", "buckets") == "rows" and world > 1:
        lo_h = sizes_h * rank // world
        hi_h = sizes_h * (rank + 1) // world
    else:
        mine_h = owner_of_bucket(parts, C) == rank
        lo_h = np.zeros(C, dtype=np.int64)
        hi_h = np.where(mine_h, sizes_h, 0)
    local_sizes = hi_h - lo_h
    lo_d, hi_d = torch.from_numpy(lo_h).to(device), torch.from_numpy(hi_h).to(device)
    off = np.zeros(C + 1, dtype=np.int64)
    off[1:] = np.cumsum(local_sizes)
    n_local = int(off[-1])
    off_d = torch.from_numpy(off).to(device)

    # pass 2: regenerate each chunk and scatter this rank's rows straight into the bucket-major layout
    t0 = time.time()
    vectors = torch.empty((max(n_local, 1), D), dtype=torch.float16, device=device)
    ids = torch.empty(max(n_local, 1), dtype=torch.int32, device=device)
    filled = torch.zeros(C, dtype=torch.int64, device=device)
    if use_repo_kernels:
        from learnedmetricindex_b200 import _native

        L = _native.lib()
    for ci, (s, m) in enumerate(plan):
        x = make_chunk(ci, m, spec, device)
        lab = preds[ci].to(torch.int64)
        rk, counts = stable_rank_in_group(lab, C + 1)
        pos = filled[lab] + rk  # position of the row inside its bucket (all ranks agree)
        mine = (pos >= lo_d[lab]) & (pos < hi_d[lab])
        dest = off_d[lab] + pos - lo_d[lab]
        sel = torch.nonzero(mine).reshape(-1)
        if sel.numel():
            dsel = dest[sel].contiguous()
            xs = x if sel.numel() == m else x[sel].contiguous()
            if use_repo_kernels:
                _native.check(L.lmi_scatter_rows(xs.data_ptr(), 1, dsel.data_ptr(), int(sel.numel()), D,
                                                 vectors.data_ptr(), D, None,
                                                 torch.cuda.current_stream().cuda_stream), "lmi_scatter_rows")
                torch.cuda.current_stream().synchronize()
            else:
                vectors[dsel] = xs
            gid = (sel + (s + 1)).to(torch.int64)  # 1-based row labels (search/search.py:190-191)
            gid = torch.where(gid >= 2 ** 31, gid - 2 ** 32, gid).to(torch.int32)
            ids[dsel] = gid
        filled += counts[:C]
        del x
    del preds
    log(f"pass 2 (bucket-major layout, {n_local} local rows, {n_local * D * 2 / 1e9:.2f} GB) {time.time() - t0:.1f}s")

    Q = syn.make_vectors(nq, spec, device=device, stream=1)
    return dict(layers=layers, vectors=vectors[:n_local], ids=ids[:n_local], bucket_off=off, sizes=sizes_h,
                local_sizes=local_sizes, Q=Q, parts=parts, spec=spec)


def ground_truth(Q, vectors, ids, world, rows_per_block=131072):
    """Exact top-10 by inner product over ALL rows (Baseline.search semantics, ids 1-based):
    per-rank brute force in fp32 with torch (harness, not the product path), merged over ranks."""
    dev = Q.device
    nq = Q.shape[0]
    best_s = torch.full((nq, K_NN), -float("inf"), device=dev)
    best_i = torch.zeros((nq, K_NN), dtype=torch.int64, device=dev)
    Qf = Q.float()
    for s in range(0, vectors.shape[0], rows_per_block):
        blk = vectors[s : s + rows_per_block]
        sim = Qf @ blk.float().t()
        kk = min(K_NN, blk.shape[0])
        ts, ti = torch.topk(sim, kk, dim=1)
        gi = ids[s : s + rows_per_block].to(torch.int64)[ti] & 0xFFFFFFFF
        cs = torch.cat([best_s, ts], 1)
        ci = torch.cat([best_i, gi], 1)
        best_s, sel = torch.topk(cs, K_NN, dim=1)
        best_i = torch.gather(ci, 1, sel)
        del sim
    if world > 1:
        all_s = [torch.empty_like(best_s) for _ in range(world)]
        all_i = [torch.empty_like(best_i) for _ in range(world)]
        dist.all_gather(all_s, best_s)
        dist.all_gather(all_i, best_i)
        cs, ci = torch.cat(all_s, 1), torch.cat(all_i, 1)
        best_s, sel = torch.topk(cs, K_NN, dim=1)
        best_i = torch.gather(ci, 1, sel)
    return best_i.cpu().numpy()


def recall_at_10(nns, gt):
    """eval/eval.py:37-45 (set intersection), vectorised."""
    hits = 0
    for a, b in zip(nns.astype(np.int64), gt):
        hits += len(set(a[:K_NN].tolist()) & set(b[:K_NN].tolist()))
    return hits / (len(gt) * K_NN)


# --------------------------------------------------------------------------------------
# CPU side: the oracle port and the unmodified reference, on the WHOLE query batch
# --------------------------------------------------------------------------------------
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def use_all_host_threads():
    """All cores of this process' affinity mask, whatever OMP_NUM_THREADS the launcher exported."""
    n = host_threads()
    torch.set_num_threads(n)
    return torch.get_num_threads()


class CpuBaseline:
    """The CPU implementations of the path over the same index, host side:
    ``run``            the oracle port (oracle/lmi_oracle.py: the reference's arithmetic without its per-round
                       groupby / .loc copies) -- the FASTER of the two, so the one the ratio is quoted against;
    ``run_unmodified`` the reference's own ``LearnedIndex.search`` (search/li/LearnedIndex.py:41-161), staged
                       byte-identical under oracle/_ref (oracle/build_ref.py), through oracle/ref_harness.py."""

    def __init__(self, w, C):
        import psutil

        from learnedmetricindex_b200 import synthetic as syn
        from oracle import lmi_oracle as orc

        self.orc = orc
        self.C = C
        self.layers = syn.layers_to_numpy(w["layers"])
        vec = w["vectors"].cpu().numpy()
        self.ids = (w["ids"].cpu().numpy().view(np.uint32)).astype(np.int64)
        self.offsets = np.asarray(w["bucket_off"], dtype=np.int64)
        need_f32 = vec.size * 4
        avail = psutil.virtual_memory().available
        self.f32 = avail > 2.5 * need_f32 + (8 << 30)
        if self.f32:
            # the reference's frames are f32 after data_X_to_torch / faiss' cast; pre-widening once is
            # generous to the CPU side (the reference re-copies every bucket every round, :357)
            vec = torch.from_numpy(vec).float().numpy()
        self.layout = orc.BucketMajor.from_grouped(vec, self.offsets, self.ids)
        self.Q = w["Q"].cpu().numpy()
        self.index = orc.OracleIndex(self.layers)
        self.threads = use_all_host_threads()
        self.ram_gb = psutil.virtual_memory().total / 2 ** 30

    def run(self, B, n_queries=None):
        q = self.Q if n_queries is None else self.Q[:n_queries]
        t0 = time.perf_counter()
        d, i, times, order = self.orc.search(self.index, q, None, q, None, [self.C], n_buckets=B, k=K_NN,
                                             layout=self.layout)
        return time.perf_counter() - t0, d, i, order

    def run_single_thread(self, B, n_queries=None):
        """The port on ONE thread -- the figure that sits beside the README's 1-core 45 QPS (README.md:54-68)."""
        n = torch.get_num_threads()
        torch.set_num_threads(1)
        try:
            return self.run(B, n_queries)
        finally:
            torch.set_num_threads(n)

    def unmodified_available(self):
        from oracle import ref_harness

        return ref_harness.reference_available()

    def run_unmodified(self, B, limit_s=600.0):
        """ONE call of the reference's own ``LearnedIndex.search`` on the whole batch.  The two DataFrames
        view this object's f32 matrix (rows are bucket-major there; ``data_prediction`` and the index labels
        are permuted alike, so the groupby yields the same groups in the same order as on the original
        frame).  Returns (seconds, dists, nns, measured_time) or raises (RAM guard / time limit)."""
        import signal

        import psutil

        from oracle import ref_harness

        vec = self.layout.vectors
        if vec.dtype != np.float32:
            raise MemoryError("vectors were not widened to f32 (host RAM too small for the reference's frames)")
        # the reference holds the frame, one regrouped copy of it per round and the bucket being scanned
        need = 1.35 * vec.nbytes + (6 << 30)
        if psutil.virtual_memory().available < need:
            raise MemoryError(f"reference search needs ~{need / 2 ** 30:.0f} GiB free, "
                              f"{psutil.virtual_memory().available / 2 ** 30:.0f} GiB available")
        li = ref_harness.make_reference_index(self.layers, "MLP-4")  # MLP pinned to the CPU (model.py:124-125)
        labels = np.repeat(np.arange(self.C, dtype=np.int64), np.diff(self.offsets)).reshape(-1, 1)

        def on_alarm(signum, frame):
            raise TimeoutError(f"reference search exceeded {limit_s:.0f} s")

        old = signal.signal(signal.SIGALRM, on_alarm)
        signal.alarm(int(limit_s))
        try:
            t0 = time.perf_counter()
            d, i, tm = ref_harness.run_reference_search(li, vec, self.Q, vec, self.Q, labels, [self.C], B, K_NN,
                                                        ids=self.ids, copy=False)
            dt = time.perf_counter() - t0
        finally:
            signal.alarm(0)
            signal.signal(signal.SIGALRM, old)
        return dt, d, i, tm

    def describe(self, B, nq):
        return (f"all {nq} of the {nq} queries against the full index, B={B}, oracle/lmi_oracle.py "
                f"(torch CPU MLP + per-bucket f32 sgemm + top-10 + stable merge), vectors "
                f"{'pre-widened to f32 once' if self.f32 else 'fp16, widened per bucket per round'}, "
                f"{self.threads} torch threads")

    # -- parity of a GPU result against this CPU result ----------------------------------------------
    def explain(self, d_cpu, i_cpu, order_cpu, d_gpu, i_gpu, order_gpu, tie_abs=4e-6, prob_tie=1e-6):
        """Every difference between the GPU result and the oracle's must be a tie the north_star tolerates:
        a bucket-order row may differ only where the swapped classes' probabilities are within 1e-6;
        on rows with the same bucket order an id may differ only where the object the GPU returned is, by an
        exact float64 inner product on the host vectors, within ``tie_abs`` of the oracle's entry at that
        position (two objects at the same distance up to f32 rounding).  Returns the parity dict."""
        nq, k = i_cpu.shape
        order_cpu = order_cpu[:, :, 0]
        rows_equal = (order_gpu == order_cpu).all(axis=1)
        unexplained_order = 0
        bad_rows = np.flatnonzero(~rows_equal)
        if bad_rows.size:
            probs, classes = self.orc.predict_proba(self.layers, self.orc.to_f32_tensor(self.Q[bad_rows]))
            for j, q in enumerate(bad_rows):
                p_of = np.empty(self.C, dtype=np.float64)
                p_of[classes[j]] = probs[j]
                diff = order_gpu[q] != order_cpu[q]
                if np.abs(p_of[order_gpu[q][diff]] - p_of[order_cpu[q][diff]]).max() > prob_tie:
                    unexplained_order += 1
        ok = rows_equal
        mism = (i_gpu != i_cpu) & ok[:, None]
        rel = np.abs(d_gpu - d_cpu) / np.maximum(np.abs(d_cpu), 1e-3)
        finite = np.isfinite(d_cpu) & (np.abs(d_cpu) < 1e38)
        dist_bad = int(((rel > 1e-3) & finite & ok[:, None] & ~mism).sum())
        unexplained = 0
        max_gap = 0.0  # largest |exact f64 distance of the GPU's object - oracle's entry| among the explained ones
        qs, js = np.nonzero(mism)
        row_of_id = None
        if qs.size:
            sorter = np.argsort(self.layout.ids, kind="stable")
            for q, j in zip(qs.tolist(), js.tolist()):
                gid = int(i_gpu[q, j])
                pos = np.searchsorted(self.layout.ids, gid, sorter=sorter)
                if pos >= sorter.size or self.layout.ids[sorter[pos]] != gid:
                    unexplained += 1
                    continue
                x = self.layout.vectors[sorter[pos]].astype(np.float64)
                true_d = 1.0 - float(np.dot(x, self.Q[q].astype(np.float64)))
                gap = abs(true_d - float(d_cpu[q, j]))
                if gap > tie_abs:
                    unexplained += 1
                else:
                    max_gap = max(max_gap, gap)
        n_ok = int(ok.sum())
        return {"bucket_order_rows_equal": float(rows_equal.mean()),
                "bucket_order_rows_unexplained": int(unexplained_order),
                "id_agreement": float((i_gpu[ok] == i_cpu[ok]).mean()) if n_ok else None,
                "id_mismatches": int(mism.sum()), "unexplained": int(unexplained + unexplained_order + dist_bad),
                "dists_beyond_rel_1e-3": dist_bad, "queries": int(nq),
                "max_abs_dist_gap_of_explained_id_mismatches": max_gap, "tie_abs": tie_abs,
                "max_rel_dist_error": float(rel[finite & ok[:, None] & ~mism].max()) if (finite & ok[:, None] & ~mism).any() else None,
                "rule": f"ids equal except where the exact f64 distance of the returned object is within {tie_abs:g} "
                        f"of the oracle's entry (f32 rounding tie); bucket rows equal except probability ties <= {prob_tie:g}"}


# --------------------------------------------------------------------------------------
_REAL_STDOUT = None


def bench_config(wl_name, n, nq, C, B, world):
    """The `config` object of the JSON line -- built by this one function for BOTH arms (ours and
    `--impl reference`), so the two lines describe the workload with identical keys and values; what only one
    arm knows (recall, sharding, generator parameters) goes into the sibling key `config_detail`."""
    return {"workload": wl_name, "n": n, "nq": nq, "d": D, "n_buckets_in_index": C, "probed_buckets": B, "k": K_NN,
            "l2_policy": "inputs larger than L2 (bucket vectors %.1f GB over %d GPU(s), every probed bucket read "
                         "once per step)" % (n * D * 2 / 1e9, world)}


def emit_json(obj):
    """The ONE line on the real stdout (libraries such as NCCL print banners to fd 1, which main() points at stderr)."""
    line = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, line)



# --------------------------------------------------------------------------------------
# multi-level workload: `--workload ml`, single GPU
# --------------------------------------------------------------------------------------
def run_multilevel(args, emit=True):
    """The reference CLI's default tree shape (`--n-categories 10 10`, search/search.py:314) in the
    notebook's arrangement: a low-dimensional navigation representation (PCA of the 768-d vectors,
    search/search.py:192-216) routes through a 2-level tree of `MLP` models, the scan runs on the
    768-d vectors.  Prints one JSON line: device-timed QPS, end-to-end QPS through
    `LearnedIndex.search` with host queries, per-stage milliseconds, recall@10, and -- the
    cpu_baseline leg -- the CPU oracle's QPS on a query sample with the agreement of the visit
    order and ids on that sample."""
    from learnedmetricindex_b200 import LearnedIndex
    from learnedmetricindex_b200 import synthetic as syn
    from learnedmetricindex_b200.layout import DeviceBucketLayout
    from learnedmetricindex_b200.navigation import flat_bucket_ids

    assert int(os.environ.get("WORLD_SIZE", "1")) == 1, "--workload ml is a single-GPU measurement"
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    cats, B = list(args.cats), args.probe

    spec = syn.MixtureSpec()
    X = syn.make_vectors(args.n, spec, device=dev)
    Q = syn.make_vectors(args.nq, spec, device=dev, stream=1)
    # navigation representation: PCA to nav_dim from a sample, re-normalised, stored fp16
    g = torch.Generator(device=dev).manual_seed(7)
    samp = X[torch.randperm(args.n, generator=g, device=dev)[:100_000]].float()
    mean = samp.mean(0, keepdim=True)
    _, _, V = torch.pca_lowrank(samp - mean, q=args.nav_dim, center=False, niter=4)
    proj = lambda t: torch.nn.functional.normalize((t.float() - mean) @ V, dim=1).to(torch.float16)
    Xn = torch.cat([proj(X[s : s + (1 << 20)]) for s in range(0, args.n, 1 << 20)])
    Qn = proj(Q)
    t0 = time.time()
    root, internal, pred, bucket_paths = syn.build_standin_index_two_level(Xn, cats, args.ml_model_type)
    log(f"stand-in 2-level build: {len(internal)} internal models, {len(bucket_paths)} buckets, {time.time() - t0:.1f}s")

    to_np = syn.layers_to_numpy
    li = LearnedIndex(to_np(root), {p: to_np(l) for p, l in internal.items()}, bucket_paths)
    labels = torch.from_numpy(flat_bucket_ids(pred.cpu().numpy(), cats)).to(dev)
    n_layout = int(np.prod(cats)) + int((labels == int(np.prod(cats))).any().item())
    li.attach_layout(DeviceBucketLayout.from_device(X, labels, n_layout))
    sizes = np.diff(li._layout.bucket_off_host)
    log(f"bucket sizes: min {sizes.min()} max {sizes.max()}")

    # exact ground truth for recall
    gt = torch.empty((args.nq, 10), dtype=torch.int64, device=dev)
    for s in range(0, args.nq, 1000):
        best_v = torch.full((min(1000, args.nq - s), 10), -2.0, device=dev)
        best_i = torch.zeros((min(1000, args.nq - s), 10), dtype=torch.int64, device=dev)
        for r in range(0, args.n, 1 << 20):
            sim = Q[s : s + 1000].float() @ X[r : r + (1 << 20)].float().t()
            v, i = sim.topk(10, dim=1)
            allv, alli = torch.cat([best_v, v], 1), torch.cat([best_i, i + r + 1], 1)
            best_v, sel = allv.topk(10, dim=1)
            best_i = torch.gather(alli, 1, sel)
        gt[s : s + 1000] = best_i
    gt = gt.cpu().numpy()

    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    stage = np.zeros(4)
    for it in range(args.warmup + args.steps):
        if it == args.warmup:
            torch.cuda.synchronize()
            t_start = torch.cuda.Event(enable_timing=True)
            t_start.record()
        d_dev, i_dev = li.search_device(Qn, Q, cats, B, 10, events=ev)
        if it >= args.warmup:
            torch.cuda.synchronize()
            stage += [ev[j].elapsed_time(ev[j + 1]) for j in range(4)]
    t_end = torch.cuda.Event(enable_timing=True)
    t_end.record()
    torch.cuda.synchronize()
    ms = t_start.elapsed_time(t_end) / args.steps
    stage /= args.steps
    nns = i_dev.cpu().numpy().view(np.uint32).astype(np.int64)
    recall = float(np.mean([len(set(a) & set(b)) for a, b in zip(nns, gt)]) / 10)

    Qn_h, Q_h = Qn.cpu().numpy(), Q.cpu().numpy()
    e2e = []
    for it in range(3 + 10):
        t0 = time.perf_counter()
        d_h, i_h, _ = li.search(None, Qn_h, None, Q_h, None, cats, n_buckets=B, k=10)
        e2e.append(time.perf_counter() - t0)
    e2e_ms = float(np.median(e2e[3:]) * 1e3)

    cpu = parity = None
    if not args.no_cpu_baseline:  # the cpu_baseline leg: the oracle as checker and as the timed CPU path
        from oracle import lmi_oracle as orc

        nc = min(args.ml_cpu_queries, args.nq)
        lay = orc.BucketMajor(X.cpu().numpy().astype(np.float32), pred.cpu().numpy())
        idx = orc.OracleIndex(to_np(root), {p: to_np(l) for p, l in internal.items()}, bucket_paths)
        t0 = time.time()
        exp_d, exp_i, _, exp_order = orc.search(idx, Qn_h[:nc], None, Q_h[:nc], None, cats, n_buckets=B, k=10,
                                                layout=lay)
        cpu_s = time.time() - t0
        got_paths, _ = li._precompute_bucket_order(Qn_h[:nc], B, cats)
        same = (got_paths == exp_order).all(axis=(1, 2))
        cpu = {"value": nc / cpu_s, "unit": "QPS", "kind": "port", "cores": torch.get_num_threads(),
               "sample": f"first {nc} queries, oracle/lmi_oracle.py, vectors pre-widened to f32"}
        parity = {"visit_order_rows_equal": float(same.mean()),
                  "id_agreement_on_equal_rows": float((i_h[:nc][same] == exp_i[same]).mean())}

    out_ml = {
        "metric": "QPS, 2-level LMI", "value": args.nq / ms * 1e3, "unit": "QPS", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "data": "synthetic",
        "dtype": "f16 in / f32 accumulate (scan), tf32 x3 split / f32 accumulate (MLPs)",
        "config": {"workload": f"{args.n} x 768 fp16, n_categories {cats}, {args.ml_model_type} models on a "
                               f"{args.nav_dim}-d PCA navigation representation, {args.nq} queries, "
                               f"{B} of {len(bucket_paths)} buckets, k=10", "recall_at_10": recall},
        "kernel_ms": {"navigation": stage[0], "invert": stage[1], "scan": stage[2], "merge": stage[3]},
        "e2e": {"value": args.nq / e2e_ms * 1e3, "unit": "QPS", "ms_per_step": e2e_ms},
        "cpu_baseline": cpu, "parity_vs_oracle": parity,
    }
    if emit:
        emit_json(out_ml)
    return out_ml


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="b", choices=sorted(WORKLOADS))
    ap.add_argument("--n", type=int, default=0)
    ap.add_argument("--nq", type=int, default=0)
    ap.add_argument("--buckets", type=int, default=0)
    ap.add_argument("--probe", type=int, default=None,
                    help="buckets probed per query (B); default 4 (10 for --workload ml); 0 = smallest B with recall>=0.90")
    ap.add_argument("--train-steps", type=int, default=300)
    ap.add_argument("--centres", type=int, default=2000)
    ap.add_argument("--noise", type=float, default=0.25)
    ap.add_argument("--latent", type=int, default=24)
    ap.add_argument("--cpu-budget", type=float, default=20.0, help="seconds of CPU work for the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-reference-extras", action="store_true",
                    help="--impl reference: skip the 1-thread call and the unmodified-reference call")
    ap.add_argument("--reference-limit", type=float, default=600.0,
                    help="time limit (s) of the ONE unmodified li.LearnedIndex.search call")
    ap.add_argument("--no-workload-c", action="store_true",
                    help="skip the appended 100M x 768 measurement (BASELINE.json configs[2])")
    ap.add_argument("--only-headline", action="store_true",
                    help="the headline measurement alone: no CPU leg, no appended workloads, sweeps or cold call "
                         "(what the ncu passes run)")
    ap.add_argument("--no-workload-ml", action="store_true",
                    help="skip the appended 2-level (n_categories 10 10) measurement, the reference CLI's default tree")
    ap.add_argument("--no-config-a", action="store_true",
                    help="skip the appended 100k x 768 measurement (BASELINE.json configs[0])")
    ap.add_argument("--no-sweeps", action="store_true",
                    help="skip the appended budget sweep (configs[3]) and batch sweep (configs[4])")
    ap.add_argument("--no-cold-call", action="store_true",
                    help="skip the cold-call measurement (first search() of a fresh index from host data)")
    ap.add_argument("--sweep", default="", help="comma list of B values: print recall/QPS per B to stderr")
    ap.add_argument("--balance", default="size2", choices=["calibrated", "size2"],
                    help="--shard buckets: LPT cost = rows^2 (queries follow the data; measured the more robust of the "
                         "two at 8 GPUs: 1.142 vs 1.165 ms/step over 3 runs each) or rows x calibrated query tiles")
    ap.add_argument("--shard", default="auto", choices=["auto", "buckets", "rows"],
                    help="multi-GPU: whole buckets per rank (calibrated LPT) or an equal row slice of every bucket; "
                         "auto = rows when a slice still holds >= 4 scan ranges of 16384 rows, else buckets")
    ap.add_argument("--cats", type=int, nargs=2, default=[10, 10], help="--workload ml: n_categories")
    ap.add_argument("--nav-dim", type=int, default=96, help="--workload ml: PCA navigation dimensionality")
    ap.add_argument("--ml-model-type", default="MLP", help="--workload ml: model type of every tree node")
    ap.add_argument("--ml-cpu-queries", type=int, default=2000, help="--workload ml: queries of the cpu_baseline leg")
    args = ap.parse_args()
    if args.only_headline:
        args.no_cpu_baseline = args.no_workload_c = args.no_workload_ml = args.no_config_a = True
        args.no_sweeps = args.no_cold_call = True
    args.default_sizes = not (args.n or args.nq or args.buckets)  # before the workload defaults are filled in
    if args.probe is None:
        args.probe = 10 if args.workload == "ml" else 4
    wl = WORKLOADS[args.workload]
    args.n = args.n or wl["n"]
    args.nq = args.nq or wl["nq"]
    args.buckets = args.buckets or wl["buckets"]
    wl_name = wl["name"] if args.default_sizes else \
        f"{args.n} x 768 fp16, {args.buckets} buckets, MLP-4, {args.nq} queries, k=10"
    assert args.warmup >= 0 and args.steps >= 1

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.workload == "ml":
        assert args.impl == "ours", "the multi-level workload has no reference arm (its cpu_baseline leg times the oracle)"
        return run_multilevel(args)
    if args.impl == "reference":
        return run_reference(args, wl_name, rank, world, local_rank)

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback in the product path)"
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    assert world == args.gpus or world == 1 and args.gpus == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    out = measure(args, wl_name, device, rank, world, local_rank, headline=True)

    # BASELINE.json configs[2] (north_star: ">= 6x scaling from 1 to 8 GPUs on 100M x 768") rides on the same
    # line at every N -- one B200 holds the 153.6 GB index, so the 1-GPU point exists -- as "workload_c"
    if args.workload == "b" and args.default_sizes and not args.no_workload_c:
        import copy
        import gc

        gc.collect()
        torch.cuda.empty_cache()
        ac = copy.copy(args)
        ac.workload, ac.n, ac.shard = "c", WORKLOADS["c"]["n"], "auto"
        ac.steps, ac.warmup = min(args.steps, 10), 3
        try:
            oc = measure(ac, WORKLOADS["c"]["name"], device, rank, world, local_rank, headline=False)
            keep = ("value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "scaling", "config", "config_detail", "e2e",
                    "kernel_ms", "roofline", "clocks", "gpu_launches", "host_enqueue_ms_per_step", "multi_gpu_parity")
            out["workload_c"] = {k_: oc[k_] for k_ in keep if k_ in oc}
        except Exception as e:  # noqa: BLE001 -- the headline stands; say why the 100M point is absent
            log(f"workload c failed: {e!r}")
            out["workload_c"] = {"error": repr(e)}
    # BASELINE.json configs[0] (the reference's own CPU-runnable size, 100k x 768) on the same line at N = 1, with
    # the unmodified reference as the checker of the timed result
    if args.workload == "b" and args.default_sizes and world == 1 and not args.no_config_a:
        import copy
        import gc

        gc.collect()
        torch.cuda.empty_cache()
        aa = copy.copy(args)
        aa.workload, aa.n, aa.shard = "a", WORKLOADS["a"]["n"], "auto"
        try:
            oa = measure(aa, WORKLOADS["a"]["name"], device, rank, world, local_rank, headline=False,
                         check_unmodified=True)
            keep = ("value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "config", "config_detail", "e2e",
                    "kernel_ms", "roofline", "gpu_launches", "host_enqueue_ms_per_step", "parity_vs_reference")
            out["config_a"] = {k_: oa[k_] for k_ in keep if k_ in oa}
        except SystemExit:
            raise
        except Exception as e:  # noqa: BLE001
            log(f"config a failed: {e!r}")
            out["config_a"] = {"error": repr(e)}
        gc.collect()
        torch.cuda.empty_cache()
    # the reference CLI's default tree shape (--n-categories 10 10, search/search.py:314) on the same line at N = 1
    if args.workload == "b" and args.default_sizes and world == 1 and not args.no_workload_ml:
        import copy
        import gc

        am = copy.copy(args)
        am.workload, am.n, am.nq, am.buckets, am.probe = "ml", WORKLOADS["ml"]["n"], WORKLOADS["ml"]["nq"], \
            WORKLOADS["ml"]["buckets"], 10
        am.steps = min(args.steps, 20)
        try:
            out["workload_ml"] = run_multilevel(am, emit=False)
        except Exception as e:  # noqa: BLE001
            log(f"workload ml failed: {e!r}")
            out["workload_ml"] = {"error": repr(e)}
        gc.collect()
        torch.cuda.empty_cache()
    if rank == 0:
        emit_json(out)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def all_ranks_ok(ok, device, world):
    """True only when every rank reports success (a rank that failed must not leave the others in a collective)."""
    if world == 1:
        return bool(ok)
    t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


def result_checksum(d, i):
    """Order-sensitive 64-bit checksum of a device result (dists f64[nq,k], nns i32[nq,k])."""
    w = torch.arange(1, d.numel() + 1, dtype=torch.int64, device=d.device)
    a = (d.reshape(-1).view(torch.int64) * w).sum()
    b = (i.reshape(-1).to(torch.int64) * (w + 7)).sum()
    return int((a ^ (b << 1)).item())


def measure(args, wl_name, device, rank, world, local_rank, headline, check_unmodified=False):
    """Build the workload ``args`` describes over the ``world`` ranks, check it, time it; returns the JSON object.
    ``headline=False`` (the appended 100M point) skips the CPU leg, the budget search and the sweeps."""
    if args.shard == "auto":
        args.shard = "rows" if world > 1 and args.n // (world * args.buckets) >= 4 * 16384 else "buckets"

    from learnedmetricindex_b200 import DeviceBucketLayout, LearnedIndex, _native
    from learnedmetricindex_b200.sharded import ShardedLearnedIndex

    _native.lib()  # fail loudly if the CUDA library is missing
    C, nq = args.buckets, args.nq
    w, err = None, None
    try:
        w = build_workload(args, device, rank, world)
    except Exception as e:  # noqa: BLE001 (e.g. out of memory on one rank)
        err = e
    if not all_ranks_ok(err is None, device, world):
        raise RuntimeError(f"workload construction failed on some rank: {err!r}")
    layout = DeviceBucketLayout(w["vectors"], torch.from_numpy(w["bucket_off"]), w["ids"])
    li = LearnedIndex(w["layers"], {}, [(i,) for i in range(C)])
    li.attach_layout(layout)
    engine = ShardedLearnedIndex(li) if world > 1 else li
    Q = w["Q"]

    t0 = time.time()
    gt = ground_truth(Q, w["vectors"], w["ids"], world)
    log(f"ground truth (exact top-10 over {args.n} rows) {time.time() - t0:.1f}s")

    def run_once(B, eng=None):
        d, i = (eng or engine).search_device(Q, Q, [C], B, K_NN)
        torch.cuda.synchronize()
        return d.cpu().numpy(), i.cpu().numpy().view(np.uint32)

    B = args.probe
    recalls = {}
    if B <= 0:
        for cand in range(1, C + 1):
            recalls[cand] = recall_at_10(run_once(cand)[1], gt)
            log(f"recall@10 with B={cand}: {recalls[cand]:.4f}")
            if recalls[cand] >= 0.90:
                B = cand
                break
    recall = recalls.get(B) or recall_at_10(run_once(B)[1], gt)
    log(f"operating point: B={B}, recall@10={recall:.4f}")
    while recall < 0.90 and B < C:
        B += 1
        recall = recall_at_10(run_once(B)[1], gt)
        log(f"recall below 0.90, raising B to {B}: recall@10={recall:.4f}")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- multi-GPU parity, asserted before anything is timed ----
    multi = None
    if world > 1:
        d_p, i_p = engine.search_device(Q, Q, [C], B, K_NN)
        d_p, i_p = d_p.clone(), i_p.clone()
        # (1) the two exchange paths (peer-memory merge kernel / NCCL all-gather + merge) are bit-identical
        eng_nccl = ShardedLearnedIndex(li, exchange="nccl")
        d_n, i_n = eng_nccl.search_device(Q, Q, [C], B, K_NN)
        modes_equal = bool(torch.equal(d_p, d_n) and torch.equal(i_p, i_n))
        # (2) every rank holds the same final [nq, k]
        cs = torch.tensor([result_checksum(d_p, i_p)], dtype=torch.int64, device=device)
        all_cs = [torch.empty_like(cs) for _ in range(world)]
        dist.all_gather(all_cs, cs)
        same_everywhere = len({int(c.item()) for c in all_cs}) == 1
        multi = {"exchange": engine.exchange, "exchange_modes_bit_equal": modes_equal,
                 "ranks_hold_same_result": same_everywhere}
        # (3) sharded == unsharded: rank 0 rebuilds the WHOLE index on its own GPU (fits at 10M) and searches it
        if args.n * D * 2 <= 40e9:
            same_unsharded = True
            if rank == 0:
                w1 = build_workload(args, device, 0, 1)
                li1 = LearnedIndex(w1["layers"], {}, [(i,) for i in range(C)])
                li1.attach_layout(DeviceBucketLayout(w1["vectors"], torch.from_numpy(w1["bucket_off"]), w1["ids"]))
                d_1, i_1 = li1.search_device(Q, Q, [C], B, K_NN)
                torch.cuda.synchronize()
                same_unsharded = bool(torch.equal(d_1, d_p) and torch.equal(i_1, i_p))
                if not same_unsharded:  # say what differs before failing
                    nd, ni = int((d_1 != d_p).sum().item()), int((i_1 != i_p).sum().item())
                    tied = int(((i_1 != i_p) & (d_1 == d_p)).sum().item())
                    log(f"sharded vs unsharded: {nd} distances and {ni} ids differ ({tied} of the ids at equal distance)")
                del w1, li1, d_1, i_1
                torch.cuda.empty_cache()
            flag = torch.tensor([1 if same_unsharded else 0], dtype=torch.int32, device=device)
            dist.broadcast(flag, 0)
            multi["sharded_equals_unsharded"] = bool(flag.item())
        del eng_nccl
        log(f"multi-GPU parity: {multi}")
        assert modes_equal, "peer-memory and NCCL exchange paths disagree"
        assert same_everywhere, "ranks hold different final results"
        assert multi.get("sharded_equals_unsharded", True), "sharded result differs from the unsharded search"

    host_enqueue_ms = [0.0]
    can_graph = world == 1 or getattr(engine, "exchange", "") == "peer"

    def timed_device_loop(Bp, steps, warmup, graph=False):
        """``steps`` back-to-back searches with device-resident queries, CUDA events around the region.
        graph=False: eager launches with an event between the stages of every step (the per-kernel split
        and the scan's launch duration come from here); graph=True: the product's default path, one CUDA
        graph replay per search (its stage events are graph nodes, read for the last step only)."""
        graph = graph and can_graph
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(5)] for _ in range(steps)]
        step = (lambda ev: engine.search_graphed(Q, Q, [C], Bp, K_NN)) if graph else \
            (lambda ev: engine.search_device(Q, Q, [C], Bp, K_NN, events=ev))
        for _ in range(warmup):
            step(None)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t_a = time.time()
        e0.record()
        for s in range(steps):
            res = step(evs[s])
        e1.record()
        host_enqueue_ms[0] = (time.time() - t_a) * 1e3 / steps  # host time to enqueue one step
        barrier()
        t_b = time.time()
        total_ms = e0.elapsed_time(e1)
        if graph:
            evs = [res[2]]
        scan_ms = [ev[2].elapsed_time(ev[3]) for ev in evs]
        parts = {name: float(np.mean([ev[a].elapsed_time(ev[b]) for ev in evs]))
                 for name, a, b in (("mlp_topn", 0, 1), ("invert", 1, 2), ("scan", 2, 3), ("merge", 3, 4))}
        if world > 1:
            t = torch.tensor([total_ms], device=device, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total_ms = float(t.item())
            per_rank = torch.zeros(world, device=device, dtype=torch.float64)
            per_rank[rank] = parts["scan"]
            dist.all_reduce(per_rank)
            parts["scan_per_rank"] = [round(v, 4) for v in per_rank.tolist()]
        return total_ms, scan_ms, parts, (t_a, t_b)

    if args.sweep and headline:
        for Bs in [int(v) for v in args.sweep.split(",")]:
            r = recall_at_10(run_once(Bs)[1], gt)
            tm, sc, parts_s, _ = timed_device_loop(Bs, 10, 3)
            log(f"sweep B={Bs}: recall@10={r:.4f} QPS={nq / (tm / 10 / 1e3):.0f} ms/step={tm / 10:.3f} parts={parts_s}")

    sampler = ClockSampler(local_rank)
    sampler.start()
    K, W = args.steps, max(args.warmup, 3)
    # the timed region: K graph replays (what `search` does per call); then the same K steps as eager
    # launches with events between the stages, for the per-kernel split and the scan's launch duration
    total_ms, _, parts_graph, win_dev = timed_device_loop(B, K, W, graph=True)
    ms_per_step = total_ms / K
    value = nq / (ms_per_step / 1e3)
    host_ms_main = host_enqueue_ms[0]
    total_ms_eager, scan_ms, parts, win_eager = timed_device_loop(B, K, W, graph=False)
    host_ms_eager = host_enqueue_ms[0]

    # ---- end to end: host queries (pinned) -> search() -> host results, every step ----
    Qh = torch.empty(Q.shape, dtype=Q.dtype).pin_memory()
    Qh.copy_(Q.cpu())
    for _ in range(W):
        engine.search(None, Qh, None, Qh, None, [C], n_buckets=B, k=K_NN)
    barrier()
    t_a = time.time()
    t0 = time.perf_counter()
    for _ in range(K):
        d_e2e, i_e2e, _ = engine.search(None, Qh, None, Qh, None, [C], n_buckets=B, k=K_NN)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    t_b = time.time()
    sampler.stop()
    e2e_qps = nq * K / e2e_s
    h2d = Qh.numel() * Qh.element_size()
    d2h = d_e2e.nbytes + i_e2e.nbytes
    assert recall_at_10(i_e2e, gt) >= recall - 1e-9
    order_gpu = li.last_bucket_order.cpu().numpy() if li.last_bucket_order is not None else None

    # ---- roofline of the scan kernel (this rank's launch) ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    clocks = sampler.summary([win_dev, win_eager, (t_a, t_b)])
    # burst peak (cuBLAS timed alone) unless the timed region itself lasts >= 1 s; the sustained figure
    # (cuBLAS looping under the power cap) rides along as a note
    long_region = total_ms_eager >= 1000.0
    tf_burst = float(peaks.get("bf16_tflops", 1590.0))
    tf_sustained = float(peaks.get("bf16_tflops_sustained", 1400.0))
    tf_peak = tf_sustained if long_region else tf_burst
    counts = li.last_plan["counts"].cpu().numpy().astype(np.int64)
    sizes_local = w["local_sizes"].astype(np.int64)
    flops = 2.0 * D * float((counts * sizes_local).sum())
    touched_rows = float(sizes_local[counts > 0].sum())
    bytes_alg = 2.0 * D * touched_rows + 2.0 * D * nq * B + 8.0 * nq * K_NN
    scan_s = float(np.mean(scan_ms)) / 1e3
    t_tensor, t_hbm = flops / (tf_peak * 1e12), bytes_alg / (hbm_peak * 1e9)
    traffic = None
    try:  # ncu dram bytes of one launch: only valid for the exact single-GPU default configuration it was taken on
        if world == 1 and args.default_sizes and B == 4:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "scan_traffic.json"))).get(args.workload)
    except Exception:
        pass
    if t_tensor >= t_hbm:
        roof = {"bound": "tensor", "achieved": flops / scan_s / 1e12, "peak": tf_peak, "unit": "TFLOP/s"}
    else:
        roof = {"bound": "hbm", "achieved": bytes_alg / scan_s / 1e9, "peak": hbm_peak, "unit": "GB/s"}
    roof["frac"] = roof["achieved"] / roof["peak"]
    roof["traffic"] = traffic
    if roof["bound"] == "tensor":
        roof["frac_of_burst_peak"] = roof["achieved"] / tf_burst
        roof["frac_of_sustained_peak"] = roof["achieved"] / tf_sustained
    roof.update(kernel="scan_topk_tc_kernel", launch_ms=scan_s * 1e3, alg_flops=flops, alg_bytes=bytes_alg,
                hbm_gbs_achieved=bytes_alg / scan_s / 1e9, tflops_achieved=flops / scan_s / 1e12,
                peak_source=("MEASURED_PEAKS.json" if peaks else "fallback") +
                            (" (sustained: the timed region lasts >= 1 s)" if long_region else " (burst)"),
                traffic_source="profiles/scan_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum of one "
                               "launch (ncu --set full)" if traffic is not None else None)

    # ---- the same launch timed ALONE (the way MEASURED_PEAKS.json's burst cuBLAS figure is taken: best of a few
    # launches, each after an idle pause, so the board is below its power cap): a note beside `frac`, which stays
    # on the mean over the timed region ----
    if headline:
        iso = []
        for _ in range(6):
            torch.cuda.synchronize()
            time.sleep(0.25)
            ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
            engine.search_device(Q, Q, [C], B, K_NN, events=ev1)
            torch.cuda.synchronize()
            iso.append(ev1[2].elapsed_time(ev1[3]))
        best = float(min(iso))
        roof["isolated_launch"] = {
            "launch_ms_best_of_6": best, "launch_ms_median": float(np.median(iso)),
            "tflops": flops / (best / 1e3) / 1e12, "frac_of_burst_peak": flops / (best / 1e3) / 1e12 / tf_burst,
            "how": "one search after 0.25 s of idle, CUDA events around the scan launch, best of 6 -- like for like "
                   "with the burst cuBLAS peak (best of 10, timed alone); `frac` above is the mean over the timed "
                   "region, where the board sits at its power cap"}

    # ---- BASELINE.json configs[3] and configs[4] on the same line (single GPU, same index): the candidate-budget
    # sweep (probed buckets -> recall@10, QPS, the scan's roofline) and the query-batch sweep (1 ... 100k queries:
    # device time of one graph replay, host-to-host latency of `search` with pinned host queries) ----
    sweep_d = sweep_e = None
    if headline and world == 1 and not args.no_sweeps:
        from learnedmetricindex_b200 import synthetic as syn

        def quick(Qx, Bx, steps=5):
            """(graph ms/step, mean scan launch ms, scan roofline dict) of searches of Qx with budget Bx."""
            nqx = Qx.shape[0]
            for _ in range(3):
                li.search_graphed(Qx, Qx, [C], Bx, K_NN)
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            a0.record()
            for _ in range(steps):
                li.search_graphed(Qx, Qx, [C], Bx, K_NN)
            a1.record()
            torch.cuda.synchronize()
            g_ms = a0.elapsed_time(a1) / steps
            evx = [[torch.cuda.Event(enable_timing=True) for _ in range(5)] for _ in range(steps)]
            li.search_device(Qx, Qx, [C], Bx, K_NN)
            for e in evx:
                li.search_device(Qx, Qx, [C], Bx, K_NN, events=e)
            torch.cuda.synchronize()
            sc_ms = float(np.mean([e[2].elapsed_time(e[3]) for e in evx]))
            cnt = li.last_plan["counts"].cpu().numpy().astype(np.int64)
            fl = 2.0 * D * float((cnt * sizes_local).sum())
            by = 2.0 * D * float(sizes_local[cnt > 0].sum()) + 2.0 * D * nqx * Bx + 8.0 * nqx * K_NN
            tt, th = fl / (tf_burst * 1e12), by / (hbm_peak * 1e9)
            rf = {"bound": "tensor" if tt >= th else "hbm", "roofline_ms": max(tt, th) * 1e3, "scan_ms": sc_ms,
                  "frac": max(tt, th) * 1e3 / sc_ms}
            return g_ms, sc_ms, rf

        try:
            sweep_d = []
            for Bs in (1, 2, 3, 4, 6, 8, 16, 48, 120):
                if Bs > C:
                    continue
                rs_ = recall_at_10(run_once(Bs)[1], gt)
                g_ms, sc_ms, rf = quick(Q, Bs)
                sweep_d.append({"probed_buckets": Bs, "recall_at_10": rs_, "qps": nq / (g_ms / 1e3), "ms_per_step": g_ms,
                                "scan": rf})
                log(f"sweep_d B={Bs}: recall {rs_:.4f} QPS {nq / (g_ms / 1e3):.0f} scan {sc_ms:.3f} ms "
                    f"({rf['bound']}, frac {rf['frac']:.3f})")
        except Exception as e:  # noqa: BLE001
            sweep_d = {"error": repr(e)}
        try:
            sweep_e = []
            Qbig = syn.make_vectors(100_000, w["spec"], device=device, stream=2)
            for nqx in (1, 10, 100, 1000, 10_000, 100_000):
                Qx = Qbig[:nqx].contiguous()
                g_ms, sc_ms, rf = quick(Qx, B)
                Qxh = torch.empty(Qx.shape, dtype=Qx.dtype).pin_memory()
                Qxh.copy_(Qx.cpu())
                lat = []
                for r_ in range(3 + 10):
                    t0 = time.perf_counter()
                    li.search(None, Qxh, None, Qxh, None, [C], n_buckets=B, k=K_NN)
                    lat.append(time.perf_counter() - t0)
                sweep_e.append({"nq": nqx, "probed_buckets": B, "device_ms_per_batch": g_ms, "qps_device": nqx / (g_ms / 1e3),
                                "host_to_host_ms_median": float(np.median(lat[3:]) * 1e3),
                                "qps_host_to_host": nqx / float(np.median(lat[3:])), "scan": rf})
                log(f"sweep_e nq={nqx}: device {g_ms:.3f} ms, host-to-host {np.median(lat[3:]) * 1e3:.3f} ms "
                    f"scan {sc_ms:.3f} ms ({rf['bound']}, frac {rf['frac']:.3f})")
            del Qbig
        except Exception as e:  # noqa: BLE001
            sweep_e = {"error": repr(e)}
        li.search_device(Q, Q, [C], B, K_NN)  # last_plan / last_bucket_order back on the headline batch
        torch.cuda.synchronize()

    # ---- cold call: the FIRST search() of a fresh index object through the reference-facing
    # signature with HOST data -- bucket layout build (the regrouping the reference pays on every call,
    # LearnedIndex.py:350,357), its H2D, weight splits, workspace allocation, graph capture and the search itself
    cold = None
    if headline and world == 1 and not args.no_cold_call:
        try:
            import psutil

            need = w["vectors"].numel() * 2
            if psutil.virtual_memory().available < 2.5 * need or torch.cuda.mem_get_info()[0] < 1.3 * need:
                cold = {"skipped": "not enough free host RAM / HBM for a second copy of the index"}
            else:
                Xh = w["vectors"].cpu().numpy()  # host matrix (bucket-major order here; any order works)
                pred_h = np.repeat(np.arange(C, dtype=np.int64), np.diff(w["bucket_off"]))[:, None]
                Qn = Qh.numpy()
                li_cold = LearnedIndex(w["layers"], {}, [(i,) for i in range(C)])
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                d_c, i_c, tm_c = li_cold.search(None, Qn, Xh, Qn, pred_h, [C], n_buckets=B, k=K_NN)
                cold_s = time.perf_counter() - t0
                t0 = time.perf_counter()
                d_w, _, _ = li_cold.search(None, Qn, Xh, Qn, pred_h, [C], n_buckets=B, k=K_NN)
                warm_s = time.perf_counter() - t0
                assert np.array_equal(d_c, d_e2e) and np.array_equal(d_w, d_e2e), "cold-call distances differ"
                cold = {"cold_call_s": cold_s, "second_call_s": warm_s, "host_bytes_uploaded": int(need),
                        "what": "first LearnedIndex.search(host data_search, data_prediction, host queries) of a new "
                                "index object: layout build + H2D + weight split + workspace + graph capture + search; "
                                "distances equal the timed steps'"}
                log(f"cold call {cold_s:.2f}s, second call {warm_s * 1e3:.2f} ms")
                del li_cold, Xh, pred_h
                import gc

                gc.collect()
                torch.cuda.empty_cache()
        except Exception as e:  # noqa: BLE001 -- the steady-state numbers stand
            cold = {"error": repr(e)}

    # ---- CPU baseline (rank 0, N == 1 only): the oracle port on the WHOLE batch, which doubles as the
    # asserted parity check of the timed GPU result at the headline size ----
    cpu = parity = None
    if headline and rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            t0 = time.time()
            cb = CpuBaseline(w, C)
            t_cpu, d_cpu, i_cpu, order_cpu = cb.run(B)
            parity = cb.explain(d_cpu, i_cpu, order_cpu, d_e2e, i_e2e, order_gpu)
            cpu = {"value": nq / t_cpu, "unit": "QPS", "cores": cb.threads, "kind": "port",
                   "sample": cb.describe(B, nq), "seconds": t_cpu, "host_cpus": host_threads(),
                   "id_agreement_with_gpu": parity["id_agreement"]}
            log(f"cpu baseline: {nq} queries in {t_cpu:.2f}s = {nq / t_cpu:.1f} QPS on {cb.threads} threads "
                f"(setup {time.time() - t0 - t_cpu:.1f}s); parity {parity}")
            del cb
        except Exception as e:  # the GPU numbers stand; say why the CPU leg is absent
            cpu = {"value": None, "unit": "QPS", "cores": 0, "kind": "port", "sample": f"failed: {e!r}"}

    # ---- config A only: the timed GPU result checked against ONE call of the UNMODIFIED reference
    # (li.LearnedIndex.search out of oracle/_ref, with oracle/faiss_shim for the absent faiss) on all queries --
    # the size at which the reference itself finishes in about a second -- and against the oracle port ----
    unmodified = None
    if check_unmodified and rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cb = CpuBaseline(w, C)
            t_port, d_p, i_p, order_p = cb.run(B)
            par_port = cb.explain(d_p, i_p, order_p, d_e2e, i_e2e, order_gpu)
            unmodified = {"oracle_port": {"seconds": t_port, "qps": nq / t_port, "cores": cb.threads,
                                          "parity": par_port}}
            if cb.unmodified_available():
                t_u, d_u, i_u, _ = cb.run_unmodified(B, limit_s=args.reference_limit)
                par_u = cb.explain(d_u, i_u.astype(np.uint32), order_p, d_e2e, i_e2e, order_gpu)
                unmodified["unmodified_reference"] = {
                    "seconds": t_u, "qps": nq / t_u, "kind": "reference", "parity": par_u,
                    "ids_equal_to_port": float((i_u.astype(np.uint32) == i_p).mean()),
                    "what": "one call of li.LearnedIndex.search (oracle/_ref, unmodified) on all queries"}
                parity = par_u if parity is None else parity
            else:
                unmodified["unmodified_reference"] = {"note": "oracle/_ref not staged (python -m oracle.build_ref)"}
                parity = par_port if parity is None else parity
            log(f"config A parity: {unmodified}")
            del cb
        except Exception as e:  # noqa: BLE001
            unmodified = {"error": repr(e)}

    # ---- the metric's other reading: the SMALLEST budget whose recall@10 is still >= 0.90;
    # the headline `value` stays on the README's budget (`--probe`, 4 of 120 buckets) ----
    min_point = None
    if headline:
        Bm, rm = B, recall
        while Bm > 1:
            r = recall_at_10(run_once(Bm - 1)[1], gt)
            if r < 0.90:
                break
            Bm, rm = Bm - 1, r
        if Bm < B:
            ks = min(K, 10)
            tm, _, _, _ = timed_device_loop(Bm, ks, 3, graph=True)
            min_point = {"probed_buckets": Bm, "recall_at_10": rm, "qps": nq / (tm / ks / 1e3), "ms_per_step": tm / ks}

    launches_per_step = getattr(li, "launches_per_search", 9) + (1 if world > 1 else 0)
    out = {
        "metric": "QPS at recall@10>=0.90", "value": value, "unit": "QPS", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f16 in / f32 accumulate (scan), tf32 x3 split / f32 accumulate (MLP)", "data": "synthetic",
        "config": bench_config(wl_name, args.n, nq, C, B, world),
        "config_detail": {
            "recall_at_10": recall,
            "sharding": (f"equal row slices of every bucket over {world} GPU(s)" if args.shard == "rows" and world > 1
                         else f"buckets over {world} GPU(s)"),
            "bucket_vectors_gb_this_gpu": w["vectors"].numel() * 2 / 1e9,
            "synthetic": {"centres": args.centres, "latent_dim": args.latent, "noise": args.noise}},
        "e2e": {"value": e2e_qps, "unit": "QPS", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_s / K * 1e3},
        "gpu_launches": launches_per_step * K,
        "host_enqueue_ms_per_step": host_ms_main,
        "execution": ("one CUDA graph replay per search" if can_graph else "eager launches (NCCL exchange)"),
        "eager_launches": {"ms_per_step": total_ms_eager / K, "host_enqueue_ms_per_step": host_ms_eager},
        "kernel_ms": parts,
        "kernel_ms_graph_last_step": parts_graph if can_graph else None,
        "roofline": roof,
        "cpu_baseline": cpu,
        "parity": parity,
        "multi_gpu_parity": multi,
        "clocks": clocks,
        "smallest_budget_with_recall_0.90": min_point,
        "cold_call": cold,
        "parity_vs_reference": unmodified,
        "sweep_d": sweep_d,
        "sweep_e": sweep_e,
    }
    if parity is not None and parity["unexplained"] > 0:
        # a fast kernel whose results differ from the reference's is not done: print the line, then fail
        if rank == 0:
            emit_json(out)
        raise SystemExit(f"parity with the CPU oracle FAILED at the headline size: {parity}")
    return out


def run_reference(args, wl_name, rank, world, local_rank):
    """The reference arm: the CPU implementation of the path on this box's host cores, on the SAME config
    (all ``nq`` queries per step, all host threads).  ``value`` is the oracle port -- the reference's
    arithmetic without its per-round pandas regrouping, i.e. the faster of the two CPU forms, so the
    GPU/CPU ratio is the conservative one; at N = 1 the line also carries ONE call of the unmodified
    reference ``LearnedIndex.search`` (oracle/_ref, staged by oracle/build_ref.py) on the same batch --
    25 such calls would not fit the driver's window at 10M -- and the port on one thread."""
    if rank != 0:
        return
    C, nq, B = args.buckets, args.nq, (args.probe if args.probe > 0 else 4)
    use_cuda = torch.cuda.is_available()
    device = torch.device("cuda", local_rank) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(local_rank)
    # setup only (data generation + stand-in index) may use the GPU through plain torch ops;
    # none of this repo's kernels run in this arm and the timed region is CPU only
    w = build_workload(args, device, 0, 1, use_repo_kernels=False)
    cb = CpuBaseline(w, C)
    del w
    if use_cuda:
        torch.cuda.empty_cache()
    K, W = args.steps, max(1, args.warmup)
    for _ in range(W):
        cb.run(B)
    t0 = time.perf_counter()
    for _ in range(K):
        _, d_port, i_port, _ = cb.run(B)
    total = time.perf_counter() - t0
    qps = nq * K / total
    sample = cb.describe(B, nq)
    log(f"reference arm (port): {K} steps x {nq} queries in {total:.1f}s = {qps:.1f} QPS on {cb.threads} threads")

    extra = {}
    if args.gpus == 1 and not args.no_reference_extras:
        # one thread: what the README's "1 core" row is comparable with
        try:
            t1, _, _, _ = cb.run_single_thread(B)
            extra["port_single_thread"] = {"value": nq / t1, "unit": "QPS", "cores": 1, "seconds": t1,
                                           "sample": f"all {nq} queries, 1 call"}
            log(f"reference arm (port, 1 thread): {t1:.1f}s = {nq / t1:.1f} QPS")
        except Exception as e:  # noqa: BLE001
            extra["port_single_thread"] = {"value": None, "note": f"failed: {e!r}"}
        # the unmodified reference on the same batch, once
        if cb.unmodified_available():
            try:
                t_ref, d_ref, i_ref, tm = cb.run_unmodified(B, limit_s=args.reference_limit)
                same = bool(np.array_equal(i_ref, i_port))
                extra["unmodified_reference"] = {
                    "value": nq / t_ref, "unit": "QPS", "kind": "reference", "cores": cb.threads, "seconds": t_ref,
                    "sample": f"all {nq} queries, 1 call of li.LearnedIndex.search (oracle/_ref, faiss.knn via "
                              f"oracle/faiss_shim: torch sgemm + topk), frames view the pre-widened f32 matrix",
                    "measured_time": {k_: float(v) for k_, v in tm.items()},
                    "ids_equal_to_port": same,
                    "max_abs_dist_diff_vs_port": float(np.nanmax(np.abs(np.where(np.isfinite(d_ref), d_ref - d_port, 0.0))))}
                log(f"reference arm (unmodified li.LearnedIndex.search): {t_ref:.1f}s = {nq / t_ref:.1f} QPS; "
                    f"ids equal to the port's: {same}")
            except Exception as e:  # noqa: BLE001 -- RAM guard / time limit: say so, the port's number stands
                extra["unmodified_reference"] = {"value": None, "kind": "reference", "note": f"not run: {e!r}"}
        else:
            extra["unmodified_reference"] = {"value": None, "kind": "reference",
                                             "note": "oracle/_ref not staged (python -m oracle.build_ref)"}
    out = {
        "impl": "reference", "metric": "QPS at recall@10>=0.90", "value": qps, "unit": "QPS", "n_gpus": args.gpus,
        "steps": K, "warmup": W, "ms_per_step": total / K * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": bench_config(wl_name, args.n, nq, C, B, args.gpus),
        "cpu_baseline": {"value": qps, "unit": "QPS", "cores": cb.threads, "kind": "port", "sample": sample,
                         "host_cpus": host_threads(), "host_ram_gib": round(cb.ram_gb, 1), **extra},
        "e2e": {"value": qps, "unit": "QPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_json(out)


if __name__ == "__main__":
    main()
