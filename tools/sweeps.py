#!/usr/bin/env python
"""BASELINE.json configs 4 and 5 on one GPU: the candidate-budget sweep (probed buckets B -> QPS vs
recall@10) and the query-batch sweep (nq = 1 ... 100k) on the 10M x 768 index.  Device-timed
(CUDA events, index and queries resident), one JSON object per line on stdout."""
import argparse, json, os, sys, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=10_000_000)
ap.add_argument("--probes", default="1,2,3,4,6,8,12,16,24,48,120")
ap.add_argument("--batches", default="1,10,100,1000,10000,100000")
ap.add_argument("--steps", type=int, default=10)
a = ap.parse_args()
C = 120
nq_max = max(10_000, max(int(x) for x in a.batches.split(",")))
args = types.SimpleNamespace(n=a.n, nq=nq_max, buckets=C, centres=2000, noise=0.25, latent=24, train_steps=300)
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
from learnedmetricindex_b200 import DeviceBucketLayout, LearnedIndex
w = bench.build_workload(args, dev, 0, 1)
li = LearnedIndex(w["layers"], {}, [(i,) for i in range(C)])
li.attach_layout(DeviceBucketLayout(w["vectors"], torch.from_numpy(w["bucket_off"]), w["ids"]))
Qall = w["Q"]
gt = bench.ground_truth(Qall[:10_000], w["vectors"], w["ids"], 1)
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
hbm, tfs = float(peaks.get("hbm_gbs", 6650.0)), float(peaks.get("bf16_tflops_sustained", 1400.0))
sizes = w["sizes"].astype(np.int64)


def run(nq, B, with_recall):
    Q = Qall[:nq].contiguous()
    for _ in range(3):
        li.search_device(Q, Q, [C], B, 10)
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(5)] for _ in range(a.steps)]
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for e in evs:
        d, i = li.search_device(Q, Q, [C], B, 10, events=e)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    scan = float(np.mean([e[2].elapsed_time(e[3]) for e in evs]))
    counts = li.last_plan["counts"].cpu().numpy().astype(np.int64)
    flops = 2.0 * 768 * float((counts * sizes).sum())
    byts = 2.0 * 768 * float(sizes[counts > 0].sum()) + 2.0 * 768 * nq * B + 8.0 * nq * 10
    t_t, t_h = flops / (tfs * 1e12), byts / (hbm * 1e9)
    out = {"nq": nq, "B": B, "ms_per_step": ms, "qps": nq / ms * 1e3, "scan_ms": scan,
           "bound": "tensor" if t_t >= t_h else "hbm", "roofline_ms": max(t_t, t_h) * 1e3,
           "scan_frac_of_roofline": max(t_t, t_h) * 1e3 / scan}
    if with_recall:
        out["recall_at_10"] = bench.recall_at_10(i.cpu().numpy().view(np.uint32), gt[:nq])
    print(json.dumps(out), flush=True)


for B in [int(x) for x in a.probes.split(",") if x]:
    run(10_000, B, True)
for nq in [int(x) for x in a.batches.split(",") if x]:
    run(nq, 4, nq <= 10_000)


def latency(nq, B=4, reps=30):
    """Host-to-host latency of `LearnedIndex.search` (pinned host queries in, numpy out): one CUDA
    graph replay vs the eager launches."""
    import time
    Qh = Qall[:nq].cpu().pin_memory()
    out = {"nq": nq, "B": B, "what": "host-to-host search latency"}
    for mode, flag in (("graph_ms", True), ("eager_ms", False)):
        li.use_graphs = flag
        ts = []
        for r in range(5 + reps):
            t0 = time.perf_counter()
            li.search(None, Qh, None, Qh, None, [C], n_buckets=B, k=10)
            ts.append(time.perf_counter() - t0)
        out[mode] = float(np.median(ts[5:]) * 1e3)
    li.use_graphs = True
    print(json.dumps(out), flush=True)


for nq in [int(x) for x in a.batches.split(",") if x]:
    if nq <= 4096:
        latency(nq)
