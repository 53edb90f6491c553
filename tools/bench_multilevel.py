"""Measurement of the multi-level path -- the reference CLI's default shape
(`--n-categories 10 10`, search/search.py:314) in the notebook's arrangement: a low-dimensional
navigation representation (PCA of the 768-d vectors, search/search.py:192-216) routes through a
2-level tree of `MLP` models, the scan runs on the 768-d vectors.

    python tools/bench_multilevel.py [--n 1000000] [--nq 10000] [--probe 10] [--nav-dim 96]

Prints one JSON line: device-timed QPS, end-to-end QPS through `LearnedIndex.search` with host
queries, per-stage milliseconds, recall@10, parity of the visit order and ids against the CPU
oracle on a query sample, and the oracle's own QPS on the host cores (`cpu_baseline`).
Not the driver's bench (that is `bench.py`, the 1-level BASELINE configs).
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from learnedmetricindex_b200 import LearnedIndex  # noqa: E402
from learnedmetricindex_b200 import synthetic as syn  # noqa: E402
from learnedmetricindex_b200.layout import DeviceBucketLayout  # noqa: E402
from learnedmetricindex_b200.navigation import flat_bucket_ids  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--nq", type=int, default=10_000)
    ap.add_argument("--cats", type=int, nargs=2, default=[10, 10])
    ap.add_argument("--probe", type=int, default=10)
    ap.add_argument("--nav-dim", type=int, default=96)
    ap.add_argument("--model-type", default="MLP")
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--cpu-queries", type=int, default=2000)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    cats, B = list(args.cats), args.probe
    log = lambda m: print(f"[ml-bench] {m}", file=sys.stderr, flush=True)

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
    root, internal, pred, bucket_paths = syn.build_standin_index_two_level(Xn, cats, args.model_type)
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

    # CPU oracle on a query sample (test infrastructure; here only as checker and baseline)
    from oracle import lmi_oracle as orc

    nc = min(args.cpu_queries, args.nq)
    Xh = X.cpu().numpy()
    off = li._layout.bucket_off_host
    lay = orc.BucketMajor(Xh.astype(np.float32), pred.cpu().numpy())
    idx = orc.OracleIndex(to_np(root), {p: to_np(l) for p, l in internal.items()}, bucket_paths)
    t0 = time.time()
    exp_d, exp_i, _, exp_order = orc.search(idx, Qn_h[:nc], None, Q_h[:nc], None, cats, n_buckets=B, k=10, layout=lay)
    cpu_s = time.time() - t0
    got_paths, _ = li._precompute_bucket_order(Qn_h[:nc], B, cats)
    order_same = float((got_paths == exp_order).all(axis=(1, 2)).mean())
    same = (got_paths == exp_order).all(axis=(1, 2))
    id_agree = float((i_h[:nc][same] == exp_i[same]).mean())

    print(json.dumps({
        "metric": "QPS, 2-level LMI", "value": args.nq / ms * 1e3, "unit": "QPS", "ms_per_step": ms,
        "config": {"workload": f"{args.n} x 768 fp16, n_categories {cats}, {args.model_type} models on a "
                               f"{args.nav_dim}-d PCA navigation representation, {args.nq} queries, "
                               f"{B} of {len(bucket_paths)} buckets, k=10", "recall_at_10": recall},
        "kernel_ms": {"navigation": stage[0], "invert": stage[1], "scan": stage[2], "merge": stage[3]},
        "e2e": {"value": args.nq / e2e_ms * 1e3, "unit": "QPS", "ms_per_step": e2e_ms},
        "cpu_baseline": {"value": nc / cpu_s, "unit": "QPS", "kind": "port", "cores": torch.get_num_threads(),
                         "sample": f"first {nc} queries, oracle/lmi_oracle.py, vectors pre-widened to f32"},
        "parity_vs_oracle": {"visit_order_rows_equal": order_same, "id_agreement_on_equal_rows": id_agree},
    }))


if __name__ == "__main__":
    main()
