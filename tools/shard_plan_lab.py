#!/usr/bin/env python
"""One rank's scan of a bucket-sharded 10M index, emulated on ONE GPU: the full index is built, restricted to the
buckets rank r of `world` would own (size^2 LPT, as bench.py shards), and the whole query batch is searched with
different work-item cuts (LMI_RANGE_ROWS / LMI_TAIL_ROWS / LMI_TAIL_PERMILLE; "auto" = plan_scan's own choice).
Prints one JSON line per (rank, setting): scan launch ms (CUDA events, mean of --steps after warm-up), the plan,
and the ideal = this rank's share of the single-GPU scan."""
import argparse, json, os, sys, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--shards", default="8:0;8:3;4:0;2:0", help="world:rank;... emulated on this GPU")
ap.add_argument("--full-nq", default="10000,1000,100", help="batch sizes searched on the FULL index")
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--steps", type=int, default=30)
ap.add_argument("--settings", default="div=6;div=4;div=3;div=2;div=1.5;div=1")
a = ap.parse_args()
C, B = 120, 4
args = types.SimpleNamespace(n=10_000_000, nq=10_000, buckets=C, centres=2000, noise=0.25, latent=24, train_steps=300)
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
from learnedmetricindex_b200 import DeviceBucketLayout, LearnedIndex
from learnedmetricindex_b200.sharded import partition_buckets
w = bench.build_workload(args, dev, 0, 1)
full = DeviceBucketLayout(w["vectors"], torch.from_numpy(w["bucket_off"]), w["ids"])
Q = w["Q"]
sizes = w["sizes"].astype(np.int64)


def scan_ms(layout, label):
    li = LearnedIndex(w["layers"], {}, [(i,) for i in range(C)])
    li.attach_layout(layout)
    for _ in range(5):
        li.search_device(Q, Q, [C], B, 10)
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(5)] for _ in range(a.steps)]
    torch.cuda.synchronize()
    for e in evs:
        li.search_device(Q, Q, [C], B, 10, events=e)
    torch.cuda.synchronize()
    st = {n: float(np.mean([e[i].elapsed_time(e[i + 1]) for e in evs])) for i, n in enumerate(("nav", "invert", "scan", "merge"))}
    counts = li.last_plan["counts"].cpu().numpy().astype(np.int64)
    loc = np.diff(layout.bucket_off_host)
    work = float((np.ceil(counts / 128.0) * loc).sum())
    plan = {k: li.last_plan[k] for k in ("range_len", "tail_range_len", "tail_permille", "rmax", "n_items")}
    plan["n_items"] = int(plan["n_items"][0].item())
    return st, work, plan


KNOBS = ("LMI_RANGE_ROWS", "LMI_TAIL_ROWS", "LMI_TAIL_PERMILLE", "LMI_RANGE_DIV")


def apply(setting):
    for k in KNOBS:
        os.environ.pop(k, None)
    if setting == "auto":
        return
    if setting.startswith("div="):
        os.environ["LMI_RANGE_DIV"] = setting[4:]
        return
    rr, tr, pm = setting.split(",")
    os.environ["LMI_RANGE_ROWS"] = rr
    if int(tr):
        os.environ["LMI_TAIL_ROWS"], os.environ["LMI_TAIL_PERMILLE"] = tr, pm
    else:
        os.environ["LMI_TAIL_PERMILLE"] = "0"


apply("auto")
st_full, work_full, plan_full = scan_ms(full, "full")
print(json.dumps({"what": "single GPU, full index", **st_full, "work_row_tiles": work_full, "plan": plan_full}), flush=True)
cases = [("full", None, None, nq) for nq in [int(x) for x in a.full_nq.split(",") if x]]
for wr in a.shards.split(";"):
    if wr:
        world, r = [int(x) for x in wr.split(":")]
        cases.append(("shard", world, r, 10_000))
Qall = Q
for kind, world, r, nq in cases:
    Q = Qall[:nq].contiguous()
    layout = full if kind == "full" else full.restrict_to_buckets(
        partition_buckets(sizes, world, weights=sizes.astype(np.float64))[r])
    for rep in range(a.reps):
        for setting in a.settings.split(";"):
            apply(setting)
            try:
                st, work, plan = scan_ms(layout, setting)
                print(json.dumps({"kind": kind, "world": world, "rank": r, "nq": nq, "rep": rep, "setting": setting, **st,
                                  "work_row_tiles": work, "plan": plan}), flush=True)
            except Exception as e:  # noqa: BLE001
                print(json.dumps({"kind": kind, "world": world, "rank": r, "nq": nq, "setting": setting, "error": repr(e)}), flush=True)
    del layout
