#!/usr/bin/env python
"""Scan-kernel ablations on the headline workload (profiling aid, not a benchmark):
times lmi_scan_topk with parts of the kernel switched off via LMI_SCAN_ABLATE
(1 = no TMA loads, 2 = no candidate processing) and for several range lengths."""
import argparse, os, sys, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=10_000_000)
ap.add_argument("--nq", type=int, default=10_000)
ap.add_argument("--probe", type=int, default=4)
ap.add_argument("--ranges", default="16384")
ap.add_argument("--ablate", default="0,1,2,3")
a = ap.parse_args()
args = types.SimpleNamespace(n=a.n, nq=a.nq, buckets=120, centres=2000, noise=0.25, latent=24, train_steps=300)
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
from learnedmetricindex_b200 import DeviceBucketLayout, LearnedIndex
w = bench.build_workload(args, dev, 0, 1)
li = LearnedIndex(w["layers"], {}, [(i,) for i in range(120)])
li.attach_layout(DeviceBucketLayout(w["vectors"], torch.from_numpy(w["bucket_off"]), w["ids"]))
Q = w["Q"]
for R in [int(x) for x in a.ranges.split(",")]:
    os.environ["LMI_RANGE_ROWS"] = str(R)
    for ab in [int(x) for x in a.ablate.split(",")]:
        os.environ["LMI_SCAN_ABLATE"] = str(ab)
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(5)] for _ in range(20)]
        for _ in range(5):
            li.search_device(Q, Q, [120], a.probe, 10)
        torch.cuda.synchronize()
        for e in evs:
            li.search_device(Q, Q, [120], a.probe, 10, events=e)
        torch.cuda.synchronize()
        scan = np.mean([e[2].elapsed_time(e[3]) for e in evs])
        if ab & 8:
            import ctypes
            from learnedmetricindex_b200 import _native
            buf = (ctypes.c_longlong * (148 * 8))()
            _native.lib().lmi_debug_scan_profile.argtypes = [ctypes.c_void_p]
            _native.lib().lmi_debug_scan_profile(buf)
            pr = np.array(buf[:]).reshape(148, 8)
            names = ["total", "sched", "q_ready", "acc_empty", "full", "issue", "stages"]
            print("   MMA-warp cycles (mean over CTAs):", {n: int(pr[:, i].mean()) for i, n in enumerate(names)},
                  "total min/max", pr[:, 0].min(), pr[:, 0].max(), "stages min/max", pr[:, 6].min(), pr[:, 6].max())
        import time
        smp = bench.ClockSampler(0, period_s=0.005); smp.start(); t_a = time.time()
        for _ in range(150):
            li.search_device(Q, Q, [120], a.probe, 10)
        torch.cuda.synchronize(); t_b = time.time(); smp.stop()
        print("   clocks over %.2f s of back-to-back steps:" % (t_b - t_a), smp.summary([(t_a + 0.3, t_b)]))
        print(f"range_rows={R} ablate={ab}: scan {scan:.3f} ms  n_items={int(li.last_plan['n_items'][0])}", flush=True)
os.environ["LMI_SCAN_ABLATE"] = "0"
