#!/usr/bin/env python
"""Scan-kernel ablations on the headline workload (profiling aid, not a benchmark; results are WRONG
when an ablation is set).  Needs the profile library (tools/build_profile_lib.sh); every configuration
runs `--steps` back-to-back searches (sustained, i.e. under the board's power cap) with the SM clock and
board power sampled, so the table gives the ENERGY per step of the kernel with parts switched off:

  LMI_SCAN_ABLATE bits: 1 no TMA loads, 2 no candidate processing, 8 MMA-warp wait cycles, 16 stages
  pre-filled with pseudo-random fp16 (MMAs without loads then multiply real data), 32 loads confined to a
  50 MB window (L2 -> SM traffic as usual, nothing from HBM)

    tools/build_profile_lib.sh && python tools/scan_ablate.py --ablate 0,2,17,32,34
"""
import argparse, json, os, sys, time, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench
from learnedmetricindex_b200 import _native

# the product package knows nothing about the profile build: point its loader at it before the first lib() call
PROFILE_LIB = os.path.join(ROOT, "tools", "liblmi_b200_profile.so")
assert os.path.isfile(PROFILE_LIB), "build tools/liblmi_b200_profile.so first (tools/build_profile_lib.sh)"
_native.LIB_PATH = PROFILE_LIB

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=10_000_000)
ap.add_argument("--nq", type=int, default=10_000)
ap.add_argument("--probe", type=int, default=4)
ap.add_argument("--ablate", default="0,2,17,32,34,0")
ap.add_argument("--impls", default="0")
ap.add_argument("--steps", type=int, default=150)
ap.add_argument("--out", default="")
a = ap.parse_args()
args = types.SimpleNamespace(n=a.n, nq=a.nq, buckets=120, centres=2000, noise=0.25, latent=24, train_steps=300)
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
from learnedmetricindex_b200 import DeviceBucketLayout, LearnedIndex
w = bench.build_workload(args, dev, 0, 1)
layout = DeviceBucketLayout(w["vectors"], torch.from_numpy(w["bucket_off"]), w["ids"])
Q = w["Q"]
rows = []
for impl in [int(x) for x in a.impls.split(",")]:
    li = LearnedIndex(w["layers"], {}, [(i,) for i in range(120)], scan_impl=impl)
    li.attach_layout(layout)
    for ab in [int(x) for x in a.ablate.split(",")]:
        os.environ["LMI_SCAN_ABLATE"] = str(ab)
        for _ in range(5):
            li.search_device(Q, Q, [120], a.probe, 10)
        torch.cuda.synchronize()
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(5)] for _ in range(a.steps)]
        smp = bench.ClockSampler(0, period_s=0.005)
        smp.start()
        t_a = time.time()
        for e in evs:
            li.search_device(Q, Q, [120], a.probe, 10, events=e)
        torch.cuda.synchronize()
        t_b = time.time()
        smp.stop()
        scan = np.array([e[2].elapsed_time(e[3]) for e in evs])
        tail = scan[len(scan) // 3:]  # the power cap has settled by then
        sel = [s for s in smp.samples if t_a + (t_b - t_a) / 3 <= s[0] <= t_b]
        mhz = float(np.median([s[1] for s in sel])) if sel else None
        watts = float(np.mean([s[3] for s in sel])) if sel else None
        row = dict(impl=impl, ablate=ab, scan_ms_settled=float(tail.mean()), scan_ms_first5=float(scan[:5].mean()),
                   sm_mhz=mhz, watts=watts, joule_per_scan=(watts * float(tail.mean()) / 1e3 if watts else None),
                   mcycles=(mhz * float(tail.mean()) / 1e3 if mhz else None))
        if ab & 8:
            import ctypes
            from learnedmetricindex_b200 import _native
            buf = (ctypes.c_longlong * (148 * 8))()
            _native.lib().lmi_debug_scan_profile.argtypes = [ctypes.c_void_p]
            _native.lib().lmi_debug_scan_profile(buf)
            pr = np.array(buf[:]).reshape(148, 8)
            names = ["total", "sched", "q_ready", "acc_empty", "full", "issue", "stages"]
            row["mma_warp_cycles"] = {n: int(pr[:, i].mean()) for i, n in enumerate(names)}
        rows.append(row)
        print(json.dumps(row), flush=True)
os.environ["LMI_SCAN_ABLATE"] = "0"
if a.out:
    with open(a.out, "w") as f:
        for r in rows:
            f.write(json.dumps(r) + "\n")
