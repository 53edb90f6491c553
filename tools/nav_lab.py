"""Navigation lab: the one-launch lmi_mlp_topn against the per-layer path (split + 2 x linear_tc + softmax/top-n)
on the MLP-4 shape of the headline config; CUDA-event means over back-to-back launches.
    python tools/nav_lab.py [--nq 1,128,1000,10000] [--reps 50] [--once fused|tc]   (--once: one launch, for ncu)"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from learnedmetricindex_b200 import NeuralNetwork  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--nq", default="1,128,1000,10000,100000")
ap.add_argument("--reps", type=int, default=50)
ap.add_argument("--B", type=int, default=4)
ap.add_argument("--once", default="")
args = ap.parse_args()
rs = np.random.RandomState(0)
d, H, C = 768, 512, 120
layers = [(rs.randn(H, d).astype(np.float32) / np.sqrt(d), rs.randn(H).astype(np.float32) * 0.1),
          (rs.randn(C, H).astype(np.float32) * (3 / np.sqrt(H)), rs.randn(C).astype(np.float32) * 0.1)]
nets = {}
for impl in ("fused", "tc"):
    nets[impl] = NeuralNetwork(layers)
    nets[impl].impl = impl
for nq in [int(v) for v in args.nq.split(",")]:
    x = torch.from_numpy(rs.randn(nq, d).astype(np.float16)).cuda()
    if args.once:
        nets[args.once].topn_device(x, args.B)
        nets[args.once].topn_device(x, args.B)
        torch.cuda.synchronize()
        continue
    row = {"nq": nq}
    for impl, net in nets.items():
        for _ in range(5):
            net.topn_device(x, args.B)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            o, p = net.topn_device(x, args.B)
        e1.record()
        torch.cuda.synchronize()
        row[impl + "_us"] = round(e0.elapsed_time(e1) / args.reps * 1e3, 2)
    print(json.dumps(row), flush=True)
