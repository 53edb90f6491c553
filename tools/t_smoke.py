"""Quick check of the tcgen05 scan kernel against the CUDA-core verification kernel on a few small
shapes (development aid; seconds on a GPU box -- run it under `timeout` before the full test suite
when changing the scan kernel's barrier protocol)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from learnedmetricindex_b200 import LearnedIndex
from tests.golden_util import seeded_layers

for (n, nq, d, C, B) in [(3000, 40, 128, 4, 2), (20000, 100, 768, 3, 2), (700, 9, 64, 5, 5), (50000, 300, 192, 4, 3)]:
    rs = np.random.RandomState(n)
    X = rs.standard_normal((n, d)).astype(np.float32); X = (X / np.linalg.norm(X, axis=1, keepdims=True)).astype(np.float16)
    Q = rs.standard_normal((nq, d)).astype(np.float32); Q = (Q / np.linalg.norm(Q, axis=1, keepdims=True)).astype(np.float16)
    layers = seeded_layers(n, "MLP-8", d, C, 4.0)
    pred = rs.randint(0, C, n).astype(np.int64).reshape(-1, 1)
    out = {}
    for impl in (1, 0):
        li = LearnedIndex(layers, {}, [(i,) for i in range(C)], scan_impl=impl)
        li.use_graphs = False
        dd, ii, _ = li.search(None, Q, X, Q, pred, [C], n_buckets=B, k=10)
        torch.cuda.synchronize()
        out[impl] = (dd, ii)
        print("done", (n, nq, d, C, B), "impl", impl, flush=True)
    same_i = (out[0][1] == out[1][1]).mean()
    err = np.abs(out[0][0] - out[1][0]).max()
    print("RESULT", (n, nq, d, C, B), "id agreement", same_i, "max |dd|", err, flush=True)
