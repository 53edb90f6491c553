#!/usr/bin/env python
"""Accuracy and time of the two Linear-layer kernels (tcgen05 tf32 split vs fp32 FFMA) against a
float64 evaluation of the same network, plus top-B order agreement with the fp32 torch oracle."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from learnedmetricindex_b200 import NeuralNetwork, synthetic as syn

dev = torch.device("cuda", 0)
nq, C, B = 10_000, 120, 4
spec = syn.MixtureSpec(n_centres=2000)
X = syn.make_vectors(200_000, spec, device=dev, stream=2)
layers, _ = syn.build_standin_index(X, C, steps=300)
Q = syn.make_vectors(nq, spec, device=dev, stream=1)
layers_np = syn.layers_to_numpy(layers)
# float64 truth
h = Q.double().cpu()
for li, (W, b) in enumerate(layers_np):
    h = h @ torch.from_numpy(W).double().t() + torch.from_numpy(b).double()
    if li + 1 < len(layers_np):
        h = torch.relu(h)
truth = h.numpy()
# fp32 torch CPU evaluation, as model.py:232-239 does (Linear/ReLU stack, softmax, topk)
with torch.no_grad():
    h32 = Q.float().cpu()
    for li, (W, b) in enumerate(layers_np):
        h32 = torch.nn.functional.linear(h32, torch.from_numpy(W), torch.from_numpy(b))
        if li + 1 < len(layers_np):
            h32 = torch.relu(h32)
    probs_t, classes_t = torch.softmax(h32, dim=1).topk(C)
classes_cpu = classes_t.numpy()
scale = np.abs(truth).max()
for impl in ("ffma", "tc"):
    os.environ["LMI_MLP_IMPL"] = impl
    net = NeuralNetwork(layers, device=dev)
    for qdtype in (torch.float16, torch.float32):
        q = Q.to(qdtype).contiguous()
        out = net.logits(q)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            out = net.logits(q)
        e1.record(); torch.cuda.synchronize()
        err = np.abs(out.double().cpu().numpy() - truth)
        order, _ = net.topn_device(q, B)
        same = (order.cpu().numpy() == classes_cpu[:, :B]).all(axis=1)
        print(f"{impl:5s} x={str(qdtype)[6:]:8s} {e0.elapsed_time(e1)/20*1e3:7.1f} us/forward  max|err|={err.max():.3e} "
              f"mean|err|={err.mean():.3e} (logit scale {scale:.1f}; fp32 eps*scale={scale*6e-8:.1e})  "
              f"top-{B} rows equal to torch-CPU fp32: {same.mean():.5f} ({(~same).sum()} differ)")
err = np.abs(h32.double().numpy() - truth)
print(f"torch CPU fp32: max|err|={err.max():.3e} mean|err|={err.mean():.3e}")
