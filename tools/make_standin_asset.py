#!/usr/bin/env python
"""Writes bench_assets/standin_mlp4_768_120.npz: the weights of bench.py's stand-in index for the default
mixture (MixtureSpec(n_centres=2000, noise=0.25, latent_dim=24), 120 buckets, MLP-4), trained ONCE on the CPU
(deterministic) from a 200k-row CPU draw of the same mixture.  bench.py loads it instead of re-training on
the GPU, so every run -- any box, any --gpus N, either arm -- searches the SAME index (GPU training was not
bit-reproducible between boxes: bucket sizes and recall moved in the fourth digit between SCALE runs)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from learnedmetricindex_b200 import synthetic as syn

torch.set_num_threads(len(os.sched_getaffinity(0)))
C, D, TRAIN_ROWS, STEPS = 120, 768, 200_000, 300
spec = syn.MixtureSpec(n_centres=2000, noise=0.25, latent_dim=24)
t0 = time.time()
xt = syn.make_vectors(TRAIN_ROWS, spec, device="cpu", stream=2).float()
g = torch.Generator(device="cpu").manual_seed(syn.SEED)
cent = syn._kmeans_centroids(xt, C, 10, g)
labels = (xt @ cent.t()).argmax(dim=1)
print(f"k-means {time.time() - t0:.1f}s; label counts min/max {torch.bincount(labels).min()}/{torch.bincount(labels).max()}", flush=True)
layers = syn.train_standin_mlp(xt, labels, C, "MLP-4", steps=STEPS)
pred = syn.torch_mlp_argmax(layers, xt.half())
acc = (pred == labels).float().mean().item()
print(f"trained in {time.time() - t0:.1f}s; train accuracy vs k-means labels {acc:.4f}; classes predicted {pred.unique().numel()}")
out = os.path.join(ROOT, "bench_assets", "standin_mlp4_768_120.npz")
np.savez_compressed(out, W1=layers[0][0].numpy(), b1=layers[0][1].numpy(), W2=layers[1][0].numpy(), b2=layers[1][1].numpy(),
                    meta=np.array([C, D, TRAIN_ROWS, STEPS, spec.n_centres, spec.latent_dim], dtype=np.int64),
                    noise=np.float64(spec.noise))
print(out, os.path.getsize(out))
