"""Synthetic LAION/CLIP-shaped data and a fast stand-in for the offline index build.

Harness code shared by ``tests/`` and ``bench.py`` (there is no network for the
SISAP'23 LAION files the reference downloads, search/search.py:38-48).  Device
agnostic: runs on CPU for the unit tests and on ``cuda`` for the 10M / 100M
benchmark configurations, where the data never exists on the host.

* ``make_vectors``: 768-d, L2-normalised (search/search.py:142-144), stored as
  fp16 (SISAP ``emb`` is fp16 on disk).  Points lie near a low-dimensional
  Gaussian-mixture manifold so that a query's true neighbours straddle bucket
  borders and recall@10 is a real function of the number of probed buckets.
* ``build_standin_index``: same contract as ``LearnedIndexBuilder.build``
  (search/li/LearnedIndexBuilder.py:43-107) at a fraction of the cost: k-means
  labels -> train the ``MLP-4``-shaped network -> ``data_prediction =
  argmax MLP(x)`` (:76).  Training is out of the hot path's scope;
  it only has to yield *some* trained weights plus the bucket assignment.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Tuple

import numpy as np
import torch

SEED = 2023  # the reference's seed (search/li/LearnedIndex.py:18-19)


@dataclass
class MixtureSpec:
    dim: int = 768
    latent_dim: int = 24
    n_centres: int = 2000
    centre_spread: float = 2.0
    noise: float = 0.25
    seed: int = SEED

    def materialise(self, device="cpu"):
        g = torch.Generator(device="cpu").manual_seed(self.seed)
        centres = torch.randn(self.n_centres, self.latent_dim, generator=g) * self.centre_spread
        lift = torch.randn(self.latent_dim, self.dim, generator=g) / math.sqrt(self.latent_dim)
        return centres.to(device), lift.to(device)


def make_vectors(
    n: int,
    spec: MixtureSpec = MixtureSpec(),
    device="cpu",
    stream: int = 0,
    chunk: int = 1 << 20,
    out: Optional[torch.Tensor] = None,
) -> torch.Tensor:
    """``n`` fp16 unit vectors drawn from ``spec``; ``stream`` separates data (0) from
    queries (1) so queries are fresh draws, not dataset members."""
    centres, lift = spec.materialise(device)
    if out is None:
        out = torch.empty((n, spec.dim), dtype=torch.float16, device=device)
    g = torch.Generator(device=device).manual_seed(spec.seed * 1000003 + stream)
    for s in range(0, n, chunk):
        m = min(chunk, n - s)
        which = torch.randint(0, spec.n_centres, (m,), generator=g, device=device)
        z = centres[which] + torch.randn(m, spec.latent_dim, generator=g, device=device)
        x = z @ lift + spec.noise * torch.randn(m, spec.dim, generator=g, device=device)
        x = torch.nn.functional.normalize(x, dim=1)
        out[s : s + m] = x.to(torch.float16)
    return out


def _kmeans_centroids(x: torch.Tensor, k: int, iters: int, g: torch.Generator) -> torch.Tensor:
    idx = torch.randperm(x.shape[0], generator=g, device=x.device)[:k]
    c = x[idx].clone()
    rows = torch.arange(x.shape[0], device=x.device)
    for _ in range(iters):
        lab = (x @ c.t()).argmax(dim=1)  # unit vectors: max inner product == nearest
        # segment sums as a one-hot GEMM: bit-reproducible run to run (index_add_ adds with float
        # atomics on CUDA, which made the stand-in index -- bucket sizes, recall -- differ between runs)
        onehot = torch.zeros((k, x.shape[0]), dtype=x.dtype, device=x.device)
        onehot[lab, rows] = 1
        sums = onehot @ x
        del onehot
        cnt = torch.bincount(lab, minlength=k).to(x.dtype).unsqueeze(1)
        c = torch.where(cnt > 0, torch.nn.functional.normalize(sums / cnt.clamp(min=1), dim=1), c)
    return c


def mlp_layer_shapes(model_type: str, in_dim: int, out_dim: int) -> List[Tuple[int, int]]:
    """Hidden widths of the reference model zoo -- search/li/model.py:29-79."""
    hidden = {
        "MLP": [128], "MLP-2": [64], "MLP-3": [256], "MLP-4": [512], "MLP-5": [256, 128],
        "MLP-6": [32], "MLP-7": [16], "MLP-8": [8], "MLP-9": [8, 16],
    }
    if model_type not in hidden:
        raise ValueError(f"Model type {model_type} not supported.")
    dims = [in_dim] + hidden[model_type] + [out_dim]
    return [(dims[i + 1], dims[i]) for i in range(len(dims) - 1)]


def train_standin_mlp(
    x_train: torch.Tensor, labels: torch.Tensor, n_classes: int, model_type: str = "MLP-4",
    steps: int = 300, lr: float = 0.01, batch: int = 8192, seed: int = SEED,
) -> List[Tuple[torch.Tensor, torch.Tensor]]:
    """Mini-batch Adam on cross-entropy (loss/optimiser as search/li/model.py:155-164).
    Returns [(W f32[out,in], b f32[out]), ...] on ``x_train.device``."""
    dev = x_train.device
    torch.manual_seed(seed)
    shapes = mlp_layer_shapes(model_type, x_train.shape[1], n_classes)
    mods: List[torch.nn.Module] = []
    for i, (o, ii) in enumerate(shapes):
        mods.append(torch.nn.Linear(ii, o))
        if i + 1 < len(shapes):
            mods.append(torch.nn.ReLU())
    net = torch.nn.Sequential(*mods).to(dev)
    opt = torch.optim.Adam(net.parameters(), lr=lr)
    loss_fn = torch.nn.CrossEntropyLoss()
    g = torch.Generator(device=dev).manual_seed(seed)
    n = x_train.shape[0]
    for _ in range(steps):
        sel = torch.randint(0, n, (min(batch, n),), generator=g, device=dev)
        loss = loss_fn(net(x_train[sel].float()), labels[sel])
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
    return [
        (m.weight.detach().float().contiguous(), m.bias.detach().float().contiguous())
        for m in net if isinstance(m, torch.nn.Linear)
    ]


def torch_mlp_argmax(layers, x: torch.Tensor, chunk: int = 1 << 18) -> torch.Tensor:
    """``NeuralNetwork.predict`` (search/li/model.py:213-224) with plain torch ops."""
    out = torch.empty(x.shape[0], dtype=torch.int64, device=x.device)
    with torch.no_grad():
        for s in range(0, x.shape[0], chunk):
            h = x[s : s + chunk].float()
            for i, (W, b) in enumerate(layers):
                h = torch.nn.functional.linear(h, W, b)
                if i + 1 < len(layers):
                    h = torch.relu(h)
            out[s : s + chunk] = h.argmax(dim=1)
    return out


def build_standin_index(
    x: torch.Tensor, n_buckets: int, model_type: str = "MLP-4", train_rows: int = 200_000,
    kmeans_iters: int = 8, steps: int = 300, lr: float = 0.01, seed: int = SEED,
):
    """Returns ``(layers, data_prediction int64[n,1])``; tensors stay on ``x.device``.

    1-level contract of ``LearnedIndexBuilder.build`` (LearnedIndexBuilder.py:67-88):
    bucket membership is the trained model's own argmax, not the k-means label."""
    dev = x.device
    g = torch.Generator(device=dev).manual_seed(seed)
    n = x.shape[0]
    if n > train_rows:
        sel = torch.randperm(n, generator=g, device=dev)[:train_rows]
        xt = x[sel].float()
    else:
        xt = x.float()
    cent = _kmeans_centroids(xt, n_buckets, kmeans_iters, g)
    labels = (xt @ cent.t()).argmax(dim=1)
    # cuBLAS TF32 must not leak into how the stand-in assigns buckets vs how tests re-check it
    layers = train_standin_mlp(xt, labels, n_buckets, model_type, steps=steps, lr=lr, seed=seed)
    pred = torch_mlp_argmax(layers, x)
    return layers, pred.reshape(-1, 1)


def layers_to_numpy(layers):
    return [(W.detach().cpu().numpy().astype(np.float32), b.detach().cpu().numpy().astype(np.float32)) for W, b in layers]


def build_standin_index_two_level(
    x_nav: torch.Tensor, n_categories, model_type: str = "MLP", train_rows: int = 200_000,
    kmeans_iters: int = 8, steps: int = 200, lr: float = 0.01, seed: int = SEED,
):
    """2-level contract of ``LearnedIndexBuilder.build`` (LearnedIndexBuilder.py:67-107, :203-280):
    root model over all rows, one internal model per level-1 class trained on the rows the root
    routes there, ``data_prediction[:, 1]`` = that model's argmax, ``bucket_paths`` =
    ``(j, 0..len(unique(predictions))-1)`` (:276-278).  A level-1 class without rows gets no model
    (the reference's builder would assert, :243-245; search handles the missing model, :269-273).
    Returns ``(root_layers, {path: layers}, data_prediction int64[n,2], bucket_paths)``."""
    c1, c2 = (int(c) for c in n_categories)
    root_layers, pred0 = build_standin_index(x_nav, c1, model_type, train_rows, kmeans_iters, steps, lr, seed)
    n = x_nav.shape[0]
    pred = torch.full((n, 2), -1, dtype=torch.int64, device=x_nav.device)
    pred[:, 0] = pred0[:, 0]
    internal, bucket_paths = {}, []
    for j in range(c1):
        rows = torch.nonzero(pred[:, 0] == j).reshape(-1)
        if rows.numel() < 2 * c2:
            continue
        layers, p = build_standin_index(x_nav[rows], c2, model_type, train_rows, kmeans_iters, steps, lr,
                                        seed + 1 + j)
        internal[(j, -1)] = layers
        pred[rows, 1] = p[:, 0]
        bucket_paths += [(j, b) for b in range(int(torch.unique(p[:, 0]).numel()))]
    return root_layers, internal, pred, bucket_paths
