"""GPU build of the index behind the reference's builder API:
``BuildConfiguration`` (search/li/BuildConfiguration.py:8-135) and
``LearnedIndexBuilder(data, config).build()`` (search/li/LearnedIndexBuilder.py:21-280)
returning ``(index, data_prediction int64[n, L], n_buckets, build_t, cluster_t)``.

What the reference does per model (``_train_model``, :119-201) and how it runs here:

* cluster the node's rows with ``faiss.Kmeans(d, k, seed=2023)`` (clustering/faiss_kmeans.py:8-24;
  faiss defaults: 20 Lloyd iterations on at most 256 points per centroid, L2) -> ``kmeans_gpu``:
  the assignment step ``argmin |x - c|^2 = argmax (x.c - |c|^2/2)`` is this package's own
  tensor-core Linear + argmax kernels (``lmi_linear_tc`` / ``lmi_argmax``), the centroid update an
  ``index_add_``;
* train ``Model`` (model.py:15-99) with Adam on cross-entropy.  ``train_batch`` (:191-211)
  iterates over all mini-batches of 256 but calls ``backward()``/``step()`` ONCE PER EPOCH, on
  the last mini-batch only -- so an "epoch" is one Adam step on one random batch.  That is what
  runs here (one step per epoch on a random batch of the same size), without the forward passes
  whose results the reference throws away; on a GPU the 10M build drops from hours to seconds;
* re-train until the model predicts every category (:183-192), ``data_prediction`` = the
  model's own argmax (:76, :268), internal models per level-1 class on the rows routed there
  (:203-280), ``bucket_paths`` numbered ``range(len(unique(predictions)))`` (:276-278).

Training uses torch autograd (device plumbing); every ``predict`` goes through the CUDA kernels.
Seeds follow the reference (2023, search/li/LearnedIndex.py:18-19) but neither faiss' nor torch's
CPU random streams are reproduced: the build is functionally, not bitwise, equivalent.
"""
from __future__ import annotations

import time
from itertools import product
from typing import Any, Dict, List, Optional, Tuple

import numpy as np
import torch

from .LearnedIndex import EMPTY_VALUE, LearnedIndex
from .model import HIDDEN_LAYERS, NeuralNetwork, get_device

SEED = 2023


class BuildConfiguration:
    """search/li/BuildConfiguration.py:8-135 (same validation and expansion rules)."""

    def __init__(self, clustering_algorithms, epochs, model_types, lrs, n_categories: List[int]):
        assert len(n_categories) > 0, "n_categories must specify at least one level"
        args = [clustering_algorithms, epochs, model_types, lrs]
        are_lists = all(isinstance(a, list) for a in args)
        are_scalars = (callable(clustering_algorithms) and isinstance(epochs, int)
                       and isinstance(model_types, str) and isinstance(lrs, float))
        assert are_lists or are_scalars, \
            "clustering_algorithms, epochs, model_types, and lrs must be lists or single values"
        for a in args:
            if isinstance(a, list):
                assert len(a) == 1 or len(a) == len(n_categories), (
                    "clustering_algorithms, epochs, model_types, and lrs must "
                    "be lists of size 1 or the same size as n_categories")
        expand = lambda a: (([a[0]] * len(n_categories)) if len(a) == 1 else a) if isinstance(a, list) \
            else [a] * len(n_categories)
        self.clustering_algorithms = expand(clustering_algorithms)
        self.epochs = expand(epochs)
        self.model_types = expand(model_types)
        self.lrs = expand(lrs)
        self.n_categories = n_categories
        self.level_configurations = [
            (self.clustering_algorithms[i], self.model_types[i], self.epochs[i], self.lrs[i], self.n_categories[i])
            for i in range(len(n_categories))]
        self.n_levels = len(n_categories)


def kmeans_gpu(x: torch.Tensor, n_clusters: int, parameters: Optional[Dict[str, Any]] = None
               ) -> Tuple[torch.Tensor, torch.Tensor]:
    """``clustering.algorithms["faiss_kmeans"]`` (clustering/faiss_kmeans.py:8-24) on the GPU.
    ``x``: device matrix fp16/f32 [n, d].  Returns (centroids f32[k, d], labels int64[n])."""
    p = {"seed": SEED, "niter": 20, "max_points_per_centroid": 256}
    p.update({k: v for k, v in (parameters or {}).items() if k in p})
    n, d = x.shape
    g = torch.Generator(device=x.device).manual_seed(int(p["seed"]))
    cap = n_clusters * int(p["max_points_per_centroid"])
    train = x if n <= cap else x[torch.randperm(n, generator=g, device=x.device)[:cap]]
    cent = train[torch.randperm(train.shape[0], generator=g, device=x.device)[:n_clusters]].float().clone()

    def assign(points: torch.Tensor, c: torch.Tensor) -> torch.Tensor:
        net = NeuralNetwork([(c, -0.5 * (c * c).sum(dim=1))], device=points.device)
        return net.predict_device(points.contiguous())

    for _ in range(int(p["niter"])):
        lab = assign(train, cent)
        sums = torch.zeros_like(cent).index_add_(0, lab, train.float())
        cnt = torch.bincount(lab, minlength=n_clusters).unsqueeze(1)
        new = sums / cnt.clamp(min=1)
        empty = (cnt == 0).reshape(-1)
        if empty.any():  # re-seed (faiss splits a large cluster; any point keeps the count at k)
            pick = torch.randint(0, train.shape[0], (int(empty.sum()),), generator=g, device=x.device)
            new[empty] = train[pick].float()
        cent = new
    return cent, assign(x, cent)


def _torch_model(in_dim: int, out_dim: int, model_type: str, device) -> torch.nn.Sequential:
    if model_type not in HIDDEN_LAYERS:
        raise ValueError(f"Model type {model_type} not supported.")  # model.py:90-92
    dims = [in_dim] + HIDDEN_LAYERS[model_type] + [out_dim]
    mods: List[torch.nn.Module] = []
    for i in range(len(dims) - 1):
        mods.append(torch.nn.Linear(dims[i], dims[i + 1]))
        if i + 2 < len(dims):
            mods.append(torch.nn.ReLU())
    return torch.nn.Sequential(*mods).to(device)


kmeans_gpu.accepts_device_tensor = True  # every other callable gets the reference's numpy contract


class LearnedIndexBuilder:
    """search/li/LearnedIndexBuilder.py:21-280.  ``data``: DataFrame / ndarray / device tensor
    [n, d]; row i carries the label i+1, as the reference assumes (:247, :270)."""

    BATCH = 256  # LearnedIndexBuilder.py:169

    def __init__(self, data, config: BuildConfiguration, device: Optional[torch.device] = None):
        self.config = config
        self.device = device or get_device()
        if isinstance(data, torch.Tensor):
            x = data
        else:
            a = data.to_numpy(copy=False) if hasattr(data, "to_numpy") else np.asarray(data)
            if a.dtype not in (np.float16, np.float32):
                a = a.astype(np.float32)
            x = torch.from_numpy(np.ascontiguousarray(a))
        self.data = x.to(self.device).contiguous()
        self.root_model: Optional[NeuralNetwork] = None
        self.internal_models: Dict[Tuple, NeuralNetwork] = {}
        self.bucket_paths: List[Tuple] = []
        self._gen = torch.Generator(device=self.device).manual_seed(SEED)

    # ---------------------------------------------------------------- one model (:119-201)
    def _train_model(self, x: torch.Tensor, params) -> Tuple[NeuralNetwork, float]:
        clustering_algorithm, model_type, epochs, lr, n_categories = params
        n = x.shape[0]
        t0 = time.time()
        if n < 2:  # _cluster :288-289
            labels = torch.zeros(n, dtype=torch.int64, device=self.device)
        else:
            k = n_categories
            if n < k:  # :291-294
                k = max(2, n // 5)
            algo = clustering_algorithm if callable(clustering_algorithm) else kmeans_gpu
            if getattr(algo, "accepts_device_tensor", False):
                _, labels = algo(x, k, None)
            else:
                # the reference's contract (clustering/faiss_kmeans.py:8-24, scikit_kmeans.py:8-29):
                # (np.ndarray f32 [n, d], n_clusters, parameters) -> (object, labels)
                _, labels = algo(x.float().cpu().numpy(), k, None)
            labels = torch.as_tensor(np.asarray(labels).reshape(-1) if not isinstance(labels, torch.Tensor) else labels,
                                     device=self.device).to(torch.int64)
        torch.cuda.synchronize(self.device)
        cluster_t = time.time() - t0
        # labels are used as class ids as they are (:166-168); the class count is the number of
        # distinct labels (:160-165)
        uniq = torch.unique(labels)
        n_categories = int(uniq.numel())
        assert int(labels.max().item()) < n_categories or n == 0, "clustering returned non-contiguous labels"

        net = _torch_model(x.shape[1], n_categories, model_type, self.device)
        opt = torch.optim.Adam(net.parameters(), lr=lr)
        loss_fn = torch.nn.CrossEntropyLoss()
        last = n - self.BATCH * ((n - 1) // self.BATCH)  # size of the sampler's final mini-batch
        iters = 0
        while True:
            for _ in range(epochs):  # train_batch :191-211: one optimiser step per epoch, on the last batch
                sel = torch.randperm(n, generator=self._gen, device=self.device)[:last]
                loss = loss_fn(net(x[sel].float()), labels[sel])
                net.zero_grad()
                loss.backward()
                opt.step()
            model = NeuralNetwork.from_reference(net, device=self.device)
            model.model_type = model_type
            pred = model.predict_device(x)
            iters += 1
            if iters > 1000:
                raise RuntimeError("The model did not converge after 1000 iterations.")  # :186-187
            if int(torch.unique(pred).numel()) == n_categories:  # :189
                return model, cluster_t

    # ---------------------------------------------------------------- build (:43-107)
    def build(self):
        s = time.time()
        cfg = self.config
        L, n = cfg.n_levels, self.data.shape[0]
        pred = torch.full((n, L), EMPTY_VALUE, dtype=torch.int64, device=self.device)
        self.root_model, cluster_t = self._train_model(self.data, cfg.level_configurations[0])
        pred[:, 0] = self.root_model.predict_device(self.data)
        if L == 1:
            self.bucket_paths = [(b,) for b in range(int(torch.unique(pred[:, 0]).numel()))]  # :78-79
        for level in range(1, L):  # _train_internal_models :203-280
            combos = [range(cfg.n_categories[l]) for l in range(level)]
            for prefix in product(*combos):
                path = tuple(prefix) + (EMPTY_VALUE,) * (L - level)
                want = torch.tensor(path, dtype=torch.int64, device=self.device)
                rows = torch.nonzero((pred == want).all(dim=1)).reshape(-1)
                assert rows.numel() != 0, "There are no data points associated with the given path."  # :243-245
                model, ct = self._train_model(self.data[rows].contiguous(), cfg.level_configurations[level])
                self.internal_models[path] = model
                cluster_t += ct
                p = model.predict_device(self.data[rows].contiguous())
                pred[rows, level] = p
                if level == L - 1:
                    for b in range(int(torch.unique(p).numel())):  # :276-278
                        self.bucket_paths.append(path[:-1] + (b,))
        index = LearnedIndex(self.root_model, self.internal_models, self.bucket_paths)
        torch.cuda.synchronize(self.device)
        return index, pred.cpu().numpy(), len(self.bucket_paths), time.time() - s, cluster_t
