"""Run the UNMODIFIED reference ``li`` package in-process (TEST INFRASTRUCTURE).

Needs the reference package: ``/root/reference/search/li`` (the build container) or its
byte-identical staged copy ``oracle/_ref/li`` (``oracle/build_ref.py``; git-ignored, ships to
the GPU box with the gpurun snapshot).  The ``-m gpu`` tests and ``smoke()`` never call this
(they use ``oracle.lmi_oracle`` and the committed ``tests/golden``); ``bench.py --impl
reference`` and its ``cpu_baseline`` leg do, to time the reference's own search on the host.

Three shims are needed to execute the reference under this image:

1. ``faiss`` is not installed -> put ``oracle/faiss_shim`` on ``sys.path``.
2. numpy >= 2 returns a tuple from ``np.ogrid[...]``; the reference assigns
   into it (search/li/LearnedIndex.py:137-138, pinned numpy 1.26.2) -> wrap
   ``np.ogrid`` so indexing returns a list.
3. ``search()`` adds ``category_L*`` columns to ``data_navigation`` in place
   (LearnedIndex.py:101-104); callers here must pass a distinct ``data_search``
   frame (``run_reference_search`` does).

The reference also moves its MLP to ``cuda:0`` when CUDA is visible
(search/li/model.py:124-125); ``force_cpu`` pins it to the CPU.
"""
import os
import sys

import numpy as np

REFERENCE_ROOT = os.environ.get("LMI_REFERENCE_ROOT", "/root/reference")
_HERE = os.path.dirname(os.path.abspath(__file__))
_SEARCH_DIRS = (os.path.join(REFERENCE_ROOT, "search"), os.path.join(_HERE, "_ref"))
_SHIM_DIR = os.path.join(_HERE, "faiss_shim")


def reference_dir():
    """Directory holding the reference's ``li`` package: the checkout, else the staged copy."""
    for d in _SEARCH_DIRS:
        if os.path.isfile(os.path.join(d, "li", "LearnedIndex.py")):
            return d
    return None


def reference_available() -> bool:
    return reference_dir() is not None


class _ListOgrid:
    """np.ogrid whose ``[...]`` result is a mutable list (numpy 1.x behaviour)."""

    def __init__(self, inner):
        self._inner = inner

    def __getitem__(self, key):
        out = self._inner[key]
        return list(out) if isinstance(out, tuple) else out


_imported = {}


def import_reference():
    """Returns the reference modules as a dict; idempotent."""
    if _imported:
        return _imported
    search_dir = reference_dir()
    if search_dir is None:
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT} nor staged under oracle/_ref "
                           "(python -m oracle.build_ref)")
    if _SHIM_DIR not in sys.path:
        sys.path.insert(0, _SHIM_DIR)
    if search_dir not in sys.path:
        sys.path.insert(0, search_dir)
    if not isinstance(np.ogrid, _ListOgrid):
        np.ogrid = _ListOgrid(np.ogrid)
    import li.LearnedIndex as m_index  # noqa: E402
    import li.model as m_model  # noqa: E402
    import li.PriorityQueue as m_pq  # noqa: E402
    import li.utils as m_utils  # noqa: E402

    # the per-bucket tqdm bar (LearnedIndex.py:350) is noise in test logs
    m_index.tqdm = lambda it, *a, **k: it
    _imported.update(index=m_index, model=m_model, pq=m_pq, utils=m_utils, source=search_dir)
    return _imported


def force_cpu(nn):
    import torch

    nn.device = torch.device("cpu")
    nn.model = nn.model.to(nn.device)
    return nn


def make_reference_network(layers, model_type):
    """Reference ``NeuralNetwork`` carrying the given weights.

    ``layers``: list of (W f32[out,in], b f32[out]) in forward order.
    """
    import torch

    ref = import_reference()
    in_dim = layers[0][0].shape[1]
    out_dim = layers[-1][0].shape[0]
    nn = ref["model"].NeuralNetwork(in_dim, out_dim, model_type=model_type)
    force_cpu(nn)
    linears = [m for m in nn.model.layers if isinstance(m, torch.nn.Linear)]
    assert len(linears) == len(layers), (len(linears), len(layers))
    with torch.no_grad():
        for lin, (W, b) in zip(linears, layers):
            assert tuple(lin.weight.shape) == tuple(W.shape)
            lin.weight.copy_(torch.from_numpy(np.asarray(W, dtype=np.float32)))
            lin.bias.copy_(torch.from_numpy(np.asarray(b, dtype=np.float32)))
    return nn


def make_reference_index(root_layers, root_type, internal=None, bucket_paths=None):
    """Reference ``LearnedIndex`` (LearnedIndex.py:23-39) from raw weights.

    ``internal``: {path tuple: (layers, model_type)} for multi-level indexes.
    """
    ref = import_reference()
    root = make_reference_network(root_layers, root_type)
    internal_models = {}
    for path, (layers, mtype) in (internal or {}).items():
        internal_models[tuple(path)] = make_reference_network(layers, mtype)
    if bucket_paths is None:
        bucket_paths = [(i,) for i in range(root_layers[-1][0].shape[0])]
    return ref["index"].LearnedIndex(root, internal_models, list(bucket_paths))


def run_reference_search(
    li, data_nav, queries_nav, data_search, queries_search, data_prediction,
    n_categories, n_buckets, k=10, ids=None, copy=None,
):
    """Calls the reference ``LearnedIndex.search`` (LearnedIndex.py:41-161).

    ``data_nav`` / ``data_search`` are ndarrays; they are wrapped in two distinct
    DataFrames whose index is ``ids`` (default 1..n, as search/search.py:190-191).
    ``copy=False`` lets both frames view the caller's array (needed at 10M rows, where a
    second and third f32 copy would not fit in RAM); the reference only ADDS and DROPS label
    columns on ``data_navigation`` (:101-104, :153-157), it never writes into the vectors.
    """
    import pandas as pd

    n = data_nav.shape[0]
    if ids is None:
        ids = np.arange(1, n + 1)
    df_nav = pd.DataFrame(np.asarray(data_nav), index=pd.Index(ids), copy=copy)
    df_search = pd.DataFrame(np.asarray(data_search), index=pd.Index(ids), copy=copy)
    dists, nns, times = li.search(
        data_navigation=df_nav,
        queries_navigation=queries_nav,
        data_search=df_search,
        queries_search=queries_search,
        data_prediction=data_prediction,
        n_categories=list(n_categories),
        n_buckets=n_buckets,
        k=k,
    )
    return dists, nns, dict(times)
