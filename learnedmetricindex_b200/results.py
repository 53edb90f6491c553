"""Result / evaluation contract of the reference's drivers, so that
``eval/eval.py`` and ``eval/plot.py`` keep working on what this package produces.

* ``store_results``  -- search/search.py:51-63: one file per (algo, params) with attributes
  ``algo, data, buildtime, querytime, size, params`` and datasets ``knns`` (the ids ``search``
  returned) and ``dists``.  HDF5 through ``h5py`` when it is importable; otherwise the same
  keys in an ``.npz`` next to the requested path (h5py is absent from the build image).
* ``get_recall``     -- eval/eval.py:37-45.
* ``write_res_csv``  -- eval/eval.py:55-77: ``res.csv`` with the columns
  ``data,size,algo,buildtime,querytime,params,recall``.
"""
from __future__ import annotations

import csv
import os
from pathlib import Path
from typing import Dict, Iterable

import numpy as np

RES_COLUMNS = ["data", "size", "algo", "buildtime", "querytime", "params", "recall"]  # eval/eval.py:59
ATTRS = ["algo", "data", "buildtime", "querytime", "size", "params"]


def store_results(dst, algo, kind, dists, anns, buildtime, querytime, params, size) -> str:
    """search/search.py:51-63.  Returns the path written."""
    os.makedirs(Path(dst).parent, exist_ok=True)
    attrs = dict(algo=algo, data=kind, buildtime=buildtime, querytime=querytime, size=size, params=params)
    try:
        import h5py
    except ImportError:
        h5py = None
    if h5py is not None:
        with h5py.File(dst, "w") as f:
            for k, v in attrs.items():
                f.attrs[k] = v
            f.create_dataset("knns", anns.shape, dtype=anns.dtype)[:] = anns
            f.create_dataset("dists", dists.shape, dtype=dists.dtype)[:] = dists
        return str(dst)
    out = str(Path(dst).with_suffix(".npz"))
    np.savez(out, knns=anns, dists=dists, **{k: np.asarray(v) for k, v in attrs.items()})
    return out


def load_results(path) -> Dict[str, object]:
    """The attributes and datasets of a file written by ``store_results`` (either container)."""
    path = str(path)
    if path.endswith(".npz"):
        z = np.load(path, allow_pickle=False)
        out = {k: (z[k].item() if z[k].ndim == 0 else z[k]) for k in z.files}
        return out
    import h5py

    with h5py.File(path, "r") as f:
        out = {k: f.attrs[k] for k in ATTRS if k in f.attrs}
        out["knns"] = np.array(f["knns"])
        out["dists"] = np.array(f["dists"])
    return out


def get_recall(I: np.ndarray, gt: np.ndarray, k: int) -> float:
    """eval/eval.py:37-45: mean over queries of |top-k(I) ∩ top-k(gt)| / k (sets, so duplicate ids
    in a row count once)."""
    assert k <= I.shape[1]
    assert len(I) == len(gt)
    n = len(I)
    hits = 0
    for i in range(n):
        hits += len(set(I[i, :k].tolist()) & set(gt[i, :k].tolist()))
    return hits / (n * k)


def write_res_csv(result_files: Iterable[str], groundtruth: Dict[str, np.ndarray], dst: str = "res.csv",
                  k: int = 10) -> str:
    """eval/eval.py:55-77 with the ground truth handed in (``{size: knns}``) instead of downloaded."""
    with open(dst, "w", newline="") as fh:
        w = csv.DictWriter(fh, fieldnames=RES_COLUMNS)
        w.writeheader()
        for path in result_files:
            r = load_results(path)
            row = {c: r.get(c, 0) for c in RES_COLUMNS}
            row["recall"] = get_recall(np.asarray(r["knns"]), groundtruth[str(r["size"])], k)
            w.writerow(row)
    return dst
