"""Generate ``tests/golden/*.npz`` by running the UNMODIFIED reference (TEST INFRASTRUCTURE).

    python -m oracle.make_golden            # build container only (/root/reference present)

Each fixture stores the exact inputs (fp16 vectors and queries, MLP weights or the
seed they are drawn from, ``data_prediction``, row ids) and what the reference's
``LearnedIndex.search`` / ``_precompute_bucket_order`` returned for them, so the
GPU box -- where /root/reference does not exist -- can check both the oracle
restatement and the CUDA path against the reference's own outputs.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import ref_harness  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

HIDDEN = {"MLP": [128], "MLP-2": [64], "MLP-3": [256], "MLP-4": [512], "MLP-5": [256, 128],
          "MLP-6": [32], "MLP-7": [16], "MLP-8": [8], "MLP-9": [8, 16]}


def seeded_layers(seed: int, model_type: str, in_dim: int, out_dim: int, scale: float = 1.0):
    """Deterministic, machine-independent weights: legacy RandomState normals scaled
    by 1/sqrt(fan_in) (plain elementwise f32 arithmetic)."""
    rs = np.random.RandomState(seed)
    dims = [in_dim] + HIDDEN[model_type] + [out_dim]
    layers = []
    for i in range(len(dims) - 1):
        W = rs.standard_normal((dims[i + 1], dims[i])).astype(np.float32)
        W = W * np.float32(scale / np.sqrt(dims[i]))
        b = rs.standard_normal(dims[i + 1]).astype(np.float32) * np.float32(0.1)
        layers.append((W, b))
    return layers


def unit_fp16(rs: np.random.RandomState, n: int, d: int, n_blobs: int = 0) -> np.ndarray:
    x = rs.standard_normal((n, d)).astype(np.float32)
    if n_blobs:
        c = rs.standard_normal((n_blobs, d)).astype(np.float32) * 2.0
        x = x + c[rs.randint(0, n_blobs, n)]
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    return x.astype(np.float16)


def one_level_case(name, seed, n, nq, d, C, model_type, runs, squash=None, id_base=1, store_weights=True,
                   weight_scale=4.0):
    """``runs``: list of (n_buckets, k).  ``squash``: {bucket: keep_rows} to force
    undersized / empty buckets by re-labelling surplus rows to bucket 0."""
    import torch

    ref = ref_harness.import_reference()
    rs = np.random.RandomState(seed)
    X = unit_fp16(rs, n, d, n_blobs=3 * C)
    Q = unit_fp16(rs, nq, d, n_blobs=3 * C)
    layers = seeded_layers(seed + 1, model_type, d, C, scale=weight_scale)
    li = ref_harness.make_reference_index(layers, model_type)
    pred = li.root_model.predict(ref["model"].data_X_to_torch(X)).astype(np.int64).reshape(-1, 1)
    if squash:
        for b, keep in squash.items():
            rows = np.flatnonzero(pred[:, 0] == b)
            pred[rows[keep:], 0] = 0 if b != 0 else 1
    ids = np.arange(id_base, id_base + n)
    out = dict(X=X, Q=Q, data_prediction=pred, ids=ids, n_categories=np.array([C]),
               model_type=np.array(model_type), weight_seed=np.array(seed + 1),
               weight_scale=np.array(weight_scale), runs=np.array(runs, dtype=np.int64))
    if store_weights:
        for i, (W, b) in enumerate(layers):
            out[f"W{i}"] = W
            out[f"b{i}"] = b
    probs, classes = li.root_model.predict_proba(ref["model"].data_X_to_torch(Q))
    out["probs"] = probs
    out["classes"] = classes
    for (B, k) in runs:
        order, _ = li._precompute_bucket_order(queries_navigation=Q, n_buckets=B, n_categories=[C])
        dists, nns, _ = ref_harness.run_reference_search(li, X, Q, X, Q, pred, [C], B, k=k, ids=ids)
        out[f"order_B{B}_k{k}"] = order
        out[f"dists_B{B}_k{k}"] = dists
        out[f"nns_B{B}_k{k}"] = nns
    path = os.path.join(GOLDEN_DIR, f"{name}.npz")
    np.savez_compressed(path, **out)
    sizes = np.bincount(pred[:, 0], minlength=C)
    print(f"{name}: n={n} nq={nq} d={d} C={C} sizes[min,max]=[{sizes.min()},{sizes.max()}] "
          f"-> {os.path.getsize(path) / 1e6:.2f} MB")


def main():
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    # small dims, MLP-4 hidden width, several budgets incl. full probe (B == C)
    one_level_case("l1_d128_c12", seed=11, n=2500, nq=160, d=128, C=12, model_type="MLP-4",
                   runs=[(1, 10), (3, 10), (12, 10), (4, 5)])
    # the 768-d shape of the BASELINE configs, narrow model so the fixture stays small
    one_level_case("l1_d768_c8", seed=23, n=700, nq=48, d=768, C=8, model_type="MLP-7",
                   runs=[(1, 10), (2, 10), (4, 10)])
    # 768 -> 512 -> C ("MLP-4", the headline network); weights regenerated from the seed
    one_level_case("l1_d768_mlp4", seed=31, n=600, nq=40, d=768, C=20, model_type="MLP-4",
                   runs=[(2, 10), (5, 10)], store_weights=False)
    # ragged: bucket 3 undersized (4 rows < 10), bucket 5 empty, ids not starting at 1
    one_level_case("l1_ragged", seed=47, n=1500, nq=200, d=64, C=10, model_type="MLP-6",
                   runs=[(1, 10), (3, 10), (10, 10), (3, 7)], squash={3: 4, 5: 0}, id_base=1001)


if __name__ == "__main__":
    main()
