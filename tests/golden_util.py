"""Load ``tests/golden/*.npz`` (written by ``oracle/make_golden.py`` from the reference)."""
import glob
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

HIDDEN = {"MLP": [128], "MLP-2": [64], "MLP-3": [256], "MLP-4": [512], "MLP-5": [256, 128],
          "MLP-6": [32], "MLP-7": [16], "MLP-8": [8], "MLP-9": [8, 16]}


def seeded_layers(seed, model_type, in_dim, out_dim, scale=1.0):
    # must stay identical to oracle/make_golden.py:seeded_layers
    rs = np.random.RandomState(seed)
    dims = [in_dim] + HIDDEN[model_type] + [out_dim]
    layers = []
    for i in range(len(dims) - 1):
        W = rs.standard_normal((dims[i + 1], dims[i])).astype(np.float32)
        W = W * np.float32(scale / np.sqrt(dims[i]))
        b = rs.standard_normal(dims[i + 1]).astype(np.float32) * np.float32(0.1)
        layers.append((W, b))
    return layers


def case_names():
    """1-level fixtures."""
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "l1_*.npz")))


def multi_level_case_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "ml_*.npz")))


class MultiLevelCase:
    """``ml_*.npz`` (oracle/make_golden.py:multi_level_case): weights are regenerated from the seeds."""

    def __init__(self, name):
        z = np.load(os.path.join(GOLDEN_DIR, f"{name}.npz"))
        self.name, self.z = name, z
        self.X, self.Q = z["X"], z["Q"]
        self.data_prediction, self.ids = z["data_prediction"], z["ids"]
        self.n_categories = [int(v) for v in z["n_categories"]]
        L = len(self.n_categories)
        d = self.X.shape[1]
        scale = float(z["weight_scale"])
        self.root_layers = seeded_layers(int(z["root_seed"]), str(z["root_type"]), d, self.n_categories[0], scale)
        self.internal = {}
        for row in z["internal_meta"]:
            path = tuple(int(v) for v in row[:L])
            self.internal[path] = seeded_layers(int(row[L]), str(z["inner_type"]), d, int(row[L + 1]), scale)
        self.bucket_paths = [tuple(int(v) for v in r) for r in z["bucket_paths"]]
        self.runs = [(int(b), int(k)) for b, k in z["runs"]]

    def expected(self, B, k):
        return self.z[f"order_B{B}_k{k}"], self.z[f"dists_B{B}_k{k}"], self.z[f"nns_B{B}_k{k}"]


class GoldenCase:
    def __init__(self, name):
        z = np.load(os.path.join(GOLDEN_DIR, f"{name}.npz"))
        self.name = name
        self.z = z
        self.X = z["X"]
        self.Q = z["Q"]
        self.data_prediction = z["data_prediction"]
        self.ids = z["ids"]
        self.n_categories = [int(v) for v in z["n_categories"]]
        self.model_type = str(z["model_type"])
        self.runs = [(int(b), int(k)) for b, k in z["runs"]]
        if "W0" in z.files:
            self.layers = []
            i = 0
            while f"W{i}" in z.files:
                self.layers.append((z[f"W{i}"], z[f"b{i}"]))
                i += 1
        else:
            self.layers = seeded_layers(int(z["weight_seed"]), self.model_type, self.X.shape[1],
                                        self.n_categories[-1], float(z["weight_scale"]))
        self.probs = z["probs"]
        self.classes = z["classes"]

    def expected(self, B, k):
        return self.z[f"order_B{B}_k{k}"], self.z[f"dists_B{B}_k{k}"], self.z[f"nns_B{B}_k{k}"]


NO_CANDIDATE = 1e38  # dist >= this marks a filler slot


def assert_search_parity(dists, nns, exp_d, exp_i, rtol=1e-3, atol=2e-6, what=""):
    """ids identical except at distance ties; dists within ``rtol`` relative (north_star)."""
    assert dists.shape == exp_d.shape, (what, dists.shape, exp_d.shape)
    assert nns.shape == exp_i.shape, (what, nns.shape, exp_i.shape)
    real = exp_d < NO_CANDIDATE
    assert np.array_equal(real, dists < NO_CANDIDATE), f"{what}: filler pattern differs"
    np.testing.assert_allclose(dists[real], exp_d[real], rtol=rtol, atol=atol, err_msg=what)
    # filler distances: +inf (empty bucket) or FLT_MAX-ish (undersized bucket) must match in kind
    assert np.array_equal(np.isinf(dists), np.isinf(exp_d)), f"{what}: inf pattern differs"
    bad = (nns != exp_i)
    if bad.any():
        # a mismatch is only allowed where the oracle's own distances tie (neighbouring slot
        # within tolerance), i.e. the two ids are interchangeable
        rows, cols = np.nonzero(bad)
        for r, c in zip(rows, cols):
            d = exp_d[r, c]
            tol = atol + rtol * 1e-3 * abs(d) + 4e-7
            near = np.abs(exp_d[r] - d) <= tol
            assert near.sum() >= 2 or not real[r, c] or c == exp_d.shape[1] - 1, (
                f"{what}: id mismatch at query {r} slot {c}: got {nns[r, c]} want {exp_i[r, c]} "
                f"(dist {dists[r, c]} vs {d}) with no tie")
            if near.sum() >= 2:
                assert set(nns[r, near].tolist()) == set(exp_i[r, near].tolist()) or c == exp_d.shape[1] - 1 \
                    or near[-1], f"{what}: tied group differs at query {r}"
