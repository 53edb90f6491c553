"""Live differential test: oracle restatement vs the UNMODIFIED reference executed
in-process.  Runs only where /root/reference exists (the build container)."""
import numpy as np
import pytest

from oracle import lmi_oracle as orc
from oracle import ref_harness
from oracle.make_golden import seeded_layers, unit_fp16
from tests.golden_util import assert_search_parity

pytestmark = pytest.mark.skipif(not ref_harness.reference_available(), reason="/root/reference not present")


@pytest.mark.parametrize("seed,n,nq,d,C,mtype,B,k", [
    (1, 900, 70, 32, 6, "MLP-6", 1, 10),
    (2, 900, 70, 32, 6, "MLP-6", 2, 10),
    (3, 1200, 50, 64, 9, "MLP-2", 4, 10),
    (4, 1200, 50, 64, 9, "MLP-9", 9, 10),
    (5, 800, 33, 96, 7, "MLP", 3, 4),
    (6, 800, 33, 96, 7, "MLP-5", 3, 10),
])
def test_one_level(seed, n, nq, d, C, mtype, B, k):
    ref = ref_harness.import_reference()
    rs = np.random.RandomState(seed)
    X = unit_fp16(rs, n, d, n_blobs=2 * C)
    Q = unit_fp16(rs, nq, d, n_blobs=2 * C)
    layers = seeded_layers(seed + 100, mtype, d, C, scale=4.0)
    li = ref_harness.make_reference_index(layers, mtype)
    pred = li.root_model.predict(ref["model"].data_X_to_torch(X)).astype(np.int64).reshape(-1, 1)
    exp_order, _ = li._precompute_bucket_order(queries_navigation=Q, n_buckets=B, n_categories=[C])
    exp_d, exp_i, exp_t = ref_harness.run_reference_search(li, X, Q, X, Q, pred, [C], B, k=k)

    assert np.array_equal(orc.predict(layers, orc.to_f32_tensor(X)), pred[:, 0])
    d_, i_, t_, order = orc.search(orc.OracleIndex(layers), Q, X, Q, pred, [C], n_buckets=B, k=k)
    assert np.array_equal(order, exp_order)
    assert d_.dtype == exp_d.dtype and i_.dtype == exp_i.dtype
    assert_search_parity(d_, i_, exp_d, exp_i, rtol=1e-6, atol=1e-7, what=f"seed {seed}")
    assert set(t_) == set(exp_t)


@pytest.mark.parametrize("seed,cats,B", [(7, [4, 3], 1), (8, [4, 3], 3), (9, [3, 3], 5), (10, [2, 2, 2], 4)])
def test_multi_level_bucket_order_and_search(seed, cats, B):
    """n_levels > 1 (LearnedIndex.py:216-325): internal models per path, PriorityQueue."""
    import itertools

    ref = ref_harness.import_reference()
    rs = np.random.RandomState(seed)
    n, nq, d = 1500, 60, 48
    X = unit_fp16(rs, n, d, n_blobs=10)
    Q = unit_fp16(rs, nq, d, n_blobs=10)
    L = len(cats)
    root_layers = seeded_layers(seed + 200, "MLP-6", d, cats[0], scale=4.0)
    internal = {}
    for depth in range(1, L):
        for j, prefix in enumerate(itertools.product(*[range(c) for c in cats[:depth]])):
            path = tuple(prefix) + (orc.EMPTY_VALUE,) * (L - depth)
            internal[path] = seeded_layers(seed + 300 + 17 * depth + j, "MLP-7", d, cats[depth], scale=4.0)
    # assign every object by descending the tree with argmax (LearnedIndexBuilder.py:203-280 contract)
    pred = np.full((n, L), orc.EMPTY_VALUE, dtype=np.int64)
    pred[:, 0] = orc.predict(root_layers, orc.to_f32_tensor(X))
    for depth in range(1, L):
        for path, layers in internal.items():
            if len(path) - path.count(orc.EMPTY_VALUE) != depth:
                continue
            rows = np.flatnonzero((pred[:, :depth] == np.array(path[:depth])).all(axis=1))
            if rows.size:
                pred[rows, depth] = orc.predict(layers, orc.to_f32_tensor(X[rows]))
    bucket_paths = sorted({tuple(int(v) for v in r) for r in pred})
    if len(bucket_paths) < B:
        pytest.skip("fewer buckets than the budget")

    li = ref_harness.make_reference_index(
        root_layers, "MLP-6", {p: (l, "MLP-7") for p, l in internal.items()}, bucket_paths)
    exp_order, _ = li._precompute_bucket_order(queries_navigation=Q, n_buckets=B, n_categories=cats)
    exp_d, exp_i, _ = ref_harness.run_reference_search(li, X, Q, X, Q, pred, cats, B, k=10)

    idx = orc.OracleIndex(root_layers, internal, bucket_paths)
    d_, i_, _, order = orc.search(idx, Q, X, Q, pred, cats, n_buckets=B, k=10)
    assert np.array_equal(order, exp_order)
    assert_search_parity(d_, i_, exp_d, exp_i, rtol=1e-6, atol=1e-7, what=f"seed {seed}")
