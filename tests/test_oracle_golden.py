"""The CPU oracle restatement reproduces what the reference itself returned
(fixtures written by oracle/make_golden.py from the unmodified reference)."""
import numpy as np
import pytest

from oracle import lmi_oracle as orc
from tests.golden_util import (GoldenCase, MultiLevelCase, assert_search_parity, case_names,
                               multi_level_case_names)


@pytest.mark.parametrize("name", case_names())
def test_predict_proba_matches_reference(name):
    c = GoldenCase(name)
    probs, classes = orc.predict_proba(c.layers, orc.to_f32_tensor(c.Q))
    np.testing.assert_allclose(probs, c.probs, rtol=1e-5, atol=1e-7)
    # class order may only differ where the reference's probabilities tie
    diff = classes != c.classes
    if diff.any():
        assert np.all(np.abs(probs[diff] - c.probs[diff]) <= 1e-6)


@pytest.mark.parametrize("name", case_names())
def test_search_matches_reference(name):
    c = GoldenCase(name)
    index = orc.OracleIndex(c.layers)
    for B, k in c.runs:
        exp_order, exp_d, exp_i = c.expected(B, k)
        d, i, times, order = orc.search(index, c.Q, c.X, c.Q, c.data_prediction, c.n_categories,
                                        n_buckets=B, k=k, ids=c.ids)
        assert order.dtype == np.int32 and order.shape == exp_order.shape
        assert np.array_equal(order, exp_order), f"{name} B={B}: bucket order differs"
        assert d.dtype == np.float64 and i.dtype == np.uint32
        assert_search_parity(d, i, exp_d, exp_i, rtol=1e-6, atol=1e-7, what=f"{name} B={B} k={k}")
        assert set(times) >= {"inference", "search_within_buckets", "seq_search", "sort", "search"}


def test_golden_covers_edge_cases():
    c = GoldenCase("l1_ragged")
    sizes = np.bincount(c.data_prediction[:, 0], minlength=c.n_categories[0])
    assert sizes[3] == 4 and sizes[5] == 0
    _, d, i = c.expected(1, 10)
    assert np.isinf(d).any(), "an empty bucket must have produced +inf fillers"
    assert ((d > 1e38) & ~np.isinf(d)).any(), "an undersized bucket must have produced FLT_MAX fillers"
    assert c.ids[0] == 1001


@pytest.mark.parametrize("name", multi_level_case_names())
def test_multi_level_search_matches_reference(name):
    """n_levels > 1: best-first navigation (LearnedIndex.py:216-325) incl. a node without a
    model, a populated path missing from bucket_paths and a model with fewer classes."""
    c = MultiLevelCase(name)
    index = orc.OracleIndex(c.root_layers, c.internal, c.bucket_paths)
    for B, k in c.runs:
        exp_order, exp_d, exp_i = c.expected(B, k)
        d, i, _, order = orc.search(index, c.Q, c.X, c.Q, c.data_prediction, c.n_categories,
                                    n_buckets=B, k=k, ids=c.ids)
        assert order.dtype == np.int32 and order.shape == exp_order.shape == (c.Q.shape[0], B, len(c.n_categories))
        assert np.array_equal(order, exp_order), f"{name} B={B}: bucket order differs"
        assert_search_parity(d, i, exp_d, exp_i, rtol=1e-6, atol=1e-7, what=f"{name} B={B} k={k}")


def test_multi_level_golden_covers_quirks():
    c = MultiLevelCase("ml_3x3x2_quirks")
    assert (2, 1, -1) not in c.internal and (0, 0, 1) not in c.bucket_paths
    assert c.internal[(1, -1, -1)][-1][0].shape[0] == 2
    assert (c.data_prediction == -1).any(), "rows under the model-less node keep an incomplete path"
    assert ((c.data_prediction == np.array([0, 0, 1])).all(axis=1)).any(), "the unlisted path is populated"
