"""The CPU oracle restatement reproduces what the reference itself returned
(fixtures written by oracle/make_golden.py from the unmodified reference)."""
import numpy as np
import pytest

from oracle import lmi_oracle as orc
from tests.golden_util import GoldenCase, case_names, assert_search_parity


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
