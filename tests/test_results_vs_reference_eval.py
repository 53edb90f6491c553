"""files written by ``learnedmetricindex_b200.results.store_results`` are consumed by the
reference's OWN evaluation scripts, run unmodified from ``/root/reference`` in this container (skipped where
the checkout is absent, e.g. on the GPU box): ``eval/eval.py`` as ``__main__`` walks ``result/``, reads ``knns`` and
the attributes and writes ``res.csv``; ``eval/plot.py`` as ``__main__`` reads that CSV.  ``h5py`` and ``matplotlib``
are not in the image: a minimal archive-backed ``h5py`` (tests/h5py_shim) and a recording mock of ``matplotlib``
stand in -- the scripts' own logic (attribute / dataset names, recall, CSV columns, the QPS axis) is what runs."""
import csv
import os
import runpy
import sys
from unittest import mock

import numpy as np
import pytest

REF = "/root/reference"
SHIM = os.path.join(os.path.dirname(os.path.abspath(__file__)), "h5py_shim")

pytestmark = pytest.mark.skipif(not os.path.isfile(os.path.join(REF, "eval", "eval.py")),
                                reason="reference checkout not present")


@pytest.fixture
def h5py_shim(monkeypatch):
    monkeypatch.syspath_prepend(SHIM)
    monkeypatch.delitem(sys.modules, "h5py", raising=False)
    import h5py

    assert h5py.__file__.startswith(SHIM)
    yield h5py
    sys.modules.pop("h5py", None)


def test_reference_eval_and_plot_consume_our_result_files(tmp_path, monkeypatch, h5py_shim):
    from learnedmetricindex_b200 import results

    rs = np.random.RandomState(0)
    nq, k, n = 200, 10, 5000
    gt = np.stack([rs.permutation(n)[:100] + 1 for _ in range(nq)]).astype(np.int32)  # ground truth: 100 ids per query
    runs = []
    for j, (B, keep) in enumerate([(1, 5), (4, 8), (16, 10)]):
        # a search result that shares `keep` of its 10 ids with the true top-10 of every query
        knns = gt[:, :k].astype(np.uint32).copy()
        knns[:, keep:] = n + 100 + np.arange(k - keep, dtype=np.uint32)
        dists = np.sort(rs.rand(nq, k), axis=1)
        path = tmp_path / "result" / "clip768v2" / "100K" / f"LMI-run{j}.h5"
        wrote = results.store_results(str(path), "LMI", "clip768v2", dists, knns, 12.5, 0.25 * (j + 1),
                                      f"index=b200;query=nprobe={B}", "100K")
        assert wrote == str(path)  # h5py importable -> the reference's container, at the reference's path
        back = results.load_results(wrote)
        assert np.array_equal(back["knns"], knns) and back["params"] == f"index=b200;query=nprobe={B}"
        runs.append((wrote, keep / k, 0.25 * (j + 1), f"nprobe={B}"))
    # the ground-truth file eval.py would download (eval/eval.py:14-24): same container, dataset 'knns'
    os.makedirs(tmp_path / "data")
    with h5py_shim.File(str(tmp_path / "data" / "groundtruth-100K.h5"), "w") as f:
        f.create_dataset("knns", gt.shape, dtype=gt.dtype)[:] = gt

    # --- eval/eval.py, unmodified, as __main__ (walks ./result, writes ./res.csv) ---
    monkeypatch.chdir(tmp_path)
    runpy.run_path(os.path.join(REF, "eval", "eval.py"), run_name="__main__")
    rows = list(csv.DictReader(open(tmp_path / "res.csv", newline="")))
    assert len(rows) == 3 and list(rows[0].keys()) == results.RES_COLUMNS
    by_params = {r["params"]: r for r in rows}
    for wrote, recall, qt, ident in runs:
        r = by_params[f"index=b200;query={ident}"]
        assert r["data"] == "clip768v2" and r["size"] == "100K" and r["algo"] == "LMI"
        assert float(r["buildtime"]) == 12.5 and float(r["querytime"]) == qt
        assert float(r["recall"]) == pytest.approx(recall, abs=1e-12)

    # our CSV writer yields the same table (row order aside)
    ours = results.write_res_csv([w for w, *_ in runs], {"100K": gt}, dst=str(tmp_path / "ours.csv"))
    ours_rows = {r["params"]: r for r in csv.DictReader(open(ours, newline=""))}
    for key, r in by_params.items():
        assert {c: str(v) for c, v in ours_rows[key].items() if c != "recall"} == \
               {c: str(v) for c, v in r.items() if c != "recall"}
        assert float(ours_rows[key]["recall"]) == float(r["recall"])

    # --- eval/plot.py, unmodified, as __main__ on the CSV (matplotlib mocked: record what it draws) ---
    mpl, plt = mock.MagicMock(name="matplotlib"), mock.MagicMock(name="pyplot")
    mpl.pyplot = plt
    monkeypatch.setitem(sys.modules, "matplotlib", mpl)
    monkeypatch.setitem(sys.modules, "matplotlib.pyplot", plt)
    monkeypatch.setattr(sys, "argv", ["plot.py", "--size", "100K", "res.csv"])
    runpy.run_path(os.path.join(REF, "eval", "plot.py"), run_name="__main__")
    assert plt.savefig.call_args[0][0] == "result_100K.png"
    (xs, ys), kw = plt.plot.call_args[0][:2], plt.plot.call_args[1]
    # pareto frontier over (recall, 10000 / querytime): all three runs are on it, fastest first
    assert list(xs) == [0.5, 0.8, 1.0]
    assert list(ys) == pytest.approx([10000 / 0.25, 10000 / 0.5, 10000 / 0.75])
