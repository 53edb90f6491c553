"""CPU-side checks: the C-ABI library builds, loads and exports every symbol the header
declares; host-side planning logic."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "lmi_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lmi_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_loads_and_exports_header_symbols():
    from learnedmetricindex_b200 import _native

    _native.build_library()
    lib = _native.lib()
    declared = _declared_symbols()
    assert len(declared) >= 12
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/lmi_b200.h but not exported"
    assert set(declared) == set(_native.EXPORTED_SYMBOLS), "ctypes signatures out of sync with the header"
    assert lib.lmi_version() == 7
    assert lib.lmi_error_string(0) == b"success"
    assert b"bad argument" in lib.lmi_error_string(-1)
    # size helpers are pure host functions
    assert lib.lmi_max_items(10000, 4, 120, 5) >= 5 * (40000 // 128 + 120)
    assert lib.lmi_invert_workspace_ints(10000, 4, 120) % 120 == 0
    import ctypes
    offs = (ctypes.c_int64 * 13)()
    total = lib.lmi_workspace_layout(10000, 4, 120, 5, 768, offs)
    assert total == offs[12] == lib.lmi_workspace_bytes(10000, 4, 120, 5) + 2 * 10000 * 768
    assert offs[12] - offs[11] >= 4 * 10000 * 4 * 10 and offs[11] % 256 == 0  # LMI_WS_GATE: uint32[2*rmax, nq*B]
    assert all(offs[i] % 256 == 0 and offs[i] <= offs[i + 1] for i in range(11))
    assert offs[9] - offs[8] >= 4 * 40000 * 10 * 10  # part_sim: P * 2*rmax * 10 floats


def test_sass_is_blackwell_native():
    """tcgen05 / TMA must be what the scan kernel compiles to (no legacy HMMA path)."""
    import shutil
    import subprocess

    from learnedmetricindex_b200 import _native

    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.isfile(cuobjdump):
        pytest.skip("cuobjdump not available")
    _native.build_library()
    sass = subprocess.run([cuobjdump, "-sass", _native.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    for needle in ("UTCHMMA", "UTMALDG", "LDTM", "STTM"):
        assert needle in sass, needle
    # the cluster-of-2 scan (lmi_scan_topk impl 2): multicast TMA loads and the multicast tcgen05.commit
    for needle in ("UTMALDG.2D.MULTICAST", "UTCBAR.MULTICAST"):
        assert needle in sass, needle
    assert "HMMA." not in sass.replace("UTCHMMA", "")


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from learnedmetricindex_b200 import _native

    monkeypatch.setattr(_native, "_lib", None)
    monkeypatch.setattr(_native, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _native.lib()


def test_no_cuda_fails_loudly():
    import torch

    from learnedmetricindex_b200 import model

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        model.get_device()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "learnedmetricindex_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("oracle/", "").lower(), f"{f} mentions the oracle"


@pytest.mark.parametrize("nq,B,C,n,maxb", [
    (10000, 4, 120, 10_000_000, 120_000), (1, 1, 120, 10_000_000, 120_000), (1, 4, 120, 10_000_000, 90_000),
    (100_000, 120, 120, 10_000_000, 120_000), (10000, 4, 120, 100_000, 1100), (7, 3, 5, 40, 17),
])
def test_plan_scan(nq, B, C, n, maxb):
    from learnedmetricindex_b200 import plan_scan
    from learnedmetricindex_b200.LearnedIndex import MAX_RANGES

    for nonempty, tile in ((None, 64), (max(1, C // 8), 64), (None, 128)):
        p = plan_scan(nq, B, C, n, maxb, n_nonempty=nonempty, rows_per_tile=tile)  # 128: lmi_scan_topk impl 3 / 4
        length, tail, rmax = p["range_len"], p["tail_range_len"], p["rmax"]
        assert length % tile == 0 and length >= tile
        assert 1 <= rmax <= MAX_RANGES
        shortest = tail or length
        assert tail % 64 == 0 and tail < length and 0 <= p["tail_permille"] <= 400
        assert (tail == 0) == (p["tail_permille"] == 0)
        assert rmax * shortest >= maxb  # every bucket is covered whichever length cuts it
        assert (rmax - 1) * shortest < maxb or rmax == 1
        if tail:
            assert -(-maxb // tail) <= 16  # <= 32 partial lists per probe: the merge's one-list-per-lane path


def test_plan_scan_sharded_shapes_get_long_items_and_a_tail():
    """A rank of a bucket-sharded index sees all queries but 1/world of the buckets: the cut must not fall back to
    thousands of short items (measured on a B200, profiles/r03_shard_plan_lab_*.jsonl: 1/8 of the 10M index scans
    in 0.70 ms with ~15k-row items and a short tail level, 0.87-1.06 ms with 3.7k-row items).  Few queries keep the
    fine cut (HBM bound: every SM has to stream)."""
    from learnedmetricindex_b200 import plan_scan

    full = plan_scan(10000, 4, 120, 10_000_000, 113_656)
    assert full["range_len"] == 24576 and full["tail_range_len"] == 8192  # the single-GPU headline cut is unchanged
    for world in (2, 4, 8):
        p = plan_scan(10000, 4, 120, 10_000_000 // world, 113_656, n_nonempty=120 // world)
        assert p["range_len"] >= 12288 and 0 < p["tail_range_len"] < p["range_len"] and p["tail_permille"] > 0, (world, p)
        assert p["rmax"] <= 16  # <= 32 partial lists per probe: the merge's one-list-per-lane path
    few = plan_scan(10, 4, 120, 10_000_000, 113_656)
    assert few["range_len"] <= 4096
    mid = plan_scan(1000, 4, 120, 10_000_000, 113_656)
    assert few["range_len"] < mid["range_len"] <= 24576


def test_flat_bucket_ids_follow_groupby_key_order():
    """Multi-level layout: flat id == rank of the path in groupby(category_L1..L) key order
    (LearnedIndex.py:343-350); incomplete paths go to one trailing, unprobeable bucket."""
    import numpy as np
    from learnedmetricindex_b200.navigation import flat_bucket_ids, unflatten_bucket_ids

    cats = [3, 4, 2]
    paths = np.array([[a, b, c] for a in range(3) for b in range(4) for c in range(2)])
    flat = flat_bucket_ids(paths, cats)
    assert np.array_equal(flat, np.arange(24))  # lexicographic == ascending flat id
    assert np.array_equal(unflatten_bucket_ids(flat, cats), paths)
    bad = np.array([[0, -1, -1], [2, 3, 1], [3, 0, 0], [1, 4, 0]])
    assert flat_bucket_ids(bad, cats).tolist() == [24, 23, 24, 24]
    assert unflatten_bucket_ids(np.array([[5, -1]]), cats).tolist() == [[[0, 2, 1], [-1, -1, -1]]]


def test_results_contract_roundtrip(tmp_path):
    """store_results / get_recall / res.csv (search/search.py:51-63, eval/eval.py:37-77)."""
    import csv
    import numpy as np
    from learnedmetricindex_b200 import results as res
    from oracle import lmi_oracle as orc

    rs = np.random.RandomState(0)
    gt = np.stack([rs.permutation(500)[:10] + 1 for _ in range(40)]).astype(np.int64)
    anns = gt.copy().astype(np.uint32)
    anns[:, 7:] = 100_000 + np.arange(3)  # 7 of 10 right
    dists = np.sort(rs.rand(40, 10), axis=1)
    p = res.store_results(str(tmp_path / "result" / "run.h5"), "learned-index", "clip768", dists, anns, 1.5, 0.25,
                          "4-of-120", "100K")
    back = res.load_results(p)
    assert back["algo"] == "learned-index" and back["size"] == "100K" and float(back["querytime"]) == 0.25
    assert np.array_equal(back["knns"], anns) and back["knns"].dtype == np.uint32
    assert res.get_recall(anns, gt, 10) == orc.get_recall(anns, gt, 10) == 0.7
    out = res.write_res_csv([p], {"100K": gt}, str(tmp_path / "res.csv"))
    rows = list(csv.DictReader(open(out)))
    assert list(rows[0]) == res.RES_COLUMNS and float(rows[0]["recall"]) == 0.7


def test_build_configuration_expands_and_validates():
    """search/li/BuildConfiguration.py:8-135."""
    import pytest
    from learnedmetricindex_b200.builder import BuildConfiguration, kmeans_gpu

    c = BuildConfiguration(kmeans_gpu, 40, "MLP-3", 0.001, [5, 10])
    assert c.n_levels == 2 and c.epochs == [40, 40] and c.model_types == ["MLP-3", "MLP-3"]
    assert c.level_configurations[1] == (kmeans_gpu, "MLP-3", 40, 0.001, 10)
    c2 = BuildConfiguration([kmeans_gpu], [40], ["MLP-3"], [0.001], [5, 10])
    assert c2.level_configurations == c.level_configurations
    with pytest.raises(AssertionError):
        BuildConfiguration(kmeans_gpu, [40], "MLP-3", 0.001, [5, 10])  # mixed scalars and lists
    with pytest.raises(AssertionError):
        BuildConfiguration([kmeans_gpu], [40, 40, 40], ["MLP"], [0.1], [5, 10])
    with pytest.raises(AssertionError):
        BuildConfiguration(kmeans_gpu, 40, "MLP", 0.1, [])
