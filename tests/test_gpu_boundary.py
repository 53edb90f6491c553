"""Boundary entry points INTEGRATION.md tells a user of the reference to call, exercised with objects
shaped exactly like the reference's (search/li/model.py:89-99, search/li/LearnedIndex.py:23-39), plus the
device-side layout builder and the fp16-storage policy.  Needs a B200: run with ``-m gpu``."""
import pickle
import warnings

import numpy as np
import pytest
import torch

from oracle import lmi_oracle as orc
from tests.golden_util import assert_search_parity

pytestmark = pytest.mark.gpu


class RefShapedModel(torch.nn.Module):
    """``li.model.Model`` (search/li/model.py:89-99): the layers live in ``self.layers``, an
    ``nn.Sequential`` of Linear/ReLU -- state-dict keys ``layers.0.weight``, ``layers.2.weight``, ..."""

    def __init__(self, in_dim, hidden, out_dim):
        super().__init__()
        dims = [in_dim] + list(hidden) + [out_dim]
        mods = []
        for i in range(len(dims) - 1):
            mods.append(torch.nn.Linear(dims[i], dims[i + 1]))
            if i + 2 < len(dims):
                mods.append(torch.nn.ReLU())
        self.layers = torch.nn.Sequential(*mods)

    def forward(self, x):
        return self.layers(x)


class RefShapedNetwork:
    """``li.model.NeuralNetwork`` (search/li/model.py:130-164) as far as the query path reads it: ``.model``
    (the module above), ``.device``, an optimiser and a loss -- all of which ``save_as_pickle`` pickles."""

    def __init__(self, in_dim, hidden, out_dim, seed):
        torch.manual_seed(seed)
        self.device = torch.device("cpu")
        self.model = RefShapedModel(in_dim, hidden, out_dim)
        self.optimizer = torch.optim.Adam(self.model.parameters(), lr=0.01)
        self.loss = torch.nn.CrossEntropyLoss()


class RefShapedIndex:
    """``li.LearnedIndex.LearnedIndex`` (search/li/LearnedIndex.py:23-39): three attributes."""

    def __init__(self, root_model, internal_models, bucket_paths):
        self.root_model, self.internal_models, self.bucket_paths = root_model, internal_models, bucket_paths


def _layers_of(net):
    sd = net.model.state_dict()
    idx = sorted({int(k.split(".")[1]) for k in sd})
    return [(sd[f"layers.{i}.weight"].numpy().copy(), sd[f"layers.{i}.bias"].numpy().copy()) for i in idx]


def _unit_fp16(rs, n, d):
    x = rs.standard_normal((n, d)).astype(np.float32)
    return (x / np.linalg.norm(x, axis=1, keepdims=True)).astype(np.float16)


def test_from_reference_on_a_pickled_reference_shaped_index(tmp_path):
    """`LearnedIndex.from_reference(pickle.load(...))` -- the swap INTEGRATION.md describes for
    search/search.py:226-259 -- on a 1-level and a 2-level index whose models carry the reference's
    `layers.N.weight` state-dict keys."""
    from learnedmetricindex_b200 import LearnedIndex

    assert sorted(RefShapedModel(8, [4], 3).state_dict()) == ["layers.0.bias", "layers.0.weight", "layers.2.bias",
                                                              "layers.2.weight"]
    rs = np.random.RandomState(3)
    n, nq, d = 6000, 200, 64
    X, Q = _unit_fp16(rs, n, d), _unit_fp16(rs, nq, d)

    # ---- 1 level, MLP-5 shape (two hidden layers: keys layers.0 / layers.2 / layers.4)
    C = 7
    root = RefShapedNetwork(d, [256, 128], C, seed=1)
    with torch.no_grad():
        for p in root.model.parameters():
            p.mul_(4.0)
    ref = RefShapedIndex(root, {}, [(i,) for i in range(C)])
    path = tmp_path / "index.pkl"
    with open(path, "wb") as f:
        pickle.dump(ref, f)  # search/li/utils.py:14-29
    with open(path, "rb") as f:
        li = LearnedIndex.from_reference(pickle.load(f))
    layers = _layers_of(root)
    for (W, b), (We, be) in zip(li.root_model.layers, layers):
        assert np.array_equal(W.cpu().numpy(), We) and np.array_equal(b.cpu().numpy(), be)
    pred = orc.predict(layers, orc.to_f32_tensor(X)).reshape(-1, 1)  # LearnedIndexBuilder.py:76
    d1, i1, _ = li.search(None, Q, X, Q, pred, [C], n_buckets=3, k=10)
    exp_d, exp_i, _, exp_order = orc.search(orc.OracleIndex(layers), Q, X, Q, pred, [C], n_buckets=3, k=10)
    same = (li.last_bucket_order.cpu().numpy() == exp_order[:, :, 0]).all(axis=1)
    assert same.mean() > 0.98
    assert_search_parity(d1[same], i1[same], exp_d[same], exp_i[same], what="from_reference 1-level")

    # ---- 2 levels [3, 2]: internal models keyed by right-padded paths (LearnedIndexBuilder.py:330-352)
    cats = [3, 2]
    root2 = RefShapedNetwork(d, [128], cats[0], seed=2)
    internal = {(j, -1): RefShapedNetwork(d, [128], cats[1], seed=10 + j) for j in range(cats[0])}
    paths = [(j, b) for j in range(cats[0]) for b in range(cats[1])]
    li2 = LearnedIndex.from_reference(pickle.loads(pickle.dumps(RefShapedIndex(root2, internal, paths))))
    l_root, l_int = _layers_of(root2), {p: _layers_of(m) for p, m in internal.items()}
    xt = orc.to_f32_tensor(X)
    pred2 = np.full((n, 2), -1, dtype=np.int64)
    pred2[:, 0] = orc.predict(l_root, xt)
    for j in range(cats[0]):
        rows = np.flatnonzero(pred2[:, 0] == j)
        pred2[rows, 1] = orc.predict(l_int[(j, -1)], xt[rows])
    d2, i2, _ = li2.search(None, Q, X, Q, pred2, cats, n_buckets=3, k=10)
    e_d, e_i, _, e_order = orc.search(orc.OracleIndex(l_root, l_int, paths), Q, X, Q, pred2, cats, n_buckets=3, k=10)
    got_paths, _ = li2._precompute_bucket_order(Q, 3, cats)
    same2 = (got_paths == e_order).all(axis=(1, 2))
    assert same2.mean() > 0.98
    assert_search_parity(d2[same2], i2[same2], e_d[same2], e_i[same2], what="from_reference 2-level")


def test_layout_from_device_equals_from_host_and_drops_unlabelled_rows():
    """`DeviceBucketLayout.from_device` (data already on the GPU; bench.py and the 100M path) builds bit
    for bit what `from_host` builds, for fp16 and f32 sources and a dimensionality that needs padding; rows
    with a negative label (no bucket -- EMPTY_VALUE paths of the reference) are dropped, not an error."""
    from learnedmetricindex_b200 import DeviceBucketLayout

    rs = np.random.RandomState(11)
    for d, dtype in ((768, np.float16), (100, np.float32), (64, np.float16)):
        n, C = 5000, 9
        X = _unit_fp16(rs, n, d).astype(dtype)
        labels = rs.randint(0, C, n).astype(np.int64)
        labels[labels == 4] = 5  # an empty bucket
        ids = rs.permutation(np.arange(100, 100 + n)).astype(np.int64)
        host = DeviceBucketLayout.from_host(X, labels, C, ids=ids)
        dev = DeviceBucketLayout.from_device(torch.from_numpy(X).cuda(), torch.from_numpy(labels).cuda(), C,
                                             ids=torch.from_numpy(ids).cuda(), chunk=1536)
        assert torch.equal(host.vectors, dev.vectors) and torch.equal(host.ids, dev.ids)
        assert np.array_equal(host.bucket_off_host, dev.bucket_off_host)
        assert (host.d, host.d_data, host.n_nonempty) == (dev.d, dev.d_data, dev.n_nonempty) == (-(-d // 64) * 64, d, C - 1)
        assert host.inexact_values == dev.inexact_values == 0  # fp16-representable data: nothing was rounded

        drop = rs.rand(n) < 0.1
        lab2 = np.where(drop, -1, labels)
        dev2 = DeviceBucketLayout.from_device(torch.from_numpy(X).cuda(), torch.from_numpy(lab2).cuda(), C,
                                              ids=torch.from_numpy(ids).cuda())
        ref2 = DeviceBucketLayout.from_host(X[~drop], labels[~drop], C, ids=ids[~drop])
        assert dev2.n == int((~drop).sum())
        assert torch.equal(dev2.vectors, ref2.vectors) and torch.equal(dev2.ids, ref2.ids)
        assert np.array_equal(dev2.bucket_off_host, ref2.bucket_off_host)


def test_f32_data_that_fp16_cannot_hold_is_surfaced_not_silently_rounded():
    """The reference driver's default (`preprocess=True`, search/search.py:142-144) L2-normalises to f32 and
    scans THAT frame with faiss in f32.  Here storage is fp16: data that is not fp16-representable is rounded
    (2^-11 relative per component).  That must never be silent: a warning by default, an error with
    allow_fp16_rounding=False, and -- opted in -- distances within the stated bound of the f32 oracle."""
    from learnedmetricindex_b200 import LearnedIndex
    from tests.golden_util import seeded_layers

    rs = np.random.RandomState(21)
    n, nq, d, C, B = 8000, 300, 128, 6, 3
    X = rs.standard_normal((n, d)).astype(np.float32)
    X /= np.linalg.norm(X, axis=1, keepdims=True)  # f32 unit vectors: NOT fp16-representable
    Q = rs.standard_normal((nq, d)).astype(np.float32)
    Q /= np.linalg.norm(Q, axis=1, keepdims=True)
    layers = seeded_layers(5, "MLP-4", d, C, scale=4.0)
    pred = orc.predict(layers, orc.to_f32_tensor(X)).reshape(-1, 1)
    paths = [(i,) for i in range(C)]

    with pytest.raises(ValueError, match="not representable in fp16"):
        LearnedIndex(layers, {}, paths, allow_fp16_rounding=False).search(None, Q, X, Q, pred, [C], n_buckets=B)
    with pytest.warns(UserWarning, match="not representable in fp16"):
        LearnedIndex(layers, {}, paths).search(None, Q, X, Q, pred, [C], n_buckets=B)
    # fp16-exact data with f32 queries: the QUERY rounding is surfaced too
    X16 = X.astype(np.float16)
    with pytest.warns(UserWarning, match="queries_search"):
        LearnedIndex(layers, {}, paths).search(None, Q, X16, Q, pred, [C], n_buckets=B)
    with warnings.catch_warnings():
        warnings.simplefilter("error")  # fp16-representable data AND queries: no warning at all
        LearnedIndex(layers, {}, paths).search(None, Q.astype(np.float16), X16, Q.astype(np.float16), pred, [C],
                                               n_buckets=B)
        li = LearnedIndex(layers, {}, paths, allow_fp16_rounding=True)
        dd, ii, _ = li.search(None, Q, X, Q, pred, [C], n_buckets=B, k=10)
    # opted in: the oracle on the f32 data (the reference's arithmetic); the MLP still sees the f32 queries, so
    # the probed buckets agree; distances move by at most the rounding bound, ids only where that reorders
    exp_d, exp_i, _, exp_order = orc.search(orc.OracleIndex(layers), Q, X, Q, pred, [C], n_buckets=B, k=10)
    same = (li.last_bucket_order.cpu().numpy() == exp_order[:, :, 0]).all(axis=1)
    assert same.mean() > 0.99
    assert np.abs(dd[same] - exp_d[same]).max() < 1e-3  # |q||x| * 2 * 2^-11 worst case; typically ~1e-5
    assert np.mean([len(set(a) & set(b)) for a, b in zip(ii[same].tolist(), exp_i[same].tolist())]) > 9.8


def test_builder_accepts_the_reference_clustering_contract():
    """`BuildConfiguration(clustering_algorithms=...)` takes the reference's callables
    (clustering/faiss_kmeans.py:8-24): (np.ndarray f32, n_clusters, parameters) -> (object, labels)."""
    from learnedmetricindex_b200 import synthetic as syn
    from learnedmetricindex_b200.builder import BuildConfiguration, LearnedIndexBuilder

    seen = {}

    def numpy_kmeans(data, n_clusters, parameters):
        assert isinstance(data, np.ndarray) and data.dtype == np.float32
        seen["shape"] = data.shape
        x = data.astype(np.float32)
        rs = np.random.RandomState(0)
        c = x[rs.choice(len(x), n_clusters, replace=False)]
        for _ in range(5):
            lab = (x @ c.T).argmax(axis=1)
            for j in range(n_clusters):
                if (lab == j).any():
                    c[j] = x[lab == j].mean(axis=0)
        return None, (x @ c.T).argmax(axis=1)

    spec = syn.MixtureSpec(n_centres=40)
    X = syn.make_vectors(6000, spec).numpy()
    cfg = BuildConfiguration(numpy_kmeans, 100, "MLP", 0.01, [5])
    li, pred, n_buckets, _, _ = LearnedIndexBuilder(X, cfg).build()
    assert seen["shape"] == (6000, 768) and n_buckets == 5 and pred.shape == (6000, 1)
    assert len(np.unique(pred[:, 0])) == 5


def test_headline_size_matches_the_oracle_on_a_query_sample():
    """The headline configuration (10M x 768, 120 buckets, MLP-4, B = 4, k = 10) against the CPU oracle on
    512 queries, with the north_star tolerances ASSERTED: probed-bucket lists equal except probability ties
    <= 1e-6, distances within 1e-3 relative, ids equal except where two objects tie within f32 rounding.
    Skipped without ~70 GB of free host RAM (the oracle scans an f32 copy of the vectors)."""
    import types

    import psutil

    import bench
    from learnedmetricindex_b200 import DeviceBucketLayout, LearnedIndex

    free, _ = torch.cuda.mem_get_info()
    if free < 40e9:
        pytest.skip("needs ~20 GB of free HBM")
    if psutil.virtual_memory().available < 70e9:
        pytest.skip("needs ~70 GB of free host RAM for the oracle's f32 copy")
    C, nq, B = 120, 512, 4
    args = types.SimpleNamespace(n=10_000_000, nq=nq, buckets=C, centres=2000, noise=0.25, latent=24, train_steps=100)
    w = bench.build_workload(args, torch.device("cuda", 0), 0, 1)
    li = LearnedIndex(w["layers"], {}, [(i,) for i in range(C)])
    li.attach_layout(DeviceBucketLayout(w["vectors"], torch.from_numpy(w["bucket_off"]), w["ids"]))
    Qh = w["Q"].cpu().numpy()
    d, i, _ = li.search(None, Qh, None, Qh, None, [C], n_buckets=B, k=10)
    order_gpu = li.last_bucket_order.cpu().numpy()
    cb = bench.CpuBaseline(w, C)
    del w
    _, d_cpu, i_cpu, order_cpu = cb.run(B)
    parity = cb.explain(d_cpu, i_cpu, order_cpu, d, i, order_gpu)
    assert parity["queries"] == nq and parity["unexplained"] == 0, parity
    assert parity["bucket_order_rows_equal"] > 0.995 and parity["id_agreement"] > 0.999, parity


def test_rank_merge_orders_equal_distances_by_probe_rank_then_gpu_rank():
    """lmi_merge_ranks (after an all-gather) with the per-entry probe ranks lmi_merge_topk emits: equal distances
    from different GPUs come out in (probe rank, GPU rank) order -- the single-GPU search's stable order over the
    rounds (LearnedIndex.py:131-134) -- and without probe ranks in GPU-rank order."""
    from learnedmetricindex_b200 import _native

    L = _native.lib()
    rs = np.random.RandomState(4)
    world, nq, k = 4, 257, 10
    # few distinct distance values -> many ties across ranks; every rank's list ascending in (dist, probe)
    d = rs.randint(0, 6, size=(world, nq, k)).astype(np.float64) / 4.0
    p = rs.randint(0, 5, size=(world, nq, k)).astype(np.uint16)
    ids = rs.randint(1, 1 << 30, size=(world, nq, k)).astype(np.uint32)
    for w in range(world):
        for q in range(nq):
            o = np.lexsort((p[w, q], d[w, q]))
            d[w, q], p[w, q], ids[w, q] = d[w, q][o], p[w, q][o], ids[w, q][o]
    dev = torch.device("cuda")
    td, ti, tp = (torch.from_numpy(a).to(dev) for a in (d, ids.view(np.int32), p.view(np.int16)))
    out_d = torch.empty((nq, k), dtype=torch.float64, device=dev)
    out_i = torch.empty((nq, k), dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    for with_probe in (True, False):
        _native.check(L.lmi_merge_ranks(td.data_ptr(), ti.data_ptr(), tp.data_ptr() if with_probe else None, nq * k,
                                        nq * k, nq * k, world, nq, k, out_d.data_ptr(), out_i.data_ptr(), st))
        torch.cuda.synchronize()
        got_d, got_i = out_d.cpu().numpy(), out_i.cpu().numpy().view(np.uint32)
        for q in range(nq):
            fd = d[:, q].reshape(-1)
            fp = p[:, q].reshape(-1).astype(np.int64) if with_probe else np.zeros(world * k, dtype=np.int64)
            fr = np.repeat(np.arange(world), k)
            fpos = np.tile(np.arange(k), world)
            o = np.lexsort((fpos, fr, fp, fd))[:k]
            np.testing.assert_array_equal(got_d[q], fd[o])
            np.testing.assert_array_equal(got_i[q], ids[:, q].reshape(-1)[o])


def test_peer_merge_and_peer_gather_kernels_on_one_gpu():
    """lmi_merge_ranks_peer and lmi_gather_peer take a device array of per-rank base pointers, so they can be
    proven on ONE GPU: `world` separately allocated blocks stand in for the peers' symmetric-memory blocks (on a
    multi-GPU box the same addresses are NVLink mappings).  The fused exchange+merge must equal lmi_merge_ranks on
    the gathered blocks -- ties by (probe rank, GPU rank), or by GPU rank without probe ranks -- and the gather
    must reproduce every block byte for byte."""
    from learnedmetricindex_b200 import _native

    L = _native.lib()
    rs = np.random.RandomState(9)
    dev = torch.device("cuda")
    st = torch.cuda.current_stream().cuda_stream
    for world, nq, k in [(2, 257, 10), (8, 1000, 10), (5, 33, 7), (32, 64, 10), (3, 1, 10)]:
        d = rs.randint(0, 6, size=(world, nq, k)).astype(np.float64) / 4.0
        p = rs.randint(0, 5, size=(world, nq, k)).astype(np.uint16)
        ids = rs.randint(1, 1 << 30, size=(world, nq, k)).astype(np.uint32)
        for w in range(world):
            for q in range(nq):
                o = np.lexsort((p[w, q], d[w, q]))
                d[w, q], p[w, q], ids[w, q] = d[w, q][o], p[w, q][o], ids[w, q][o]
        off_i, off_p = nq * k * 8, nq * k * 12
        per_rank = -(-(off_p + nq * k * 2) // 16) * 16
        blocks = []
        for w in range(world):
            raw = np.zeros(per_rank, dtype=np.uint8)
            raw[:off_i] = d[w].reshape(-1).view(np.uint8)
            raw[off_i:off_p] = ids[w].reshape(-1).view(np.uint8)
            raw[off_p:off_p + nq * k * 2] = p[w].reshape(-1).view(np.uint8)
            blocks.append(torch.from_numpy(raw).to(dev))
        base = torch.tensor([b.data_ptr() for b in blocks], dtype=torch.int64, device=dev)

        # gather: dst[w * bytes + i] = rank_base[w][i]
        gathered = torch.full((world * per_rank,), 0xAB, dtype=torch.uint8, device=dev)
        _native.check(L.lmi_gather_peer(base.data_ptr(), per_rank, world, gathered.data_ptr(), st), "lmi_gather_peer")
        torch.cuda.synchronize()
        assert torch.equal(gathered, torch.cat(blocks)), (world, nq, k)

        td, ti, tp = (torch.from_numpy(a).to(dev) for a in (d, ids.view(np.int32), p.view(np.int16)))
        for with_probe in (True, False):
            exp_d = torch.empty((nq, k), dtype=torch.float64, device=dev)
            exp_i = torch.empty((nq, k), dtype=torch.int32, device=dev)
            _native.check(L.lmi_merge_ranks(td.data_ptr(), ti.data_ptr(), tp.data_ptr() if with_probe else None,
                                            nq * k, nq * k, nq * k, world, nq, k, exp_d.data_ptr(), exp_i.data_ptr(),
                                            st), "lmi_merge_ranks")
            got_d = torch.full((nq, k), -1.0, dtype=torch.float64, device=dev)
            got_i = torch.full((nq, k), -1, dtype=torch.int32, device=dev)
            _native.check(L.lmi_merge_ranks_peer(base.data_ptr(), off_i, off_p if with_probe else -1, world, nq, k,
                                                 got_d.data_ptr(), got_i.data_ptr(), st), "lmi_merge_ranks_peer")
            torch.cuda.synchronize()
            assert torch.equal(got_d, exp_d) and torch.equal(got_i, exp_i), (world, nq, k, with_probe)
            # and against the definition, independent of either kernel
            gd, gi = got_d.cpu().numpy(), got_i.cpu().numpy().view(np.uint32)
            for q in range(0, nq, max(1, nq // 40)):
                fd = d[:, q].reshape(-1)
                fp = p[:, q].reshape(-1).astype(np.int64) if with_probe else np.zeros(world * k, dtype=np.int64)
                o = np.lexsort((np.tile(np.arange(k), world), np.repeat(np.arange(world), k), fp, fd))[:k]
                np.testing.assert_array_equal(gd[q], fd[o])
                np.testing.assert_array_equal(gi[q], ids[:, q].reshape(-1)[o])
