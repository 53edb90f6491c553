"""The N>1 path on CPU: world_size-2 gloo run of the bucket-sharded search's host logic
(bucket partitioning, packed all-gather buffers, gather, rank-ordered merge).  The
per-rank compute is the oracle here (no GPU in this container); on a GPU box the same
plumbing carries ``LearnedIndex.search_device`` + ``lmi_merge_ranks`` (tests/test_gpu_parity.py)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_buckets_balances_and_is_deterministic():
    from learnedmetricindex_b200.sharded import owner_of_bucket, partition_buckets

    rs = np.random.RandomState(1)
    sizes = rs.randint(50_000, 120_000, 120)
    for world in (1, 2, 4, 8):
        parts = partition_buckets(sizes, world)
        assert sorted(b for p in parts for b in p) == list(range(120))
        loads = np.array([sizes[p].sum() for p in parts], dtype=np.float64)
        assert loads.max() / loads.mean() < 1.05  # LPT on 120 similar buckets is near-perfect
        assert parts == partition_buckets(sizes, world)
        own = owner_of_bucket(parts, 120)
        assert all(own[b] == r for r, p in enumerate(parts) for b in p)
    # weighted by expected probes
    w = rs.rand(120) + 0.1
    parts = partition_buckets(sizes, 4, w)
    loads = np.array([(sizes[p] * w[p]).sum() for p in parts])
    assert loads.max() / loads.mean() < 1.1
    # more ranks than buckets: some ranks own nothing, nothing is lost
    parts = partition_buckets([5, 3], 4)
    assert sorted(b for p in parts for b in p) == [0, 1]
    # ranks of different speed (GPUs of one box differ under their power caps): finish times are balanced
    speeds = np.array([1.0, 0.9, 1.1, 1.0, 0.92, 1.05, 1.0, 0.95])
    parts = partition_buckets(sizes, 8, speeds=speeds)
    assert sorted(b for p in parts for b in p) == list(range(120))
    t = np.array([sizes[p].sum() for p in parts]) / speeds
    assert t.max() / t.mean() < 1.05
    t_blind = np.array([sizes[p].sum() for p in partition_buckets(sizes, 8)]) / speeds
    assert t.max() < t_blind.max()


def test_slice_bounds_cover_every_bucket_in_rank_order():
    from learnedmetricindex_b200.sharded import slice_bounds

    sizes = np.array([0, 1, 7, 10, 83_333, 1_000_003])
    for shares in (None, [1.0, 0.9, 1.1, 1.0], [3.0, 1.0, 1.0, 1.0]):
        lo, hi = slice_bounds(sizes, 4, shares)
        assert (lo[0] == 0).all() and (hi[-1] == sizes).all()
        assert (lo[1:] == hi[:-1]).all() and (hi >= lo).all()  # contiguous, in rank order, nothing lost
        if shares is not None:
            frac = (hi - lo)[:, -1] / sizes[-1]
            np.testing.assert_allclose(frac, np.array(shares) / np.sum(shares), atol=1e-5)


def _numpy_merge(buf, out_d, out_i):
    from oracle import lmi_oracle as orc

    ds = [buf.rank_views(w)[0].numpy() for w in range(buf.world)]
    ns = [buf.rank_views(w)[1].numpy().view(np.uint32) for w in range(buf.world)]
    d, i = orc.merge_ranks(ds, ns, buf.k)
    out_d.copy_(torch.from_numpy(d))
    out_i.copy_(torch.from_numpy(i.view(np.int32)))


def _case():
    from tests.golden_util import seeded_layers
    from oracle import lmi_oracle as orc

    rs = np.random.RandomState(11)
    n, nq, d, C = 3000, 97, 64, 9
    X = rs.standard_normal((n, d)).astype(np.float32)
    X = (X / np.linalg.norm(X, axis=1, keepdims=True)).astype(np.float16)
    Q = rs.standard_normal((nq, d)).astype(np.float32)
    Q = (Q / np.linalg.norm(Q, axis=1, keepdims=True)).astype(np.float16)
    layers = seeded_layers(3, "MLP-6", d, C, 4.0)
    pred = orc.predict(layers, orc.to_f32_tensor(X)).reshape(-1, 1)
    pred[pred[:, 0] == 4, 0] = 5  # an empty bucket
    return X, Q, layers, pred, C


def _worker(rank, world, port, B, k, ret):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from learnedmetricindex_b200.sharded import GatherBuffers, gather_and_merge, owner_of_bucket, partition_buckets
        from oracle import lmi_oracle as orc

        X, Q, layers, pred, C = _case()
        sizes = np.bincount(pred[:, 0], minlength=C)
        parts = partition_buckets(sizes, world)
        own = owner_of_bucket(parts, C)
        mine = own[pred[:, 0]] == rank
        ids = np.arange(1, X.shape[0] + 1)
        d, i, _, _ = orc.search(orc.OracleIndex(layers), Q, X[mine], Q, pred[mine], [C], n_buckets=B, k=k,
                                ids=ids[mine])
        kk = d.shape[1]
        buf = GatherBuffers(Q.shape[0], kk, world, rank, "cpu")
        ld, li, _ = buf.local_views()  # no probe ranks from the CPU oracle: ties keep the lower rank
        ld.copy_(torch.from_numpy(d))
        li.copy_(torch.from_numpy(i.view(np.int32)))
        out_d, out_i = gather_and_merge(buf, None, merge_fn=_numpy_merge)
        ret[rank] = (out_d.numpy().copy(), out_i.numpy().view(np.uint32).copy())
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("B,k", [(1, 10), (3, 10), (4, 7)])
def test_sharded_search_world2_gloo(B, k):
    from oracle import lmi_oracle as orc
    from tests.golden_util import assert_search_parity

    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), B, k, ret), nprocs=world, join=True)
    X, Q, layers, pred, C = _case()
    exp_d, exp_i, _, _ = orc.search(orc.OracleIndex(layers), Q, X, Q, pred, [C], n_buckets=B, k=k)
    for r in range(world):
        d, i = ret[r]
        assert d.shape == exp_d.shape
        if B == 1:
            # one probe per query: the owner's list verbatim (fillers included)
            np.testing.assert_array_equal(d, exp_d)
            real = exp_d < 1e38
            assert np.array_equal(i[real], exp_i[real])
        else:
            # real candidates sort before any filler on every rank, so wherever the
            # single-process result holds a real candidate the sharded one is identical
            real = exp_d < 1e38
            np.testing.assert_array_equal(d[real], exp_d[real])
            assert_search_parity(np.where(real, d, exp_d), np.where(real, i, exp_i), exp_d, exp_i, what=f"rank {r}")
    np.testing.assert_array_equal(ret[0][0], ret[1][0])
    np.testing.assert_array_equal(ret[0][1], ret[1][1])


def _order_worker(rank, world, port, nq, B, ret):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from learnedmetricindex_b200.sharded import gather_bucket_order, query_slice

        full = torch.arange(nq * B, dtype=torch.int32).reshape(nq, B) % 17
        start, stop, per = query_slice(nq, world, rank)
        got = gather_bucket_order(full[start:stop].contiguous(), nq, B)
        ret[rank] = got.numpy().copy()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("nq,B", [(97, 3), (2, 1), (64, 4)])
def test_query_sharded_navigation_gathers_full_order(nq, B):
    """Each rank contributes the bucket order of its query slice; every rank ends up with all of it."""
    from learnedmetricindex_b200.sharded import query_slice

    world = 2
    covered = []
    for r in range(world):
        a, b, per = query_slice(nq, world, r)
        covered += list(range(a, b))
        assert b - a <= per
    assert covered == list(range(nq))
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_order_worker, args=(world, _free_port(), nq, B, ret), nprocs=world, join=True)
    full = (np.arange(nq * B, dtype=np.int32).reshape(nq, B) % 17)
    for r in range(world):
        np.testing.assert_array_equal(ret[r], full)
