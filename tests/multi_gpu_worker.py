"""Worker of tests/test_gpu_multi.py (run under torchrun, one process per GPU): the bucket-sharded
search with both exchange paths -- peer-memory merge (lmi_merge_ranks_peer) and NCCL all-gather +
lmi_merge_ranks -- must equal the unsharded search on the same index, bit for bit."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    from learnedmetricindex_b200 import DeviceBucketLayout, LearnedIndex
    from learnedmetricindex_b200 import synthetic as syn
    from learnedmetricindex_b200.sharded import ShardedLearnedIndex, partition_buckets

    n, nq, C = 40_000, 1000, 16
    spec = syn.MixtureSpec(n_centres=300)
    X = syn.make_vectors(n, spec)  # CPU generator: identical on every rank
    Q = syn.make_vectors(nq, spec, stream=1).to(dev)
    layers, pred = syn.build_standin_index(X, C, steps=60)
    layers = syn.layers_to_numpy(layers)
    full = DeviceBucketLayout.from_host(X.numpy(), pred.numpy()[:, 0], C, device=dev)
    parts = partition_buckets(np.diff(full.bucket_off_host), world)
    ref = LearnedIndex(layers, {}, [(i,) for i in range(C)])
    ref.attach_layout(full)
    for B, k in [(1, 10), (3, 10), (5, 7)]:
        exp_d, exp_i = ref.search_device(Q, Q, [C], B, k)
        for exchange in ("peer", "nccl"):
            li = LearnedIndex(layers, {}, [(i,) for i in range(C)])
            li.attach_layout(full.restrict_to_buckets(parts[rank]))
            eng = ShardedLearnedIndex(li, exchange=exchange)
            for rep in range(3):  # repeated: the peer path alternates between its two blocks
                d, i = eng.search_device(Q, Q, [C], B, k)
                assert torch.equal(d, exp_d) and torch.equal(i, exp_i), (exchange, B, k, rep, rank)
            if exchange == "peer":  # the whole sharded search as one graph replay (two alternating instances)
                for rep in range(3):
                    dg, ig, _ = eng.search_graphed(Q, Q, [C], B, k)
                    assert torch.equal(dg, exp_d) and torch.equal(ig, exp_i), ("graph", B, k, rep, rank)
            # host-facing API (queries uploaded 1/world per rank and all-gathered; graph replay on the peer path)
            dh, ih, _ = eng.search(None, Q.cpu().numpy(), None, Q.cpu().numpy(), None, [C], n_buckets=B, k=k)
            assert np.array_equal(dh, exp_d.cpu().numpy()) and np.array_equal(ih.view(np.int32), exp_i.cpu().numpy())
    # graph cache of the sharded search: keyed on shapes only (a fresh device tensor per call replays the same
    # graph on every rank), bounded, and an entry that was evicted and returns -- after more batch shapes than
    # the local workspace caches hold -- is captured again on live workspaces
    from learnedmetricindex_b200.sharded import SHARDED_GRAPH_CACHE

    li = LearnedIndex(layers, {}, [(i,) for i in range(C)])
    li.attach_layout(full.restrict_to_buckets(parts[rank]))
    eng = ShardedLearnedIndex(li, exchange="peer")
    exp_d, exp_i = ref.search_device(Q, Q, [C], 3, 10)
    for rep in range(4):
        q = Q.clone() if (rep + rank) % 2 else Q  # ranks disagree on which calls see a new address
        dg, ig, _ = eng.search_graphed(q, q, [C], 3, 10)
        assert torch.equal(dg, exp_d) and torch.equal(ig, exp_i), ("graph key", rep, rank)
        assert len(eng._graphs) == 1
    for m in range(1, SHARDED_GRAPH_CACHE + 3):
        q = Q[: nq - m].contiguous()
        e_d, e_i = ref.search_device(q, q, [C], 3, 10)
        dg, ig, _ = eng.search_graphed(q, q, [C], 3, 10)
        assert torch.equal(dg, e_d) and torch.equal(ig, e_i), ("graph cycle", m, rank)
    assert len(eng._graphs) == SHARDED_GRAPH_CACHE
    for rep in range(3):
        dg, ig, _ = eng.search_graphed(Q, Q, [C], 3, 10)
        assert torch.equal(dg, exp_d) and torch.equal(ig, exp_i), ("graph replay after eviction", rep, rank)
    del eng, li
    # a batch that does not divide by the number of ranks (padded navigation / upload slices)
    for exchange in ("peer", "nccl"):
        li = LearnedIndex(layers, {}, [(i,) for i in range(C)])
        li.attach_layout(full.restrict_to_buckets(parts[rank]))
        eng = ShardedLearnedIndex(li, exchange=exchange)
        Qodd = Q[:999].contiguous()
        exp_d, exp_i = ref.search_device(Qodd, Qodd, [C], 3, 10)
        for rep in range(2):
            dh, ih, _ = eng.search(None, Qodd.cpu().numpy(), None, Qodd.cpu().numpy(), None, [C], n_buckets=3, k=10)
            assert np.array_equal(dh, exp_d.cpu().numpy()) and np.array_equal(ih.view(np.int32), exp_i.cpu().numpy())
    # equal row slices of every bucket: same result for k <= 10 (distances always; ids on real candidates)
    li = LearnedIndex(layers, {}, [(i,) for i in range(C)])
    li.attach_layout(full.restrict_to_slice(rank, world))
    eng = ShardedLearnedIndex(li)
    for B, k in [(1, 10), (4, 10), (3, 5)]:
        exp_d, exp_i = ref.search_device(Q, Q, [C], B, k)
        d, i = eng.search_device(Q, Q, [C], B, k)
        assert torch.equal(d, exp_d), ("rows", B, k, rank)
        real = exp_d < 1e38
        assert torch.equal(i[real], exp_i[real]), ("rows ids", B, k, rank)
    # exact distance ties ACROSS ranks: every vector exists twice, the copies live in neighbouring buckets (so mostly on
    # different GPUs), and both copies of a near neighbour reach a query's top-k with the same f32 distance.  The rank
    # merge must order them as the unsharded search does (by probe rank) -- bit-equal ids, not just equal distances
    X2 = torch.cat([X, X]).numpy()
    pred2 = np.concatenate([pred.numpy()[:, 0], (pred.numpy()[:, 0] + 1) % C])
    full2 = DeviceBucketLayout.from_host(X2, pred2, C, device=dev)
    parts2 = partition_buckets(np.diff(full2.bucket_off_host), world)
    ref2 = LearnedIndex(layers, {}, [(i,) for i in range(C)])
    ref2.attach_layout(full2)
    for B, k in [(4, 10), (6, 7)]:
        exp_d, exp_i = ref2.search_device(Q, Q, [C], B, k)
        ties = int((exp_d[:, 1:] == exp_d[:, :-1]).sum().item())
        assert ties > nq // 4, f"the duplicated index should produce exact ties ({ties})"
        for exchange in ("peer", "nccl"):
            li = LearnedIndex(layers, {}, [(i,) for i in range(C)])
            li.attach_layout(full2.restrict_to_buckets(parts2[rank]))
            eng = ShardedLearnedIndex(li, exchange=exchange)
            d, i = eng.search_device(Q, Q, [C], B, k)
            assert torch.equal(d, exp_d) and torch.equal(i, exp_i), ("ties", exchange, B, k, rank)
            if exchange == "peer":
                dg, ig, _ = eng.search_graphed(Q, Q, [C], B, k)
                assert torch.equal(dg, exp_d) and torch.equal(ig, exp_i), ("ties graph", B, k, rank)
    dist.barrier()
    if rank == 0:
        print("MULTI_GPU_OK", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
