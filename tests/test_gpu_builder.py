"""GPU build path: ``LearnedIndexBuilder(data, config).build()`` with the
reference's contract (search/li/LearnedIndexBuilder.py:43-107), then search on what it built."""
import numpy as np
import pytest
import torch

from oracle import lmi_oracle as orc

pytestmark = pytest.mark.gpu


def _data(n, nq, seed=0, centres=60):
    from learnedmetricindex_b200 import synthetic as syn

    spec = syn.MixtureSpec(n_centres=centres, seed=2023 + seed)
    return syn.make_vectors(n, spec).numpy(), syn.make_vectors(nq, spec, stream=1).numpy()


def test_kmeans_assignment_is_nearest_centroid():
    from learnedmetricindex_b200.builder import kmeans_gpu

    X, _ = _data(20_000, 1)
    x = torch.from_numpy(X).cuda()
    cent, lab = kmeans_gpu(x, 16, None)
    assert cent.shape == (16, 768) and lab.shape == (20_000,) and lab.dtype == torch.int64
    d2 = torch.cdist(x.float(), cent)
    exp = d2.argmin(dim=1)
    agree = (exp == lab).float().mean().item()
    assert agree > 0.999  # differences only at (near-)equidistant points
    assert torch.unique(lab).numel() == 16
    # Lloyd iterations reduce the quantisation error vs the random initialisation
    g = torch.Generator(device="cuda").manual_seed(1)
    rnd = x[torch.randperm(x.shape[0], generator=g, device="cuda")[:16]].float()
    assert d2.min(dim=1).values.pow(2).mean() < torch.cdist(x.float(), rnd).min(dim=1).values.pow(2).mean()


def test_build_one_level_and_search():
    from learnedmetricindex_b200.builder import BuildConfiguration, LearnedIndexBuilder, kmeans_gpu

    n, nq, C = 30_000, 400, 12
    X, Q = _data(n, nq)
    cfg = BuildConfiguration(kmeans_gpu, 150, "MLP-4", 0.01, [C])
    li, pred, n_buckets, build_t, cluster_t = LearnedIndexBuilder(X, cfg).build()
    assert pred.shape == (n, 1) and pred.dtype == np.int64 and n_buckets == C and build_t >= cluster_t >= 0
    assert li.bucket_paths == [(b,) for b in range(C)]
    assert np.array_equal(pred[:, 0], li.root_model.predict(X))  # LearnedIndexBuilder.py:76
    assert len(np.unique(pred[:, 0])) == C  # trained until every category is predicted (:189)
    # the built index is a regular index: parity with the oracle on ITS weights, and useful recall
    layers = [(W.cpu().numpy(), b.cpu().numpy()) for W, b in li.root_model.layers]
    d, i, _ = li.search(None, Q, X, Q, pred, [C], n_buckets=3, k=10)
    exp_d, exp_i, _, exp_order = orc.search(orc.OracleIndex(layers), Q, X, Q, pred, [C], n_buckets=3, k=10)
    same = (li.last_bucket_order.cpu().numpy() == exp_order[:, :, 0]).all(axis=1)
    assert same.mean() > 0.99
    assert (i[same] == exp_i[same]).mean() > 0.999
    gt_d, gt_i = orc.brute_force(Q, X)
    assert orc.get_recall(i.astype(np.int64), gt_i, 10) > 0.6
    d_all, i_all, _ = li.search(None, Q, X, Q, pred, [C], n_buckets=C, k=10)
    assert orc.get_recall(i_all.astype(np.int64), gt_i, 10) > 0.999


def test_build_two_levels_and_search():
    from learnedmetricindex_b200.builder import BuildConfiguration, LearnedIndexBuilder, kmeans_gpu

    n, nq, cats = 30_000, 300, [4, 5]
    X, Q = _data(n, nq, seed=1)
    cfg = BuildConfiguration([kmeans_gpu], [120], ["MLP"], [0.01], cats)
    li, pred, n_buckets, _, _ = LearnedIndexBuilder(X, cfg).build()
    assert pred.shape == (n, 2) and (pred >= 0).all()
    assert set(li.internal_models) == {(j, -1) for j in range(cats[0])}
    assert n_buckets == len(li.bucket_paths) == len({tuple(r) for r in pred.tolist()})
    # bucket_paths are numbered 0..len(unique)-1 per parent (LearnedIndexBuilder.py:276-278)
    for j in range(cats[0]):
        kids = sorted(p[1] for p in li.bucket_paths if p[0] == j)
        assert kids == list(range(len(kids)))
    B = 6
    d, i, _ = li.search(None, Q, X, Q, pred, cats, n_buckets=B, k=10)
    to_np = lambda m: [(W.cpu().numpy(), b.cpu().numpy()) for W, b in m.layers]
    idx = orc.OracleIndex(to_np(li.root_model), {p: to_np(m) for p, m in li.internal_models.items()}, li.bucket_paths)
    exp_d, exp_i, _, exp_order = orc.search(idx, Q, X, Q, pred, cats, n_buckets=B, k=10)
    got_paths, _ = li._precompute_bucket_order(Q, B, cats)
    same = (got_paths == exp_order).all(axis=(1, 2))
    assert same.mean() > 0.99
    assert (i[same] == exp_i[same]).mean() > 0.999
    gt_d, gt_i = orc.brute_force(Q, X)
    assert orc.get_recall(i.astype(np.int64), gt_i, 10) > 0.6
