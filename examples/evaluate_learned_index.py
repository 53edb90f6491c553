#!/usr/bin/env python
"""The reference driver's `evaluate_learned_index` (search/search.py:173-283) on this package:
DataFrames with a 1-based index -> `BuildConfiguration` -> `LearnedIndexBuilder.build()` ->
`li.search(...)` per bucket budget -> `store_results` -> `res.csv`, on synthetic CLIP-shaped data
(no network for the SISAP LAION files, search.py:36-48) with a separate low-dimensional navigation
representation as in the notebook.  The body reads like the reference's; only the imports differ.

    python examples/evaluate_learned_index.py [--size 100000] [--n-categories 10 10] [-bp 5 10 20]
"""
import argparse
import logging
import os
import sys
import time

import numpy as np
import pandas as pd
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from learnedmetricindex_b200 import synthetic as syn  # noqa: E402
from learnedmetricindex_b200.builder import BuildConfiguration, LearnedIndexBuilder, kmeans_gpu  # noqa: E402
from learnedmetricindex_b200.results import get_recall, store_results, write_res_csv  # noqa: E402

logging.basicConfig(level=logging.INFO, format="[%(asctime)s][%(levelname)-5.5s] %(message)s")
LOG = logging.getLogger("evaluate")
algorithms = {"faiss_kmeans": kmeans_gpu}  # li.clustering.algorithms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=100_000)
    ap.add_argument("--queries", type=int, default=10_000)
    ap.add_argument("--k", default=10, type=int)
    ap.add_argument("--n-categories", nargs="+", default=[10, 10], type=int)
    ap.add_argument("--epochs", nargs="+", default=[100], type=int)
    ap.add_argument("--model-type", nargs="+", default=["MLP"])
    ap.add_argument("--lr", nargs="+", default=[0.01], type=float)
    ap.add_argument("-bp", "--buckets-perc", nargs="+", default=[5, 10, 20], type=int)
    ap.add_argument("--nav-dim", type=int, default=96)
    ap.add_argument("--out", default="result")
    args = ap.parse_args()
    kind, size = f"synthetic-pca{args.nav_dim}", str(args.size)

    # ---- data (normalised, fp16 like SISAP `emb`) and its navigation representation ---- #
    spec = syn.MixtureSpec()
    data_search_np = syn.make_vectors(args.size, spec).numpy()
    queries_search = syn.make_vectors(args.queries, spec, stream=1).numpy()
    xs = torch.from_numpy(data_search_np[: min(args.size, 100_000)]).float()
    mean = xs.mean(0, keepdim=True)
    _, _, V = torch.pca_lowrank(xs - mean, q=args.nav_dim, center=False, niter=4)
    project = lambda a: torch.nn.functional.normalize((torch.from_numpy(a).float() - mean) @ V, dim=1).numpy()
    data, queries = project(data_search_np), project(queries_search)

    s = time.time()
    data_pd = pd.DataFrame(data)
    data_pd.index += 1
    data_search = pd.DataFrame(data_search_np)
    data_search.index += 1
    config = BuildConfiguration([algorithms["faiss_kmeans"]], args.epochs, args.model_type, args.lr, args.n_categories)
    builder = LearnedIndexBuilder(data_pd, config)
    li, data_prediction, n_buckets_in_index, build_t, cluster_t = builder.build()
    LOG.info("index built: %d buckets", n_buckets_in_index)
    LOG.info("  k-means assignment  %.2f s", cluster_t)
    LOG.info("  training + routing  %.2f s", build_t)
    LOG.info("  build, wall clock   %.2f s", time.time() - s)

    n_buckets = sorted({b for b in (int((p / 100) * n_buckets_in_index) for p in args.buckets_perc) if b > 0})
    LOG.info("probe budgets to evaluate: %s", n_buckets)

    # exact ground truth (Baseline.search semantics, search/li/Baseline.py:11-16: ids + 1)
    dev = li.device
    X, Q = torch.from_numpy(data_search_np).to(dev).float(), torch.from_numpy(queries_search).to(dev).float()
    gt = torch.cat([(Q[i : i + 2048] @ X.t()).topk(args.k, dim=1).indices + 1 for i in range(0, len(Q), 2048)]).cpu().numpy()

    files = []
    for bucket in n_buckets:
        LOG.info("--- budget: %s of %d buckets per query", bucket, n_buckets_in_index)
        dists, nns, measured_time = li.search(
            data_navigation=data_pd,
            queries_navigation=queries,
            data_search=data_search,
            queries_search=queries_search,
            data_prediction=data_prediction,
            n_categories=args.n_categories,
            n_buckets=bucket,
            k=args.k,
        )
        LOG.info("  navigation (MLP + top-B)   %.6f s", measured_time["inference"])
        LOG.info("  whole call, host to host   %.6f s", measured_time["search"])
        LOG.info("  inversion + scan + merge   %.6f s", measured_time["search_within_buckets"])
        LOG.info("    of which the bucket scan %.6f s", measured_time["seq_search"])
        LOG.info("    of which the merge       %.6f s", measured_time["sort"])
        LOG.info(f"Recall@{args.k}: {get_recall(nns.astype(np.int64), gt, args.k):.4f}")
        identifier = f"{kind}-learned-index-{'_'.join(map(str, args.n_categories))}-buckets={bucket}"
        files.append(store_results(os.path.join(args.out, kind, size, f"{identifier}.h5"), "Learned-index", kind,
                                   dists, nns, build_t, measured_time["search"], identifier, size))
    csv_path = write_res_csv(files, {size: gt}, os.path.join(args.out, "res.csv"), k=args.k)
    LOG.info(f"Wrote {csv_path}")
    print(open(csv_path).read())


if __name__ == "__main__":
    main()
