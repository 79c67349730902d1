"""Fuzz of the CUDA trainer against the oracle on random small problems (random sizes, depths, bag weights, subset
sizes, min-instance settings): trees must be identical.  Usage: fuzz_trainer.py [iterations]"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import fdb200 as fd  # noqa: E402,F401
from test_training_cpu import _csr, _oracle_forest, _same_tree, _toy  # noqa: E402

training = importlib.import_module("fraud-detection-spark-kafka-llm_b200.training")
lib = importlib.import_module("fraud-detection-spark-kafka-llm_b200._lib")
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 40
h = lib.Handle(0)
rng = np.random.default_rng(99)
bad = 0
for it in range(iters):
    n = int(rng.integers(50, 1200))
    f0 = int(rng.integers(12, 120))
    n_trees = int(rng.integers(1, 6))
    depth = int(rng.integers(1, 10))
    rows_idx, rows_val, labels, f = _toy(n, f0, seed=int(rng.integers(0, 1 << 30)))
    if rng.random() < 0.3:   # label noise: deeper, less pure trees
        flip = rng.random(n) < 0.25
        labels = np.where(flip, 1 - labels, labels).astype(labels.dtype)
    bag = rng.poisson(1.0, size=(n_trees, n)).astype(np.uint8) if rng.random() < 0.8 else np.ones((n_trees, n), np.uint8)
    k = int(rng.integers(1, f + 1))
    seed_s = int(rng.integers(0, 1 << 30))
    subsets = {}

    def subset_fn(t, node, _k=k, _s=seed_s, _f=f, _c=subsets):
        return _c.setdefault((t, node), np.random.default_rng([_s, t, node]).permutation(_f)[:_k].astype(np.int32))

    tr = training.ForestTrainer(training.CudaTreeBackend(h), f, n_trees=n_trees, max_depth=depth, feature_subset_strategy="sqrt", seed=1)
    rowptr, idx, val = (x.cuda() for x in _csr(rows_idx, rows_val))
    trees = tr.fit(rowptr, idx, val, torch.from_numpy(np.asarray(labels)).cuda(), bag_weights=torch.from_numpy(bag).cuda(), subset_fn=subset_fn)
    _, fi, fv = _csr(rows_idx, rows_val)
    thresholds = training.find_splits(fi.numpy(), fv.numpy(), f, float(n), min(32, n) - 1)
    exp = _oracle_forest(rows_idx, rows_val, np.asarray(labels), f, thresholds, bag, subset_fn, depth)
    try:
        for a, o in zip(trees, exp):
            _same_tree(a, o)
    except AssertionError:
        bad += 1
        print("MISMATCH iteration", it, "n", n, "f", f, "trees", n_trees, "depth", depth, "k", k, flush=True)
print("iterations", iters, "mismatches", bad, flush=True)
sys.exit(1 if bad else 0)
