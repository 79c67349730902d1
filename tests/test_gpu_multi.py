"""Two ranks, NCCL, the real CUDA trainer: rows sharded over two B200s must grow exactly the trees (and fit exactly the IDF
vector) that one GPU grows from all the rows, given the same bag weights and feature subsets.  Histograms and document
frequencies are integers, so the cross-rank sums are exact.  Skipped on a box with fewer than two GPUs."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import multi_rank_worker as w

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_nccl_ranks_grow_the_single_gpu_forest(fd, tmp_path):
    out = str(tmp_path / "two_ranks.npz")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29571", os.path.join(ROOT, "tests", "multi_rank_worker.py"), out]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-3000:]
    two = np.load(out)
    pk, lab, bag = w.workload(fd, torch.device("cuda", 0))
    one = w.fit(fd, pk, lab, bag, 0, w.N_ROWS, 0)
    assert np.array_equal(one.stages[-2].idf, two["idf"]), "IDF vector (docFreq all-reduce) differs"
    trees = one.stages[-1].trees
    assert int(two["n_trees"]) == len(trees) and max(t.n_nodes for t in trees) > 15
    for i, t in enumerate(trees):
        for name in ("feature", "threshold", "left", "right", "prediction", "stats", "gain"):
            assert np.array_equal(np.asarray(getattr(t, name)), two[f"t{i}_{name}"]), (i, name)
