"""Worker of tests/test_gpu_multi.py, started once per rank by torch.distributed.run with the NCCL backend.

Every rank builds the same seeded dialogues, keeps its contiguous shard of the rows, and fits
Tokenizer -> StopWordsRemover -> HashingTF -> IDF -> RandomForest through ``Pipeline.fit`` with the real CUDA backend
(``CudaTreeBackend``): document frequencies and histograms are summed over the ranks with NCCL all-reduce.  Bag weights
and feature subsets are given explicitly (the parity contract's inputs), so the trees must equal the one-GPU fit.
Rank 0 writes the fitted IDF vector and the trees to ``argv[1]`` (npz)."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

N_ROWS, NF, N_TREES, DEPTH, K = 6000, 1 << 12, 5, 6, 40


def workload(fd, device):
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    corpus = synth.SynthCorpus(n_words=3000, seed=9)
    planted = {"ids": list(range(corpus.n_stop + 100, corpus.n_stop + 115)), "rate": 0.08}
    pk, lab = corpus.sample(N_ROWS, tokens=50, device=device, seed=4, planted=planted)
    bag = np.random.default_rng(6).poisson(1.0, size=(N_TREES, N_ROWS)).astype(np.uint8)
    return pk, lab, bag


def subset_fn(t, node):
    return np.random.default_rng([17, t, node]).permutation(NF)[:K].astype(np.int32)


def stages(fd):
    return [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
            fd.HashingTF(inputCol="filtered_words", outputCol="raw_features", numFeatures=NF),
            fd.IDF(inputCol="raw_features", outputCol="features"),
            fd.RandomForestClassifier(featuresCol="features", labelCol="labels", numTrees=N_TREES, maxDepth=DEPTH, seed=1,
                                      featureSubsetStrategy="sqrt")]


def fit(fd, pk, lab, bag, lo, hi, device):
    shard = fd.PackedText(pk.data, pk.offs[lo:hi + 1])
    model = fd.Pipeline(stages=stages(fd)).fit(shard, labels=lab[lo:hi], device=device,
                                               bag_weights=torch.from_numpy(np.ascontiguousarray(bag[:, lo:hi])), subset_fn=subset_fn)
    return model


def dump(model, path):
    rf = model.stages[-1]
    arrs = {"idf": model.stages[-2].idf, "n_trees": np.asarray(len(rf.trees))}
    for i, t in enumerate(rf.trees):
        for name in ("feature", "threshold", "left", "right", "prediction", "stats", "gain"):
            arrs[f"t{i}_{name}"] = np.asarray(getattr(t, name))
    np.savez(path, **arrs)


def main():
    import fdb200 as fd
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    pk, lab, bag = workload(fd, dev)
    lo, hi = N_ROWS * rank // world, N_ROWS * (rank + 1) // world
    model = fit(fd, pk, lab, bag, lo, hi, local)
    stats = model.stages[-1].trainer_stats
    assert stats["allreduce_calls"] >= 1 and stats["n_total"] == N_ROWS
    if rank == 0:
        dump(model, sys.argv[1])
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
