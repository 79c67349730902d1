#!/usr/bin/env python
"""Secondary measurements for the BASELINE.json configs that are not the bench.py line (C3, C4, C5).

  C3  RandomForest 100 trees x depth 10 predict, rows sharded over ranks (run under torchrun for N > 1)
  C4  GBT 200 iters x depth 6 on Kafka-shaped micro-batches of 4096 skewed dialogues: p50 / p99 batch latency
  C5  RandomForest training (scaled: see --train-*), histogram all-reduce over ranks

Prints one JSON object per config; timing with CUDA events after warm-up (C4 host path: perf_counter around the
synchronous C-ABI call)."""
import argparse
import ctypes as C
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import fdb200 as fd  # noqa: E402

synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
P = lambda t: C.c_void_p(t.data_ptr())


def feature_stages(nf):
    return [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
            fd.HashingTF(inputCol="filtered_words", outputCol="raw_features", numFeatures=nf)]


def fitted_idf(stages, pk, nf, ns=1 << 16):
    ns = min(ns, pk.n_rows)
    rowptr, idx, _ = fd.PipelineModel(stages).featurize(fd.PackedText(pk.data, pk.offs[:ns + 1]), cleaned=False)
    df = torch.bincount(idx.long(), minlength=nf).cpu().numpy()
    return fd.IDF(inputCol="raw_features", outputCol="features").fit_from_doc_freq(df, ns), float(rowptr[-1].item()) / ns


def timed(fn, steps, dev):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c3,c4,c5")
    ap.add_argument("--c3-rows", type=int, default=10_000_000, help="total dialogues (split over ranks)")
    ap.add_argument("--train-rows", type=int, default=1_000_000)
    ap.add_argument("--train-trees", type=int, default=20)
    ap.add_argument("--train-depth", type=int, default=6)
    ap.add_argument("--train-profile", action="store_true")
    ap.add_argument("--train-weak", action="store_true", help="--train-rows is per GPU (weak scaling) instead of the total")
    a = ap.parse_args()
    import torch.distributed as dist
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL prints its version banner on stdout when the communicator is created: point fd 1 at stderr until that has
        # happened, so that stdout carries the JSON line(s) and nothing else
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.all_reduce(torch.zeros(1, device=dev))
            torch.cuda.synchronize(dev)
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    nf = 1 << 18
    corpus = synth.SynthCorpus()
    top = synth.hash_terms(corpus.words[corpus.n_stop:corpus.n_stop + 2000] + [b""], nf)
    rng = np.random.default_rng(1)
    want = a.configs.split(",")

    def maxreduce(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if "c1" in want and rank == 0:
        # C1: the reference's own shape -- the SHIPPED model (HashingTF 10000 + IDF + LogisticRegression) on 1,600
        # dialogues of about 375 tokens (the CSV itself is not distributable), plus 1M such dialogues for throughput
        import tarfile, tempfile
        d = tempfile.mkdtemp()
        with tarfile.open(os.path.join(ROOT, "tests", "golden", "shipped_model.tar.gz")) as tf:
            tf.extractall(d, filter="data")
        model = fd.PipelineModel.load(os.path.join(d, "dialogue_classification_model"))
        h = model.handle(local)
        small, _ = corpus.sample(1600, tokens=375, device="cpu", seed=11)
        hd, ho = small.data.pin_memory(), small.offs.pin_memory()
        hp, hq, hr = (torch.empty(s_, dtype=torch.float64).pin_memory() for s_ in ((1600,), (1600, 2), (1600, 2)))
        lat = []
        for i in range(60):
            t0 = time.perf_counter()
            h.call("fd_transform_host", P(hd), P(ho), 1600, 1, P(hp), P(hq), P(hr))
            lat.append((time.perf_counter() - t0) * 1e3)
        lat = lat[10:]
        big, _ = corpus.sample(1_000_000, tokens=375, device=dev, seed=12)
        n = big.n_rows
        out = [torch.empty(n, dtype=torch.float64, device=dev), torch.empty((n, 2), dtype=torch.float64, device=dev),
               torch.empty((n, 2), dtype=torch.float64, device=dev)]
        sp = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        step = lambda: h.call("fd_transform", P(big.data), P(big.offs), n, 1, P(out[0]), P(out[1]), P(out[2]), sp)
        for _ in range(3):
            step()
        ms = timed(step, 5, dev)
        print(json.dumps({"config": "C1 shipped LR model, 375-token dialogues", "batch_1600_host_path_ms": {"p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99))},
                          "batch_1600_dialogues_per_s": 1600 / float(np.percentile(lat, 50)) * 1e3, "rows_1M_ms_per_pass": ms,
                          "rows_1M_dialogues_per_s": n / ms * 1e3, "L_bar": float(big.offs[-1].item()) / n,
                          "algorithmic_GBps": n * (float(big.offs[-1].item()) / n + 48) / ms / 1e6}))
        del big, out, model

    if "c3" in want:
        n = a.c3_rows // world
        pk, _ = corpus.sample(n, tokens=200, device=dev, seed=100 + rank)
        st = feature_stages(nf)
        idf, nnz_bar = fitted_idf(st, pk, nf)
        trees = [synth.random_tree(rng, 10, top, idf.idf) for _ in range(100)]
        model = fd.PipelineModel(st + [idf, fd.RandomForestClassificationModel(trees, nf)])
        h = model.handle(local)
        out = [torch.empty(n, dtype=torch.float64, device=dev), torch.empty((n, 2), dtype=torch.float64, device=dev),
               torch.empty((n, 2), dtype=torch.float64, device=dev)]
        sp = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        step = lambda: h.call("fd_transform", P(pk.data), P(pk.offs), n, 1, P(out[0]), P(out[1]), P(out[2]), sp)
        for _ in range(3):
            step()
        if world > 1:
            dist.barrier()
        ms = maxreduce(timed(step, 5, dev))
        if rank == 0:
            lbar = float(pk.offs[-1].item()) / n
            print(json.dumps({"config": "C3 RF 100x depth 10 predict", "n_gpus": world, "rows_total": n * world, "ms_per_pass": ms,
                              "dialogues_per_s": n * world / ms * 1e3, "scaling": "strong", "L_bar": lbar, "nnz_bar": nnz_bar,
                              "algorithmic_GBps_per_gpu": n * (lbar + 48) / ms / 1e6, "nodes": sum(t.n_nodes for t in trees),
                              "label1_fraction": float(out[0].mean().item())}))
        del pk, out, model

    if "c4" in want and rank == 0:
        nb, bs = 64, 4096
        pk, _ = corpus.sample(nb * bs, tokens=200, device=dev, seed=7, skew=True)
        st = feature_stages(nf)
        idf, _ = fitted_idf(st, pk, nf)
        trees = [synth.random_tree(rng, 6, top, idf.idf, regression=True) for _ in range(200)]
        model = fd.PipelineModel(st + [idf, fd.GBTClassificationModel(trees, [1.0] + [0.1] * 199, nf)])
        h = model.handle(local)
        hdata = pk.data.cpu().pin_memory()
        hoffs = pk.offs.cpu().pin_memory()
        hp = torch.empty(bs, dtype=torch.float64).pin_memory()
        hq = torch.empty((bs, 2), dtype=torch.float64).pin_memory()
        hr = torch.empty((bs, 2), dtype=torch.float64).pin_memory()
        dp, dq, dr = (torch.empty_like(x, device=dev) for x in (hp, hq, hr))
        sp = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        lat_host, lat_dev = [], []
        for rep in range(3):
            for b in range(nb):
                offs_b = hoffs[b * bs:(b + 1) * bs + 1]
                t0 = time.perf_counter()
                h.call("fd_transform_host", P(hdata), P(offs_b), bs, 1, P(hp), P(hq), P(hr))
                t1 = time.perf_counter()
                doffs = pk.offs[b * bs:(b + 1) * bs + 1]
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                h.call("fd_transform", P(pk.data), P(doffs), bs, 1, P(dp), P(dq), P(dr), sp)
                e1.record()
                torch.cuda.synchronize(dev)
                if rep > 0:
                    lat_host.append((t1 - t0) * 1e3)
                    lat_dev.append(e0.elapsed_time(e1))
        q = lambda v, p: float(np.percentile(v, p))
        print(json.dumps({"config": "C4 GBT 200x depth 6, micro-batches of 4096 (LogNormal lengths, median 200, clip 4096)",
                          "batches": len(lat_host), "host_path_ms": {"p50": q(lat_host, 50), "p99": q(lat_host, 99)},
                          "device_resident_ms": {"p50": q(lat_dev, 50), "p99": q(lat_dev, 99)},
                          "mean_batch_bytes": float(pk.offs[-1].item()) / nb,
                          "host_path_dialogues_per_s_p50": bs / q(lat_host, 50) * 1e3}))
        del pk, model

    if "c5" in want:
        n = a.train_rows if a.train_weak else a.train_rows // world
        planted = {"ids": list(range(corpus.n_stop + 300, corpus.n_stop + 320)), "rate": 0.05}
        pk, lab = corpus.sample(n, tokens=200, device=dev, seed=500 + rank, planted=planted)
        st = feature_stages(nf)
        if world > 1:  # NCCL builds its communicator on the first collective: keep that out of the timed region
            dist.all_reduce(torch.zeros(1, device=dev))
            dist.barrier()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        pipe = fd.Pipeline(stages=st + [fd.IDF(inputCol="raw_features", outputCol="features"),
                                        fd.RandomForestClassifier(featuresCol="features", labelCol="labels", numTrees=a.train_trees,
                                                                  maxDepth=a.train_depth, seed=42, featureSubsetStrategy="auto")])
        if a.train_profile:
            import importlib as _il
            _tr = _il.import_module("fraud-detection-spark-kafka-llm_b200.training")
            _orig = _tr.ForestTrainer.__init__
            def _init(self, *aa, **kw):
                _orig(self, *aa, **kw)
                self.profile = True
            _tr.ForestTrainer.__init__ = _init
        model = pipe.fit(pk, labels=lab)
        torch.cuda.synchronize(dev)
        secs = maxreduce(time.perf_counter() - t0)
        rf = model.stages[-1]
        test, tlab = corpus.sample(20000, tokens=200, device=dev, seed=999, planted=planted)
        out = model.transform(test)
        from sklearn.metrics import roc_auc_score
        auc = roc_auc_score(tlab.cpu().numpy(), out["rawPrediction"][:, 1].cpu().numpy())
        if rank == 0:
            print(json.dumps({"config": f"C5 RF train {a.train_trees} trees x depth {a.train_depth}, {n * world} rows x 2^18 ({'weak' if a.train_weak else 'strong'} scaling)",
                              "n_gpus": world, "fit_seconds_wall": secs, "trees_per_s": a.train_trees / secs,
                              "nodes": sum(t.n_nodes for t in rf.trees), "test_auc": float(auc), "trainer": getattr(rf, "trainer_stats", {})}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
