#!/usr/bin/env python
"""Measurements for the BASELINE.json configs other than the headline one (C2 is bench.py's own line).

  C1  the SHIPPED model (HashingTF 10000 + IDF + LogisticRegression) on 1,600 dialogues of ~375 tokens through the
      host-buffer call (p50 / p99 per batch), plus 1M such dialogues device-resident
  C3  RandomForest 100 trees x depth 10 predict on 10M dialogues, rows sharded over ranks (strong scaling)
  C4  GBT 200 iters x depth 6 on Kafka-shaped micro-batches of 4096 skewed dialogues: p50 / p99 batch latency
  C5  RandomForest training 100 trees x depth 10 on 10M x 2^18 sparse TF-IDF, histogram all-reduce over ranks:
      the same 10M rows split over the ranks (strong) and 10M rows per rank (weak)

Each ``run_cX`` returns one JSON-able dict; ``bench.py`` calls them after its headline measurement, in the same
process and at the same number of ranks.  Run directly (``python scripts/bench_configs.py --configs c3,c5``, under
torchrun for N > 1) it prints one JSON object per config.  Timing: CUDA events after warm-up, maximum over ranks;
host-path latencies: perf_counter around the synchronous C-ABI call."""
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
NF = 1 << 18


def feature_stages(nf):
    return [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
            fd.HashingTF(inputCol="filtered_words", outputCol="raw_features", numFeatures=nf)]


def fitted_idf(stages, pk, nf, ns=1 << 16):
    ns = min(ns, pk.n_rows)
    rowptr, idx, _ = fd.PipelineModel(stages).featurize(fd.PackedText(pk.data, pk.offs[:ns + 1]), cleaned=False)
    df = torch.bincount(idx.long(), minlength=nf).cpu().numpy()
    return fd.IDF(inputCol="raw_features", outputCol="features").fit_from_doc_freq(df, ns), float(rowptr[-1].item()) / ns, df


def timed(fn, steps, dev):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) / steps


class Ctx:
    """What every config needs: ranks, device, the synthetic corpus and the split-feature pool of SURVEY.md 8(d)."""

    def __init__(self):
        import torch.distributed as dist
        self.dist = dist
        self.world, self.rank, self.local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
        self.dev = torch.device("cuda", self.local)
        self.corpus = synth.SynthCorpus()
        self.top = synth.hash_terms(self.corpus.words[self.corpus.n_stop:self.corpus.n_stop + 2000] + [b""], NF)
        self.rng = np.random.default_rng(1)

    def maxreduce(self, x):
        t = torch.tensor([x], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize(self.dev)


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"])
    return 6650.0


def run_c1(ctx, rows_big=1_000_000):
    """C1: the reference's own shape.  Rank 0 only (returns None elsewhere)."""
    if ctx.rank != 0:
        return None
    import tarfile
    import tempfile
    dev = ctx.dev
    d = tempfile.mkdtemp()
    with tarfile.open(os.path.join(ROOT, "tests", "golden", "shipped_model.tar.gz")) as tf:
        tf.extractall(d, filter="data")
    model = fd.PipelineModel.load(os.path.join(d, "dialogue_classification_model"))
    h = model.handle(ctx.local)
    small, _ = ctx.corpus.sample(1600, tokens=375, device="cpu", seed=11)
    hd, ho = small.data.pin_memory(), small.offs.pin_memory()
    hp, hq, hr = (torch.empty(s_, dtype=torch.float64).pin_memory() for s_ in ((1600,), (1600, 2), (1600, 2)))
    lat = []
    for i in range(110):
        t0 = time.perf_counter()
        h.call("fd_transform_host", P(hd), P(ho), 1600, 1, P(hp), P(hq), P(hr))
        lat.append((time.perf_counter() - t0) * 1e3)
    lat = lat[10:]
    big, _ = ctx.corpus.sample(rows_big, tokens=375, device=dev, seed=12)
    n = big.n_rows
    out = [torch.empty(n, dtype=torch.float64, device=dev), torch.empty((n, 2), dtype=torch.float64, device=dev),
           torch.empty((n, 2), dtype=torch.float64, device=dev)]
    sp = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    step = lambda: h.call("fd_transform", P(big.data), P(big.offs), n, 1, P(out[0]), P(out[1]), P(out[2]), sp)
    for _ in range(3):
        step()
    ms = timed(step, 5, dev)
    lbar = float(big.offs[-1].item()) / n
    gbs = n * (lbar + 48) / ms / 1e6
    return {"config": "C1 shipped LogisticRegression model (HashingTF 10000), 375-token dialogues",
            "batch_1600_host_path_ms": {"p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99))},
            "batch_1600_dialogues_per_s": 1600 / float(np.percentile(lat, 50)) * 1e3, "rows_device_resident": n, "ms_per_pass": ms,
            "dialogues_per_s": n / ms * 1e3, "L_bar": lbar,
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm_peak(), "unit": "GB/s", "frac": gbs / hbm_peak(),
                         "bytes_per_dialogue": lbar + 48, "kernel": "k_main<LINEAR,clean>"}}


def run_c3(ctx, rows_total=10_000_000):
    """C3: RandomForest 100 x depth 10 predict, the SAME total rows split over the ranks (strong scaling)."""
    dev = ctx.dev
    n = rows_total // ctx.world
    pk, _ = ctx.corpus.sample(n, tokens=200, device=dev, seed=100 + ctx.rank)
    st = feature_stages(NF)
    idf, nnz_bar, df = fitted_idf(st, pk, NF)
    rng = np.random.default_rng(3)  # the same forest on every rank
    trees = [synth.random_tree(rng, 10, ctx.top, idf.idf, weights=df[ctx.top]) for _ in range(100)]
    model = fd.PipelineModel(st + [idf, fd.RandomForestClassificationModel(trees, NF)])
    h = model.handle(ctx.local)
    out = [torch.empty(n, dtype=torch.float64, device=dev), torch.empty((n, 2), dtype=torch.float64, device=dev),
           torch.empty((n, 2), dtype=torch.float64, device=dev)]
    sp = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    step = lambda: h.call("fd_transform", P(pk.data), P(pk.offs), n, 1, P(out[0]), P(out[1]), P(out[2]), sp)
    for _ in range(3):
        step()
    ctx.barrier()
    ms = ctx.maxreduce(timed(step, 5, dev))
    lbar = float(pk.offs[-1].item()) / n
    res = {"config": "C3 RandomForest 100 trees x depth 10 predict (split features drawn by document frequency from the 2,000 most frequent buckets)",
           "n_gpus": ctx.world, "rows_total": n * ctx.world, "ms_per_pass": ms, "dialogues_per_s": n * ctx.world / ms * 1e3,
           "scaling": "strong", "L_bar": lbar, "nnz_bar": nnz_bar, "nodes": sum(t.n_nodes for t in trees),
           "roofline_per_gpu": {"bound": "hbm", "achieved": n * (lbar + 48) / ms / 1e6, "peak": hbm_peak(), "unit": "GB/s",
                                "frac": n * (lbar + 48) / ms / 1e6 / hbm_peak(), "kernel": "k_main<RF,clean>"},
           "label1_fraction": float(out[0].mean().item())}
    del pk, out, model
    torch.cuda.empty_cache()
    return res if ctx.rank == 0 else None


def run_c4(ctx):
    """C4: GBT 200 x depth 6 on micro-batches of 4096 skewed dialogues, latency per batch.  Rank 0 only."""
    if ctx.rank != 0:
        return None
    dev = ctx.dev
    nb, bs = 64, 4096
    pk, _ = ctx.corpus.sample(nb * bs, tokens=200, device=dev, seed=7, skew=True)
    st = feature_stages(NF)
    idf, _, df = fitted_idf(st, pk, NF)
    rng = np.random.default_rng(4)
    trees = [synth.random_tree(rng, 6, ctx.top, idf.idf, regression=True, weights=df[ctx.top]) for _ in range(200)]
    model = fd.PipelineModel(st + [idf, fd.GBTClassificationModel(trees, [1.0] + [0.1] * 199, NF)])
    h = model.handle(ctx.local)
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
    res = {"config": "C4 GBT 200 iters x depth 6, micro-batches of 4096 dialogues (LogNormal lengths, median 200 tokens, clipped at 4096)",
           "batches": len(lat_host), "host_path_ms": {"p50": q(lat_host, 50), "p99": q(lat_host, 99)},
           "device_resident_ms": {"p50": q(lat_dev, 50), "p99": q(lat_dev, 99)}, "mean_batch_bytes": float(pk.offs[-1].item()) / nb,
           "host_path_dialogues_per_s_p50": bs / q(lat_host, 50) * 1e3}
    del pk, model
    torch.cuda.empty_cache()
    return res


def _c5_fit(ctx, n_local, trees, depth, profile, seed_base=500, release=True):
    """One Pipeline.fit of the RandomForest on this rank's n_local synthetic rows -> (seconds max over ranks, model, stats).
    ``release``: hand the caching allocator's blocks back to the driver afterwards, so that the next fit starts from an
    empty pool and pays every cudaMalloc again."""
    dev = ctx.dev
    planted = {"ids": list(range(ctx.corpus.n_stop + 300, ctx.corpus.n_stop + 320)), "rate": 0.05}
    pk, lab = ctx.corpus.sample(n_local, tokens=200, device=dev, seed=seed_base + ctx.rank, planted=planted)
    st = feature_stages(NF)
    tr = importlib.import_module("fraud-detection-spark-kafka-llm_b200.training")
    tr.PROFILE_DEFAULT = bool(profile)
    if ctx.world > 1:  # NCCL sets its channels up on first use: keep that out of the timed region
        ctx.dist.all_reduce(torch.zeros(1 << 20, dtype=torch.int32, device=dev))
        ctx.dist.all_reduce(torch.zeros(8, dtype=torch.float64, device=dev))
        ctx.dist.all_reduce(torch.zeros(8, dtype=torch.int64, device=dev))
    ctx.barrier()
    t0 = time.perf_counter()
    pipe = fd.Pipeline(stages=st + [fd.IDF(inputCol="raw_features", outputCol="features"),
                                    fd.RandomForestClassifier(featuresCol="features", labelCol="labels", numTrees=trees, maxDepth=depth,
                                                              seed=42, featureSubsetStrategy="auto")])
    model = pipe.fit(pk, labels=lab)
    torch.cuda.synchronize(dev)
    secs = ctx.maxreduce(time.perf_counter() - t0)
    tr.PROFILE_DEFAULT = False
    del pk, lab
    if release:
        torch.cuda.empty_cache()
    return secs, model, planted


def run_c5(ctx, rows=10_000_000, trees=100, depth=10, weak=True):
    """C5: RandomForest training.  ``strong``: `rows` in total, split over the ranks; ``weak``: `rows` per rank (at one rank
    the two are the same run).  Each is fitted three times plain (`fit_seconds` = their median, all three in
    `fit_seconds_runs`: the first starts from an empty allocator pool, the others reuse it), then once with the trainer's
    phase timers (they synchronise the device, so only the phase split of that run is reported)."""
    from sklearn.metrics import roc_auc_score
    out = {"config": f"C5 RandomForest training {trees} trees x depth {depth} on {rows} x 2^18 sparse TF-IDF, histogram all-reduce over ranks",
           "n_gpus": ctx.world}
    modes = [("strong", rows // ctx.world)] + ([("weak", rows)] if weak and ctx.world > 1 else [])
    _c5_fit(ctx, 50_000, 4, 3, profile=False)  # warm-up: kernels loaded, collectives set up (steady state is what is timed)
    for name, n_local in modes:
        secs, model, planted = _c5_fit(ctx, n_local, trees, depth, profile=False, release=False)
        rf = model.stages[-1]
        # the same fit twice more with the allocator pool kept: `fit_seconds` is the median of the three (the first pays every
        # cudaMalloc again -- with eight processes allocating tens of GB each that alone moves it by +-0.02 s)
        runs = [secs]
        for _ in range(2):
            w_secs, wmodel, _ = _c5_fit(ctx, n_local, trees, depth, profile=False, release=False)
            runs.append(w_secs)
            del wmodel
        cold_secs, warm_secs = runs[0], min(runs[1:])
        secs = float(np.median(runs))
        test, tlab = ctx.corpus.sample(20000, tokens=200, device=ctx.dev, seed=999, planted=planted)
        auc = roc_auc_score(tlab.cpu().numpy(), model.transform(test)["rawPrediction"][:, 1].cpu().numpy())
        stats = dict(getattr(rf, "trainer_stats", {}))
        stats.pop("phases_s", None)
        psecs, pmodel, _ = _c5_fit(ctx, n_local, trees, depth, profile=True)
        pstats = getattr(pmodel.stages[-1], "trainer_stats", {})
        phases = dict(pstats.get("phases_s", {}))
        phases.update({"pipeline:" + k: v for k, v in pstats.get("pipeline_phases_s", {}).items()})
        if ctx.world > 1:  # the same phases on every rank (a phase that waits for a collective shows the slowest rank's time)
            keys = sorted(phases)
            mine = torch.tensor([phases[k] for k in keys], dtype=torch.float64, device=ctx.dev)
            allp = [torch.zeros_like(mine) for _ in range(ctx.world)]
            ctx.dist.all_gather(allp, mine)
            spread = {k: [round(float(allp[r][i]), 4) for r in range(ctx.world)] for i, k in enumerate(keys)}
        else:
            spread = None
        out[name] = {"rows_total": n_local * ctx.world, "rows_per_gpu": n_local, "fit_seconds": secs, "rows_per_s": n_local * ctx.world / secs,
                     "trees_per_s": trees / secs, "nodes": sum(t.n_nodes for t in rf.trees), "test_auc": float(auc),
                     "allreduce_calls": stats.get("allreduce_calls"), "hist_bytes": stats.get("hist_bytes"),
                     "phases_s": {k: round(v, 4) for k, v in phases.items()}, "fit_seconds_with_phase_timers": psecs,
                     "fit_seconds_runs": runs, "fit_seconds_cold_pool": cold_secs, "fit_seconds_warm_pool": warm_secs,
                     "phases_s_by_rank": spread}
        del model, pmodel
    if ctx.world == 1:
        out["weak"] = out["strong"]
    return out if ctx.rank == 0 else None


RUNNERS = {"c1": run_c1, "c3": run_c3, "c4": run_c4, "c5": run_c5}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c1,c3,c4,c5")
    ap.add_argument("--c3-rows", type=int, default=10_000_000, help="total dialogues (split over ranks)")
    ap.add_argument("--c1-rows", type=int, default=1_000_000, help="dialogues of C1's device-resident pass")
    ap.add_argument("--train-rows", type=int, default=10_000_000)
    ap.add_argument("--train-trees", type=int, default=100)
    ap.add_argument("--train-depth", type=int, default=10)
    a = ap.parse_args()
    ctx = Ctx()
    torch.cuda.set_device(ctx.local)
    if ctx.world > 1:
        # NCCL prints its version banner on stdout when the communicator is created: point fd 1 at stderr until that has
        # happened, so that stdout carries the JSON line(s) and nothing else
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            ctx.dist.init_process_group("nccl", device_id=ctx.dev)
            ctx.dist.all_reduce(torch.zeros(1, device=ctx.dev))
            torch.cuda.synchronize(ctx.dev)
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    for name in a.configs.split(","):
        if name == "c3":
            res = run_c3(ctx, a.c3_rows)
        elif name == "c1":
            res = run_c1(ctx, a.c1_rows)
        elif name == "c5":
            res = run_c5(ctx, a.train_rows, a.train_trees, a.train_depth)
        else:
            res = RUNNERS[name](ctx)
        if res is not None:
            print(json.dumps(res), flush=True)
    if ctx.world > 1:
        ctx.dist.destroy_process_group()


if __name__ == "__main__":
    main()
