#!/usr/bin/env python
"""bench.py -- dialogues/sec of the fused PipelineModel.transform on B200 (BASELINE.json metric).

Workload (config C2, the one the metric is quoted on): synthetic 1M dialogues x 200 tokens, Tokenizer ->
StopWordsRemover -> HashingTF(2^18) -> IDF -> DecisionTree(depth 5) predict.  A "step" is one pass of the
hot path over the whole batch.  `value` = dialogues/s with inputs resident in HBM; `e2e` = the same through
the C-ABI call with HOST buffers (H2D + kernels + D2H inside the timed region); `roofline` = algorithmic
bytes of the dominant kernel / its measured duration against the measured HBM copy peak; `cpu_baseline` =
the multi-threaded C port of the CPU oracle on this box's host cores (Spark local[*] itself: unavailable
offline).  Launch: `python bench.py --gpus N --steps K --warmup W` (N > 1 under torchrun, row-sharded,
no collective, weak scaling).  `--impl reference` times the CPU port alone.
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "dialogues/sec TF-IDF+tree classify"
UNIT = "dialogues/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=1_000_000, help="dialogues per GPU")
    ap.add_argument("--tokens", type=int, default=200)
    ap.add_argument("--cpu-rows", type=int, default=1_000_000, help="bounded sample for the CPU baseline")
    ap.add_argument("--stages", action="store_true", help="(kept for old command lines: the stage timings are always taken)")
    ap.add_argument("--configs", default="c1,c3,c4,c5", help="other BASELINE.json configs measured after the headline one ('' = none)")
    ap.add_argument("--config-rows", type=int, default=10_000_000, help="rows of C3 (total) and C5 (total / per GPU)")
    return ap.parse_args()


def spark_probe():
    """look for the reference's own runtime on this box instead of assuming it is absent."""
    import shutil
    out = {"java": shutil.which("java"), "spark_submit": shutil.which("spark-submit"), "pyspark": None}
    try:
        import importlib.util
        spec = importlib.util.find_spec("pyspark")
        out["pyspark"] = None if spec is None else (spec.origin or "found")
    except Exception as e:  # a broken install counts as absent
        out["pyspark"] = None
        out["error"] = type(e).__name__
    out["available"] = bool(out["java"] and out["pyspark"])
    return out


def kernel_source_hash():
    """sha256 over the sources of the dialogue kernels: profiles/traffic.json is only quoted while it matches."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "fraud-detection-spark-kafka-llm_b200", "csrc")
    for f in ("fd_main.cu", "fd_rows_dev.cuh", "fd_rows.cuh", "fd_common.cuh"):
        with open(os.path.join(d, f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._halt = index, [], set(), None, threading.Event()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                     "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}
            while not self._halt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
                time.sleep(0.05)
        except Exception as e:  # NVML missing: report it rather than fail the bench
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def build_workload(fd, rows, tokens, device, seed, featurizer="gpu"):
    """-> (model, PackedText on device, corpus, stats).  ``featurizer`` says who counts the document frequencies of the
    IDF fit and hashes the tree's split terms: "gpu" = this repo's CUDA path, "oracle" = the CPU oracle (the reference
    arm uses that one, so that nothing of the CUDA library runs anywhere in it; both give the same model, bit for bit)."""
    import importlib

    import torch
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    nf = 1 << 18
    corpus = synth.SynthCorpus()
    pk, _ = corpus.sample(rows, tokens=tokens, device=device, seed=seed)
    feat_stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"),
                   fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
                   fd.HashingTF(inputCol="filtered_words", outputCol="raw_features", numFeatures=nf)]
    # IDF fitted on a 65,536-dialogue sample of the same corpus
    ns = min(rows, 1 << 16)
    # the tree splits on features drawn from the 2,000 most frequent buckets (plus the empty token), in
    # proportion to their document frequency, as a trained tree prefers words that occur -- so that paths actually branch
    words = corpus.words[corpus.n_stop:corpus.n_stop + 2000] + [b""]
    if featurizer == "gpu":
        tf_model = fd.PipelineModel(feat_stages)
        sample = fd.PackedText(pk.data, pk.offs[:ns + 1])
        rowptr, idx, _ = tf_model.featurize(sample, cleaned=False)
        df = torch.bincount(idx.long(), minlength=nf).cpu().numpy()
        nnz_bar = float(rowptr[-1].item()) / ns
        top = synth.hash_terms(words, nf)
        del tf_model
    else:
        from oracle import c_oracle
        from oracle import spark_oracle as so
        co = c_oracle.COracle(feat_stages[1].getOrDefault("stopWords"), False, nf, False, None, None)
        offs_np = pk.offs[:ns + 1].cpu().numpy()
        data_np = pk.data[:int(offs_np[-1]) + 32].cpu().numpy()
        rp, ix, _ = co.featurize(data_np, offs_np, True)
        df = np.bincount(ix[:int(rp[-1])], minlength=nf)
        nnz_bar = float(rp[-1]) / ns
        top = np.asarray([so.hashing_tf_index(w, nf) for w in words], dtype=np.int64)
    idf_model = fd.IDF(inputCol="raw_features", outputCol="features").fit_from_doc_freq(df, ns)
    rng = np.random.default_rng(seed)
    tree = synth.random_tree(rng, 5, top, idf_model.idf, weights=np.asarray(df)[top])
    model = fd.PipelineModel(feat_stages + [idf_model, fd.DecisionTreeClassificationModel(tree, nf)])
    total_bytes = int(pk.offs[-1].item() - pk.offs[0].item())
    stats = {"L_bar": total_bytes / rows, "nnz_bar": nnz_bar, "text_bytes": total_bytes}
    return model, pk, corpus, stats


def workload_config(n, tokens, stats, parallelism):
    """The ``config`` object of the JSON line -- the same for both arms."""
    return {"workload": f"C2: {n} dialogues/GPU x {tokens} tokens, HashingTF 2^18 + IDF + DecisionTree depth 5 predict",
            "L_bar_bytes": stats["L_bar"], "nnz_bar": stats["nnz_bar"], "parallelism": parallelism,
            "l2": "inputs (%.2f GB/GPU) larger than L2" % (stats["text_bytes"] / 1e9), "seed": 20261106}


def cpu_port(model, data_np, offs_np, threads):
    """Time the C port of the oracle (oracle/oracle.c) over the given dialogues -> (rows/s, seconds)."""
    from oracle import c_oracle
    from oracle import spark_oracle as so
    p = model._plan
    c = p.classifier
    trees = [so.Tree(t.feature, t.threshold, t.left, t.right, t.stats[:, :2]) for t in c.trees]
    co = c_oracle.COracle(p.stop.getOrDefault("stopWords"), False, int(p.hashing.getOrDefault("numFeatures")), False,
                          p.idf.idf, {"kind": c.kind, "trees": trees, "weights": list(c.treeWeights)})
    n = offs_np.size - 1
    co.transform(data_np, offs_np[:min(n, 20000) + 1], True, threads)  # warm
    t0 = time.perf_counter()
    out = co.transform(data_np, offs_np, True, threads)
    dt = time.perf_counter() - t0
    return n / dt, dt, out


def main():
    a = parse()
    import torch
    import torch.distributed as dist

    import fdb200 as fd

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cores = os.cpu_count() or 1

    if a.impl == "reference":
        # CPU arm: the reference's own implementation (Spark local[*]) is unavailable offline; the C port of the
        # oracle stands in, all host threads, on a bounded sample of the same workload.  Rank 0 only.
        if rank != 0:
            return
        dev = "cuda" if torch.cuda.is_available() else "cpu"  # torch only generates the synthetic text here
        rows = min(a.rows, a.cpu_rows)
        model, pk, _, stats = build_workload(fd, rows, a.tokens, dev, 20261106, featurizer="oracle")
        data_np, offs_np = pk.data.cpu().numpy(), pk.offs.cpu().numpy()
        for _ in range(a.warmup):
            cpu_port(model, data_np, offs_np[:20001], cores)
        t = []
        for _ in range(a.steps):
            t.append(cpu_port(model, data_np, offs_np, cores)[1])
        ms = 1000.0 * float(np.mean(t))
        v = rows / (ms / 1000.0)
        cfg = workload_config(rows, a.tokens, stats, f"rows over {cores} host threads")
        probe = spark_probe()
        cfg["note"] = ("Spark local[*] " + ("found on this box but not wired in" if probe["available"] else "unavailable on this box (no JVM/pyspark)")
                       + "; CPU port of the oracle, oracle/oracle.c")
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
                          "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                          "dtype": "u8/f64", "data": "synthetic", "config": cfg,
                          "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                           "sample": f"{rows} of the workload's dialogues per step"},
                          "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "spark_probe": probe}))
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
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
    # ---- latency configs first (C1's 1,600-row batch, C4's micro-batches): a call of a fraction of a millisecond is
    #      measured in a process that holds only its own model, as a serving process does -- the same call takes up to twice
    #      as long once this process also holds the gigabytes of pinned and device buffers of the throughput measurements
    want = [c for c in a.configs.split(",") if c]
    cfgs = {}
    bc = ctx = None
    if want:
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import bench_configs as bc
        ctx = bc.Ctx()

    def run_config(name):
        t0 = time.perf_counter()
        try:
            if name == "c3":
                res = bc.run_c3(ctx, a.config_rows)
            elif name == "c5":
                res = bc.run_c5(ctx, a.config_rows)
            else:
                res = bc.RUNNERS[name](ctx)
        except Exception as e:  # a failed side measurement must not take the headline line with it
            res = {"error": f"{type(e).__name__}: {e}"[:300]} if rank == 0 else None
        if res is not None:
            res["wall_s"] = round(time.perf_counter() - t0, 1)
            cfgs[name] = res

    for name in want:
        if name in ("c1", "c4"):
            run_config(name)
    model, pk, corpus, stats = build_workload(fd, a.rows, a.tokens, dev, 20261106 + rank)
    h = model.handle(local)
    lib = h.lib
    n = pk.n_rows
    pred = torch.empty(n, dtype=torch.float64, device=dev)
    prob = torch.empty((n, 2), dtype=torch.float64, device=dev)
    raw = torch.empty((n, 2), dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev)
    sp = C.c_void_p(stream.cuda_stream)
    P = lambda t: C.c_void_p(t.data_ptr())

    def step():
        h.call("fd_transform", P(pk.data), P(pk.offs), n, 1, P(pred), P(prob), P(raw), sp)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(a.warmup, 3)):
        step()
    h.call("fd_status", sp)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = h.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(a.steps):
        step()
    ev1.record(stream)
    barrier()
    launches = h.launch_count() - l0
    el_ms = ev0.elapsed_time(ev1)
    t = torch.tensor([el_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    el_ms = float(t.item())
    ms_per_step = el_ms / a.steps
    value = world * n / (ms_per_step / 1000.0)

    # ---- end to end through the host-buffer C-ABI call (pinned host memory in, host memory out)
    numa = {}
    with fd.numa_local(local, numa):  # pinned pages on the GPU's own NUMA node (first touch happens inside)
        hdata = torch.empty(pk.data.numel(), dtype=torch.uint8, pin_memory=True)
        hdata.copy_(pk.data)
        hoffs = torch.empty(pk.offs.numel(), dtype=torch.int64, pin_memory=True)
        hoffs.copy_(pk.offs)
        hp = torch.zeros(n, dtype=torch.float64, pin_memory=True)
        hq = torch.zeros((n, 2), dtype=torch.float64, pin_memory=True)
        hr = torch.zeros((n, 2), dtype=torch.float64, pin_memory=True)
        torch.cuda.synchronize(dev)

    def e2e_step():
        h.call("fd_transform_host", P(hdata), P(hoffs), n, 1, P(hp), P(hq), P(hr))

    for _ in range(2):
        e2e_step()
    barrier()
    e2e_steps = max(3, min(a.steps, 5))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize(dev)
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * n / float(te.item())
    clocks = sampler.stop()
    assert torch.equal(hp, pred.cpu()), "host-buffer path and device path disagree"

    # ---- what the host side of this box delivers at this N: the same bytes as bare pinned copies, all ranks at once
    dst = torch.empty_like(pk.data)
    barrier()
    t0 = time.perf_counter()
    for _ in range(3):
        dst.copy_(hdata, non_blocking=True)
    torch.cuda.synchronize(dev)
    th = torch.tensor([(time.perf_counter() - t0) / 3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(th, op=dist.ReduceOp.MAX)
    h2d_gbs = hdata.numel() / float(th.item()) / 1e9     # per rank, slowest rank
    e2e_gbs = (stats["text_bytes"] + 8 * (n + 1)) / float(te.item()) / 1e9
    del dst

    stages = stage_times(model, pk, dev, stats, peaks()[0])
    del pred, prob, raw
    torch.cuda.empty_cache()

    out = None
    if rank == 0:
        peak, peak_src = peaks()
        bytes_per_row = stats["L_bar"] + 8.0 + 40.0  # text + offset + three result columns (fused)
        achieved = n * bytes_per_row / (ms_per_step / 1000.0) / 1e9
        traffic = None
        issue = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            with open(tp) as f:
                cap = json.load(f)
            # the capture is only quoted while the kernel sources are the ones it was taken from
            if cap.get("kernel_source_sha") == kernel_source_hash():
                per_dialogue = cap.get("k_main_dt_clean_bytes_per_dialogue")
                traffic = None if per_dialogue is None else per_dialogue * n  # DRAM bytes of one launch over this batch
                ipd = cap.get("k_main_dt_clean_warp_instructions_per_dialogue")
                if ipd is not None and clocks.get("sm_mhz"):
                    # what actually bounds the kernel: warp-instruction issue (148 SMs x 4 schedulers x 1 instruction per
                    # clock); instructions per dialogue come from the committed ncu capture, the rate from this run
                    sms = torch.cuda.get_device_properties(dev).multi_processor_count
                    peak_issue = sms * 4 * clocks["sm_mhz"] * 1e6
                    ach_issue = ipd * n / (ms_per_step / 1000.0)
                    issue = {"bound": "warp-instruction issue", "warp_instructions_per_dialogue": ipd, "achieved": ach_issue / 1e9,
                             "peak": peak_issue / 1e9, "unit": "G warp-instr/s", "frac": ach_issue / peak_issue,
                             "source": "instructions per dialogue from the ncu capture in profiles/traffic.json (" + str(cap.get("capture")) + ")"}
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
               "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/f64",
               "data": "synthetic",
               "config": workload_config(n, a.tokens, stats, f"rows sharded x{world}, no collective"),
               "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                            "kernel": "k_main<DT,clean>", "bytes_per_dialogue": bytes_per_row, "peak_source": peak_src,
                            "duration_ms": ms_per_step},
               "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(stats["text_bytes"] + 8 * (n + 1)),
                       "d2h_bytes_per_step": int(40 * n), "ms_per_step": 1000.0 * float(te.item()), "pinned_numa": numa,
                       "h2d_gbs_per_gpu": e2e_gbs, "h2d_ceiling_gbs": h2d_gbs, "frac_of_ceiling": e2e_gbs / h2d_gbs,
                       "ceiling": "bare pinned cudaMemcpyAsync of the same text bytes, all ranks at once, slowest rank"},
               "gpu_launches": int(launches), "clocks": clocks, "issue_roofline": issue, "stages": stages,
               "spark_probe": spark_probe()}
        # ---- CPU baseline on this box's host cores (bounded sample), rank 0 at N = 1 only
        if world == 1:
            rows_cpu = min(n, a.cpu_rows)
            data_np, offs_np = hdata.numpy(), hoffs.numpy()[:rows_cpu + 1]
            v, secs, cpu_out = cpu_port(model, data_np, offs_np, cores)
            agree = bool(np.array_equal(cpu_out[0], hp.numpy()[:rows_cpu]) and np.array_equal(cpu_out[2], hr.numpy()[:rows_cpu]))
            out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                   "sample": f"first {rows_cpu} dialogues of the workload, {secs:.2f} s, oracle/oracle.c with {cores} threads; "
                                             "Spark local[*]: " + ("found on this box but not wired in" if out["spark_probe"]["available"] else "unavailable on this box"),
                                   "labels_and_raw_equal_gpu": agree}
    # ---- the other BASELINE.json configs, same process, same N (C3 and C5 run on every rank)
    del hdata, hoffs, hp, hq, hr, pk, model
    torch.cuda.empty_cache()
    for name in want:
        if name not in ("c1", "c4"):
            run_config(name)
    if out is not None and want:
        out["configs"] = {k: cfgs[k] for k in want if k in cfgs}
    if out is not None:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def stage_times(model, pk, dev, stats, peak):
    """featurize -> CSR and CSR -> score timed separately through the C-ABI with preallocated outputs
    (3 repetitions each after one warm-up), each with its own roofline object (SURVEY.md 8d algorithmic bytes)."""
    import torch
    res = {}
    n = pk.n_rows
    h = model.handle(dev.index)
    P = lambda t: C.c_void_p(t.data_ptr())
    sp = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    rowptr, idx, val = model.featurize(pk, cleaned=False)
    nnz = int(idx.numel())
    cnt = C.c_int64(0)

    def feat():
        h.call("fd_featurize", P(pk.data), P(pk.offs), n, 1, P(rowptr), P(idx), P(val), nnz, C.byref(cnt), sp)

    feat()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        feat()
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / 3
    b = stats["text_bytes"] + 8 * n + 12 * nnz + 8 * n
    res["featurize_csr"] = {"ms": ms, "dialogues_per_s": n / ms * 1e3, "nnz": nnz,
                            "roofline": {"bound": "hbm", "achieved": b / ms / 1e6, "peak": peak, "unit": "GB/s", "frac": b / ms / 1e6 / peak,
                                         "bytes_per_dialogue": b / n, "traffic": None,
                                         "kernels": "k_main<CSR,clean> + k_scan_* + k_gather_csr (whole fd_featurize call, incl. its host read-back of nnz)"}}
    pred = torch.empty(n, dtype=torch.float64, device=dev)
    prob = torch.empty((n, 2), dtype=torch.float64, device=dev)
    raw = torch.empty((n, 2), dtype=torch.float64, device=dev)

    def score():
        h.call("fd_score", P(rowptr), P(idx), P(val), n, P(pred), P(prob), P(raw), sp)

    score()
    torch.cuda.synchronize(dev)
    e0.record()
    for _ in range(3):
        score()
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / 3
    b = 12 * nnz + 48 * n
    res["score_csr"] = {"ms": ms, "dialogues_per_s": n / ms * 1e3,
                        "roofline": {"bound": "hbm", "achieved": b / ms / 1e6, "peak": peak, "unit": "GB/s", "frac": b / ms / 1e6 / peak,
                                     "bytes_per_dialogue": b / n, "traffic": None, "kernels": "k_score_csr"}}
    return res


if __name__ == "__main__":
    main()
