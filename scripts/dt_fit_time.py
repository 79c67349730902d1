"""DecisionTree (featureSubsetStrategy = all) Pipeline.fit time on synthetic dialogues.  Usage: dt_fit_time.py [rows] [depth]"""
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import torch  # noqa: E402
import fdb200 as fd  # noqa: E402
import bench_configs as bc  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
depth = int(sys.argv[2]) if len(sys.argv) > 2 else 5
ctx = bc.Ctx()
torch.cuda.set_device(ctx.local)
tr = importlib.import_module("fraud-detection-spark-kafka-llm_b200.training")
planted = {"ids": list(range(ctx.corpus.n_stop + 300, ctx.corpus.n_stop + 320)), "rate": 0.05}
for n, prof in ((50_000, False), (rows, False), (rows, True)):
    pk, lab = ctx.corpus.sample(n, tokens=200, device=ctx.dev, seed=500, planted=planted)
    tr.PROFILE_DEFAULT = prof
    st = bc.feature_stages(bc.NF) + [fd.IDF(inputCol="raw_features", outputCol="features"),
                                     fd.DecisionTreeClassifier(featuresCol="features", labelCol="labels", maxDepth=depth)]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    model = fd.Pipeline(stages=st).fit(pk, labels=lab)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    stats = getattr(model.stages[-1], "trainer_stats", {})
    print(n, "rows depth", depth, "fit", round(dt, 4), "s nodes", model.stages[-1].trees[0].n_nodes,
          {k: round(v, 4) for k, v in stats.get("phases_s", {}).items()} if prof else "", flush=True)
    del pk, lab, model
    torch.cuda.empty_cache()
