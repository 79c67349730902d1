"""Time the two-step path (fd_featurize -> fd_score) against the fused transform for the forest configs."""
import ctypes as C, importlib, json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fdb200 as fd
sys.path.insert(0, os.path.join(ROOT, "scripts"))
from bench_configs import feature_stages, fitted_idf, timed
synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
dev = torch.device("cuda", 0)
nf = 1 << 18
corpus = synth.SynthCorpus()
top = synth.hash_terms(corpus.words[corpus.n_stop:corpus.n_stop + 2000] + [b""], nf)
rng = np.random.default_rng(1)
n = 1_000_000
pk, _ = corpus.sample(n, tokens=200, device=dev, seed=100)
st = feature_stages(nf)
idf, _ = fitted_idf(st, pk, nf)
for name, cl in (("dt d5", fd.DecisionTreeClassificationModel(synth.random_tree(rng, 5, top, idf.idf), nf)),
                 ("rf 100x10", fd.RandomForestClassificationModel([synth.random_tree(rng, 10, top, idf.idf) for _ in range(100)], nf)),
                 ("gbt 200x6", fd.GBTClassificationModel([synth.random_tree(rng, 6, top, idf.idf, regression=True) for _ in range(200)], [1.0] + [0.1] * 199, nf))):
    model = fd.PipelineModel(st + [idf, cl])
    rowptr, idx, val = model.featurize(pk, cleaned=False)
    fused = timed(lambda: model.transform(pk, cleaned=False), 3, dev)
    score = timed(lambda: model.score(rowptr, idx, val), 3, dev)
    a, b = model.transform(pk, cleaned=False), model.score(rowptr, idx, val)
    print(json.dumps({"model": name, "rows": n, "fused_transform_ms": fused, "score_csr_ms": score,
                      "equal": bool(torch.equal(a["rawPrediction"], b["rawPrediction"]))}))
