"""The C port of the oracle (oracle/oracle.c, bench.py's cpu_baseline) agrees with the pinned Python oracle."""
import importlib

import numpy as np

from helpers import messy_texts, oracle_for
from oracle import c_oracle


def _c_for(model):
    o = oracle_for(model)
    p = model._plan
    sw = p.stop.getOrDefault("stopWords") if p.stop is not None else []
    return o, c_oracle.COracle(sw, False, o.num_features, o.binary, o.idf, o.classifier)


def _pack(fd, rows):
    return fd.pack_strings(rows)


def test_c_oracle_matches_python_on_shipped_model(fd, shipped_model_dir, golden):
    model = fd.PipelineModel.load(shipped_model_dir)
    py, c = _c_for(model)
    rows = [r["text"].encode() for r in golden["provisional_oracle_outputs"]] + [t.encode() for t in messy_texts(300, seed=1)]
    data, offs = _pack(fd, rows)
    for clean in (True, False):
        src = rows if clean else [r.decode().lower().encode() for r in rows]
        d, o = _pack(fd, src)
        rp, idx, val = c.featurize(d, o, do_clean=clean, n_threads=3)
        for i, r in enumerate(src):
            eix, evl, _, _ = py.featurize(r, do_clean=clean)
            assert idx[rp[i]:rp[i + 1]].tolist() == eix and val[rp[i]:rp[i + 1]].tolist() == evl
    pred, prob, raw = c.transform(data, offs, n_threads=2)
    e = py.transform(rows)
    assert np.array_equal(pred, e[0]) and np.array_equal(raw, e[2]) and np.array_equal(prob, e[1])


def test_c_oracle_matches_python_on_tree_models(fd):
    tg = importlib.import_module("test_gpu_parity")
    for kind, nt, depth in (("dt", 1, 5), ("rf", 20, 6), ("gbt", 30, 4)):
        model, corpus = tg._tree_model(fd, kind, nt, depth)
        py, c = _c_for(model)
        pk, _ = corpus.sample(150, tokens=80, device="cpu", seed=2)
        pred, prob, raw = c.transform(pk.data.numpy(), pk.offs.numpy())
        d, o = pk.data.numpy(), pk.offs.numpy()
        e = py.transform([bytes(d[o[i]:o[i + 1]]) for i in range(150)])
        assert np.array_equal(pred, e[0]) and np.array_equal(raw, e[2]) and np.array_equal(prob, e[1])
