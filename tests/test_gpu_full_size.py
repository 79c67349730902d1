"""BASELINE.json's full size (config C2: 1,000,000 dialogues x 200 tokens, HashingTF 2^18 + IDF + DecisionTree depth 5) on
the GPU: every row against the threaded C port of the oracle, plus properties that need no oracle at all --
row-permutation equivariance, cleaning idempotence (raw text through the cleaning kernel == pre-cleaned text through
the Tokenizer-only kernel), conservation of token mass (the counts of a dialogue sum to its surviving tokens), CSR
well-formedness, and featurize -> score == fused transform."""
import importlib

import numpy as np
import pytest
import torch

from helpers import oracle_for

pytestmark = pytest.mark.gpu
N_ROWS = 1_000_000


@pytest.fixture(scope="module")
def c2(fd):
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    nf = 1 << 18
    corpus = synth.SynthCorpus()
    pk, _ = corpus.sample(N_ROWS, tokens=200, device="cuda:0", seed=20261106)
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
              fd.HashingTF(inputCol="filtered_words", outputCol="raw_features", numFeatures=nf)]
    # IDF fitted on a 65,536-dialogue sample (document frequencies counted on the GPU), as bench.py does
    ns = 1 << 16
    _, idx, _ = fd.PipelineModel(stages).featurize(fd.PackedText(pk.data, pk.offs[:ns + 1]), cleaned=False)
    df = torch.bincount(idx.long(), minlength=nf).cpu().numpy()
    idf = fd.IDF(inputCol="raw_features", outputCol="features").fit_from_doc_freq(df, ns)
    top = synth.hash_terms(corpus.words[corpus.n_stop:corpus.n_stop + 2000] + [b""], nf)  # the 2,000 most frequent buckets, as bench.py
    tree = synth.random_tree(np.random.default_rng(20261106), 5, top, idf.idf, weights=df[top])  # drawn by document frequency: both branches taken
    model = fd.PipelineModel(stages + [idf, fd.DecisionTreeClassificationModel(tree, nf)])
    return model, pk


def test_c2_every_row_against_the_c_oracle(fd, c2):
    from oracle import c_oracle
    model, pk = c2
    res = model.transform(pk, cleaned=False)
    o = oracle_for(model)
    co = c_oracle.COracle(model._plan.stop.getOrDefault("stopWords"), False, o.num_features, o.binary, o.idf, o.classifier)
    pred, prob, raw = co.transform(pk.data.cpu().numpy(), pk.offs.cpu().numpy(), True)
    assert np.array_equal(res["prediction"].cpu().numpy(), pred)
    assert np.array_equal(res["rawPrediction"].cpu().numpy(), raw)
    np.testing.assert_allclose(res["probability"].cpu().numpy(), prob, rtol=1e-9, atol=1e-300)
    assert len(np.unique(raw, axis=0)) >= 8 and 0.001 < pred.mean() < 0.999, "degenerate workload: the rows reach too few leaves"


def test_c2_csr_properties_and_two_step_path(fd, c2):
    model, pk = c2
    rowptr, idx, val = model.featurize(pk, cleaned=False)
    n = pk.n_rows
    assert rowptr.numel() == n + 1 and int(rowptr[0]) == 0 and int(rowptr[-1]) == idx.numel() == val.numel()
    nnz = rowptr[1:] - rowptr[:-1]
    assert int(nnz.min()) >= 1 and int(nnz.max()) <= 200
    # indices strictly ascending inside every row: steps down happen only where a new row starts
    step_down = (idx[1:] <= idx[:-1]).nonzero().flatten() + 1
    is_row_start = torch.zeros(idx.numel() + 1, dtype=torch.bool, device=idx.device)
    is_row_start[rowptr] = True
    assert bool(is_row_start[step_down].all())
    assert int(idx.min()) >= 0 and int(idx.max()) < (1 << 18)
    # token mass: counts (= val / idf) of a dialogue add up to its surviving tokens, which are at most its 200 tokens
    idf = torch.from_numpy(model._plan.idf.idf).to(val.device)
    w = idf[idx.long()]
    cnt = torch.where(w != 0, val / w, torch.zeros_like(val))
    assert torch.all((cnt - cnt.round()).abs() < 1e-9)
    rows = torch.repeat_interleave(torch.arange(n, device=val.device), nnz)
    mass = torch.zeros(n, dtype=torch.float64, device=val.device).index_add_(0, rows, cnt.round())
    total = int(pk.offs[-1])
    cs = torch.zeros(total + 1, dtype=torch.int64, device=val.device)
    torch.cumsum((pk.data[:total] == 0x20).to(torch.int64), 0, out=cs[1:])
    tokens = cs[pk.offs[1:]] - cs[pk.offs[:-1]] + 1   # Java split: separators + 1
    assert bool((mass <= tokens.to(torch.float64)).all()) and float(mass.mean()) > 60.0
    # CSR -> score equals the fused kernel, bit for bit
    fused = model.transform(pk, cleaned=False)
    two = model.score(rowptr, idx, val)
    for k in ("prediction", "probability", "rawPrediction"):
        assert torch.equal(fused[k], two[k])


def test_c2_row_permutation_and_cleaning_idempotence(fd, c2):
    model, pk = c2
    n = 200_000  # a slice keeps the gathers small; the kernels see the same row mix
    offs = pk.offs[: n + 1].clone()
    lens = offs[1:] - offs[:-1]
    perm = torch.randperm(n, device=offs.device, generator=torch.Generator(device=offs.device).manual_seed(7))
    plen = lens[perm]
    poffs = torch.zeros(n + 1, dtype=torch.int64, device=offs.device)
    torch.cumsum(plen, 0, out=poffs[1:])
    total = int(poffs[-1])
    # byte gather: destination byte j of permuted row r comes from source row perm[r]
    row_of = torch.repeat_interleave(torch.arange(n, device=offs.device), plen)
    src = offs[:-1][perm][row_of] + (torch.arange(total, device=offs.device) - poffs[:-1][row_of])
    pdata = torch.cat([pk.data[src], torch.zeros(32, dtype=torch.uint8, device=offs.device)])
    base = fd.PackedText(pk.data, offs)
    shuf = fd.PackedText(pdata, poffs)
    a = model.transform(base, cleaned=False)
    b = model.transform(shuf, cleaned=False)
    for k in ("prediction", "probability", "rawPrediction"):
        assert torch.equal(a[k][perm], b[k])
    # the synthetic text is already lower-case letters and single/double spaces: cleaning must be the identity on it
    ra, ia, va = model.featurize(base, cleaned=False)
    rb, ib, vb = model.featurize(base, cleaned=True)
    assert torch.equal(ra, rb) and torch.equal(ia, ib) and torch.equal(va, vb)
