"""GPU parity: the CUDA path, called through the C-ABI, against the CPU oracle on the same inputs.

Bar (BASELINE.json north_star): hashed feature indices, term counts and predicted labels bit-exact;
TF-IDF values bit-exact (one fp64 multiply); margins / votes bit-exact (same fp64 operation order);
class probabilities within 1e-9 relative (they pass through exp(), whose last bit may differ between
libm and CUDA)."""
import importlib

import numpy as np
import pytest

from helpers import csr_rows, messy_texts, oracle_for

pytestmark = pytest.mark.gpu
PROB_RTOL = 1e-9


def _check_scores(res, exp, exact_raw=True):
    pred, prob, raw = (res[k].cpu().numpy() if hasattr(res[k], "cpu") else res[k] for k in ("prediction", "probability", "rawPrediction"))
    epred, eprob, eraw = exp
    assert np.array_equal(pred, epred)
    if exact_raw:
        assert np.array_equal(raw, eraw)
    else:
        np.testing.assert_allclose(raw, eraw, rtol=PROB_RTOL, atol=0)
    np.testing.assert_allclose(prob, eprob, rtol=PROB_RTOL, atol=1e-300)


def test_shipped_lr_model_golden_texts(fd, shipped_model_dir, golden):
    model = fd.PipelineModel.load(shipped_model_dir)
    recs = golden["provisional_oracle_outputs"]
    texts = [r["text"] for r in recs]
    res = model.transform(texts)
    raw = res["rawPrediction"].cpu().numpy()
    for i, r in enumerate(recs):
        assert raw[i, 1] == r["margin"], r["text"][:40]
        assert res["prediction"][i].item() == r["prediction"]
    _check_scores(res, oracle_for(model).transform([t.encode() for t in texts]))
    # the agent-shaped call of utils/agent_api.py:155-175
    long = [r for r in recs if r["n_tokens"] == 375][0]
    out = fd.predict_and_get_label(model, long["text"])
    assert out["prediction"] == 1.0 and abs(out["confidence"] - long["probability"][1]) <= 1e-9 * long["probability"][1]
    rowptr, idx, val = model.featurize(texts)
    for (ix, vl), r in zip(csr_rows(rowptr, idx, val), recs):
        assert ix == r["idx"] and vl == r["val"]


@pytest.mark.parametrize("num_features,binary", [(10000, False), (262144, False), (1000003, True), (1, False)])
def test_featurize_messy_text_bit_exact(fd, num_features, binary):
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered"),
              fd.HashingTF(inputCol="filtered", outputCol="raw_features", numFeatures=num_features, binary=binary)]
    rng = np.random.default_rng(5)
    idf = rng.random(num_features) * 3.0
    idf[rng.integers(0, num_features, size=max(1, num_features // 10))] = 0.0
    stages.append(fd.IDFModel(idf, inputCol="raw_features", outputCol="features"))
    model = fd.PipelineModel(stages)
    texts = messy_texts(700, seed=num_features % 97)
    ora = oracle_for(model)
    for cleaned in (False, True):
        rowptr, idx, val = model.featurize(texts, cleaned=cleaned)
        rows = csr_rows(rowptr, idx, val)
        assert len(rows) == len(texts)
        for t, (ix, vl) in zip(texts, rows):
            src = t.lower().encode() if cleaned else t.encode()  # Tokenizer-only mode: non-ASCII case folding is the host's job
            eix, evl, _, _ = ora.featurize(src, do_clean=not cleaned)
            assert ix == eix, repr(t)
            assert vl == evl, repr(t)


@pytest.mark.parametrize("case_sensitive", [False, True])
def test_custom_stop_word_lists(fd, case_sensitive):
    """Stop words longer than 12 bytes (byte-compared from the pool), an empty-string stop word, mixed case and a
    few hundred entries (cuckoo placement) -- against the oracle."""
    rng = np.random.default_rng(3)
    many = ["w" + "".join(chr(97 + int(c)) for c in rng.integers(0, 26, size=int(rng.integers(1, 16)))) for _ in range(600)]
    words = ["the", "Hello", "antidisestablishmentarianism", "internationalisation", "", "social", "xx", "q"] + many
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"),
              fd.StopWordsRemover(inputCol="words", outputCol="filtered", stopWords=words, caseSensitive=case_sensitive),
              fd.HashingTF(inputCol="filtered", outputCol="features", numFeatures=4096)]
    model = fd.PipelineModel(stages)
    texts = messy_texts(400, seed=12) + [" ".join(many[:300]) + " keep me", "a  b   antidisestablishmentarianism internationalisations"]
    ora = oracle_for(model)
    rowptr, idx, val = model.featurize(texts)
    for t, (ix, vl) in zip(texts, csr_rows(rowptr, idx, val)):
        eix, evl, _, _ = ora.featurize(t.encode())
        assert ix == eix and vl == evl, repr(t[:50])


def _long_rows():
    rng = np.random.default_rng(11)
    words = [bytes(rng.integers(97, 123, size=int(rng.integers(1, 9)), dtype=np.uint8)).decode() for _ in range(5000)]
    rows = []
    rows.append(" ".join(words[:3000]))                        # > 384 distinct buckets: wide-table kernel
    rows.append(" ".join(words[i % 40] for i in range(4000)))  # many windows, few buckets
    rows.append("x" * 5000 + " tail the end")                  # one token spanning several windows
    rows.append("1" * 3000 + "ab" + "2" * 2000 + " cd")        # multi-window token that cleans down to "ab"
    rows.append("7" * 2500 + " a b")                           # multi-window EMPTY token followed by letters
    rows.append("a " + "9" * 2500)                             # trailing multi-window empty token (dropped)
    rows.append(" " * 3000)                                    # separators only -> no tokens
    rows.append(" " * 1500 + "z" + " " * 1500)                 # interior empties kept, trailing dropped
    rows.append("the " * 1200)                                 # stop words only
    rows.append("ok")
    rows.append("word " * 205)                                 # token exactly at a window edge region
    for k in (1008, 1009, 1023, 1024, 1025, 1040, 2047, 2048, 2049):
        rows.append(("ab " * 400)[:k])
        rows.append("q" * k)
        rows.append("q" * (k - 1) + " r")
    return rows


def test_long_rows_windows_wide_table_and_giant_tokens(fd, shipped_model_dir):
    model = fd.PipelineModel.load(shipped_model_dir)
    rows = _long_rows()
    ora = oracle_for(model)
    rowptr, idx, val = model.featurize(rows)
    for t, (ix, vl) in zip(rows, csr_rows(rowptr, idx, val)):
        eix, evl, _, _ = ora.featurize(t.encode())
        assert ix == eix and vl == evl, repr(t[:30]) + f" len={len(t)}"
    _check_scores(model.transform(rows), ora.transform([t.encode() for t in rows]))


def test_last_window_extension_every_length(fd, shipped_model_dir):
    """A dialogue's last window may run 64 bytes past the 1 KiB window (handled one word per lane): every byte
    length from just below one window to just past the extension, and around the second window's end, with
    messy content (punctuation, case, U+0130 / U+212A, runs of separators) and every 16-byte alignment."""
    model = fd.PipelineModel.load(shipped_model_dir)
    ora = oracle_for(model)
    rng = np.random.default_rng(23)
    pool = " ".join(messy_texts(400, seed=31, max_tokens=30))
    rows = []
    for k in list(range(985, 1125)) + list(range(2030, 2150)):
        o = int(rng.integers(0, len(pool) - k - 1))
        rows.append(pool[o:o + k])
        rows.append(("x" * int(rng.integers(1, 40)) + " ") * 60)   # spacer of varying size: shifts the next row's alignment
        rows[-1] = rows[-1][:int(rng.integers(1, 900))]
    for k in range(1000, 1100, 3):  # U+0130 / U+212A straddling the window end, spaces and letters in the extension
        rows.append("a" * k + "İ K b" + " " * (k % 7) + "c")
        rows.append("Ab " * (k // 3) + "İİKK" + "  zz" * (k % 5))
    for cleaned in (False, True):
        rowptr, idx, val = model.featurize(rows, cleaned=cleaned)
        for t, (ix, vl) in zip(rows, csr_rows(rowptr, idx, val)):
            src = t.lower().encode() if cleaned else t.encode()
            eix, evl, _, _ = ora.featurize(src, do_clean=not cleaned)
            assert ix == eix and vl == evl, f"cleaned={cleaned} len={len(t.encode())} " + repr(t[:20])
    _check_scores(model.transform(rows), ora.transform([t.encode() for t in rows]))


def _tree_model(fd, kind, n_trees, depth, num_features=262144, seed=5):
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    rng = np.random.default_rng(seed)
    corpus = synth.SynthCorpus(n_words=3000, seed=77)
    feats = synth.hash_terms(corpus.words[corpus.n_stop:corpus.n_stop + 400] + [b""], num_features)
    idf = np.log(1001.0 / (rng.integers(0, 1000, size=num_features) + 1.0))
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
              fd.HashingTF(inputCol="filtered_words", outputCol="raw_features", numFeatures=num_features),
              fd.IDFModel(idf, inputCol="raw_features", outputCol="features")]
    if kind == "dt":
        cl = fd.DecisionTreeClassificationModel(synth.random_tree(rng, depth, feats, idf), num_features)
    elif kind == "rf":
        cl = fd.RandomForestClassificationModel([synth.random_tree(rng, depth, feats, idf, full=False) for _ in range(n_trees)], num_features)
    else:
        trees = [synth.random_tree(rng, depth, feats, idf, regression=True, full=False) for _ in range(n_trees)]
        cl = fd.GBTClassificationModel(trees, [1.0] + [0.1] * (n_trees - 1), num_features)
    return fd.PipelineModel(stages + [cl]), corpus


@pytest.mark.parametrize("kind,n_trees,depth", [("dt", 1, 5), ("dt", 1, 9), ("rf", 100, 10), ("rf", 33, 4), ("gbt", 200, 6), ("gbt", 7, 3)])
def test_tree_models_fused_and_csr_paths(fd, kind, n_trees, depth, tmp_path):
    model, corpus = _tree_model(fd, kind, n_trees, depth)
    pk, _ = corpus.sample(300, tokens=120, device="cpu", seed=9)
    data, offs = pk.data.numpy(), pk.offs.numpy()
    rows = [bytes(data[offs[i]:offs[i + 1]]) for i in range(300)] + [t.encode() for t in messy_texts(100, seed=4)]
    exp = oracle_for(model).transform(rows)
    _check_scores(model.transform(rows), exp)
    rowptr, idx, val = model.featurize(rows)
    _check_scores(model.score(rowptr, idx, val), exp)
    # Spark-format round trip of the tree stage (writer + loader) leaves the scores unchanged
    model.save(str(tmp_path / "m"))
    again = fd.PipelineModel.load(str(tmp_path / "m"))
    _check_scores(again.transform(rows), exp)
    assert len(np.unique(exp[2][:, 1])) > 3, "degenerate test: every row reached the same leaves"
    if (kind, n_trees) in (("dt", 1), ("rf", 100), ("gbt", 200)):
        assert not np.all(exp[0] == exp[0][0]), "degenerate test: all rows got one label"


@pytest.mark.parametrize("featuriser", ["hashing", "hashing-binary", "vocabulary"])
@pytest.mark.parametrize("kind,n_trees,depth", [("dt", 1, 5), ("dt", 1, 8), ("rf", 20, 6)])
def test_trees_on_long_and_messy_rows(fd, kind, n_trees, depth, featuriser):
    """Tree models over the window-edge rows (many windows, multi-window tokens, hundreds of distinct buckets) and
    messy text, with HashingTF (counts and binary) and with a CountVectorizerModel in front.  A depth-5 tree takes
    the split-feature-counter path of the fused kernel (no dialogue is ever deferred there); depth 8 and the forest
    take the full count table, where the first long row overflows into the wide-table kernel."""
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    rng = np.random.default_rng(29)
    rows = _long_rows() + messy_texts(300, seed=12, max_tokens=120)
    import collections
    df = collections.Counter(w for r in rows for w in set("".join(c if c.isalpha() or c == " " else "" for c in r.lower()).split(" ")) if w.isascii())
    stop = set(fd.load_default_stop_words())
    pick = [w for w, _ in df.most_common() if 0 < len(w) <= 14 and w not in stop][:300]   # most frequent first
    if featuriser == "vocabulary":
        vocab = [""] + pick
        nf = len(vocab)
        front = fd.CountVectorizerModel(vocab, inputCol="filtered_words", outputCol="raw_features")
        feats = np.arange(41)
    else:
        nf = 1 << 16
        front = fd.HashingTF(inputCol="filtered_words", outputCol="raw_features", numFeatures=nf, binary=featuriser.endswith("binary"))
        feats = synth.hash_terms([w.encode() for w in pick[:40]] + [b""], nf)   # trees split on frequent words: both branches taken
    idf = np.log(301.0 / (rng.integers(0, 300, size=nf) + 1.0))
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
              front, fd.IDFModel(idf, inputCol="raw_features", outputCol="features")]
    if kind == "dt":
        cl = fd.DecisionTreeClassificationModel(synth.random_tree(rng, depth, feats, idf), nf)
    else:
        cl = fd.RandomForestClassificationModel([synth.random_tree(rng, depth, feats, idf, full=False) for _ in range(n_trees)], nf)
    model = fd.PipelineModel(stages + [cl])
    exp = oracle_for(model).transform([t.encode() for t in rows])
    _check_scores(model.transform(rows), exp)
    rowptr, idx, val = model.featurize(rows)
    _check_scores(model.score(rowptr, idx, val), exp)
    assert len(np.unique(exp[2][:, 1])) > 1, "degenerate test: every row reached the same leaf"


def _forest_with_idf(fd, kind, idf, depths, num_features, feats, seed):
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    rng = np.random.default_rng(seed)
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
              fd.HashingTF(inputCol="filtered_words", outputCol="raw_features", numFeatures=num_features),
              fd.IDFModel(idf, inputCol="raw_features", outputCol="features")]
    trees = [synth.random_tree(rng, d, feats, idf, regression=(kind == "gbt"), full=bool(i % 2)) for i, d in enumerate(depths)]
    if kind == "rf":
        cl = fd.RandomForestClassificationModel(trees, num_features)
    else:
        cl = fd.GBTClassificationModel(trees, [1.0] + [0.3] * (len(trees) - 1), num_features)
    return fd.PipelineModel(stages + [cl])


@pytest.mark.parametrize("kind", ["rf", "gbt"])
@pytest.mark.parametrize("weights", ["plain", "zeros", "negative"])
def test_tree_walk_corner_cases(fd, kind, weights):
    """Forests of every small shape (a single leaf, depths 1-4, odd and even, ragged) so that each arrangement of the
    two-level node records occurs (leaf child, leaf grandchild on either side); idf vectors with zero entries
    (threshold compare is 0 <= thr whatever the count) and with negative entries (the integer walk is switched off
    and the kernels compare tf-idf values in fp64)."""
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    num_features = 4096
    corpus = synth.SynthCorpus(n_words=600, seed=5)
    feats = synth.hash_terms(corpus.words[corpus.n_stop:corpus.n_stop + 150] + [b""], num_features)
    rng = np.random.default_rng(17)
    idf = np.log(501.0 / (rng.integers(0, 500, size=num_features) + 1.0))
    if weights == "zeros":
        idf[feats[::3]] = 0.0
    elif weights == "negative":
        idf[feats[::4]] *= -1.0
    depths = [0, 1, 2, 3, 4, 1, 2, 3, 4, 5, 0, 2, 4, 3, 1, 6]
    model = _forest_with_idf(fd, kind, idf, depths, num_features, feats, seed=3)
    pk, _ = corpus.sample(400, tokens=60, device="cpu", seed=2)
    data, offs = pk.data.numpy(), pk.offs.numpy()
    rows = [bytes(data[offs[i]:offs[i + 1]]) for i in range(400)] + ["", " ", "zzz"]
    exp = oracle_for(model).transform([r if isinstance(r, bytes) else r.encode() for r in rows])
    _check_scores(model.transform(rows), exp)
    rowptr, idx, val = model.featurize(rows)
    _check_scores(model.score(rowptr, idx, val), exp)
    assert len(np.unique(exp[2][:, 1])) > 3


@pytest.mark.parametrize("kind", ["dt", "rf", "gbt", "lr"])
@pytest.mark.parametrize("case_sensitive", [False, True])
def test_classifier_features_that_stop_words_map_to(fd, kind, case_sensitive):
    """The fused kernels skip the stop-word lookup for terms the classifier cannot see and probe it only for buckets
    that both the classifier sees and a stop word maps to.  Here the classifier splits on (or weighs) exactly such
    buckets: the buckets of 'the', 'a', 'of', 'to' next to those of ordinary words, with a stop-word list that also
    holds a word longer than 12 bytes.  A removed stop word must not be counted; with a case-sensitive list of
    capitalised words nothing matches the lower-cased tokens and everything is counted."""
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    nf = 1 << 12
    rng = np.random.default_rng(41)
    stop = ["the", "a", "of", "to", "and", "is", "i", "antidisestablishmentarianism"]
    if case_sensitive:
        stop = [w.capitalize() for w in stop] + ["hello"]
    words = [b"the", b"a", b"of", b"to", b"hello", b"social", b"number", b"verify", b"antidisestablishmentarianism", b"q", b""]
    feats = synth.hash_terms(words, nf)
    idf = np.log(401.0 / (rng.integers(0, 400, size=nf) + 1.0))
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"),
              fd.StopWordsRemover(inputCol="words", outputCol="filtered_words", stopWords=stop, caseSensitive=case_sensitive),
              fd.HashingTF(inputCol="filtered_words", outputCol="raw_features", numFeatures=nf),
              fd.IDFModel(idf, inputCol="raw_features", outputCol="features")]
    if kind == "dt":
        cl = fd.DecisionTreeClassificationModel(synth.random_tree(rng, 5, feats, idf), nf)
    elif kind == "rf":
        cl = fd.RandomForestClassificationModel([synth.random_tree(rng, 4, feats, idf, full=False) for _ in range(12)], nf)
    elif kind == "gbt":
        trees = [synth.random_tree(rng, 3, feats, idf, regression=True, full=False) for _ in range(9)]
        cl = fd.GBTClassificationModel(trees, [1.0] + [0.1] * 8, nf)
    else:
        coef = np.zeros(nf)
        coef[feats] = rng.normal(0.0, 1.0, size=len(feats))
        cl = fd.LogisticRegressionModel(coef, -0.3)
    model = fd.PipelineModel(stages + [cl])
    rows = messy_texts(600, seed=21, max_tokens=90) + _long_rows()[:6]
    exp = oracle_for(model).transform([t.encode() for t in rows])
    _check_scores(model.transform(rows), exp)
    rowptr, idx, val = model.featurize(rows)
    _check_scores(model.score(rowptr, idx, val), exp)
    assert len(np.unique(exp[2][:, 1])) > 3


@pytest.mark.parametrize("kind,n_trees,depth", [("rf", 40, 7), ("gbt", 24, 5), ("dt", 1, 6)])
def test_forests_in_the_large_batch_shape(fd, kind, n_trees, depth):
    """Batches of at least 65,536 short dialogues run the 256-slot-table shape of the kernel (16 warps per CTA); the
    other tree tests are small batches and run the 512-slot shape.  Checked against the C port of the oracle."""
    from oracle import c_oracle
    model, corpus = _tree_model(fd, kind, n_trees, depth)
    pk, _ = corpus.sample(66000, tokens=40, device="cpu", seed=13)
    res = model.transform(pk, cleaned=False)
    o = oracle_for(model)
    co = c_oracle.COracle(model._plan.stop.getOrDefault("stopWords"), False, o.num_features, o.binary, o.idf, o.classifier)
    _check_scores(res, co.transform(pk.data.numpy(), pk.offs.numpy(), True))


@pytest.mark.parametrize("num_features", [4096, 10000])
def test_sparse_lr_coefficients_bit_exact(fd, num_features):
    """LogisticRegression whose coefficient vector is mostly exact zeros (+0.0 and -0.0): the fused kernel drops those
    terms before counting them, which must not move a single bit of the margin -- also with zero idf entries (products
    of -0.0), rows whose terms all have zero coefficients (margin = intercept) and empty rows.  A dense vector (no
    filter) and the CSR scoring path are checked alongside."""
    rng = np.random.default_rng(num_features)
    idf = rng.random(num_features) * 3.0
    idf[rng.integers(0, num_features, size=num_features // 8)] = 0.0
    texts = messy_texts(600, seed=77, max_tokens=90) + ["", "   ", "zzzz qqqq", "the of and"]
    for density in (0.05, 0.4, 1.0):
        coef = rng.normal(0, 1, num_features)
        zero = rng.random(num_features) >= density
        coef[zero] = 0.0
        coef[zero & (rng.random(num_features) < 0.5)] = -0.0
        stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered"),
                  fd.HashingTF(inputCol="filtered", outputCol="raw_features", numFeatures=num_features),
                  fd.IDFModel(idf, inputCol="raw_features", outputCol="features"),
                  fd.LogisticRegressionModel(coef, 0.25, featuresCol="features")]
        model = fd.PipelineModel(stages)
        exp = oracle_for(model).transform([t.encode() for t in texts])
        _check_scores(model.transform(texts), exp)
        rowptr, idx, val = model.featurize(texts)
        _check_scores(model.score(rowptr, idx, val), exp)
        if density < 1.0:
            assert (exp[2][:, 1] == 0.25).sum() >= 3   # rows with no surviving non-zero term sit exactly at the intercept


def test_host_buffer_transform_matches_device(fd):
    model, corpus = _tree_model(fd, "dt", 1, 5)
    pk, _ = corpus.sample(20000, tokens=50, device="cpu", seed=21)
    dev = model.transform(pk, cleaned=False)
    host = model.transform(pk, cleaned=False, output="host")
    for k in ("prediction", "probability", "rawPrediction"):
        assert np.array_equal(dev[k].cpu().numpy(), host[k])
    assert model.handle().launch_count() > 0


def test_pinned_host_buffers_are_read_in_place(fd, shipped_model_dir):
    """A micro-batch whose buffers are pinned host memory is read and written by the kernels in place (no staging copies):
    same columns as the copying path (pageable buffers) and as the oracle -- long rows, deferred rows, empty rows, a text
    buffer that ends exactly at its last (unaligned) byte, and a batch that starts in the middle of the buffer."""
    import torch
    from helpers import messy_texts
    model = fd.PipelineModel.load(shipped_model_dir)
    ora = oracle_for(model)
    rows = _long_rows() + ["", " ", "a"] + messy_texts(60, seed=9) + ["", "trailing token"]
    pk = fd.PackedText.from_strings(rows)
    n = pk.n_rows
    total = int(pk.offs[-1])
    assert total % 16 != 0
    exp = ora.transform([t.encode() for t in rows])
    h = model.handle()
    P = importlib.import_module("fraud-detection-spark-kafka-llm_b200._lib").ptr
    data = torch.from_numpy(np.ascontiguousarray(pk.data[:total])).pin_memory()   # ends at the last byte of the text
    offs = torch.from_numpy(pk.offs.copy()).pin_memory()
    outs = [torch.full(s_, float("nan"), dtype=torch.float64).pin_memory() for s_ in ((n,), (n, 2), (n, 2))]
    h.call("fd_transform_host", P(data), P(offs), n, 1, P(outs[0]), P(outs[1]), P(outs[2]))
    got = dict(zip(("prediction", "probability", "rawPrediction"), (o.numpy() for o in outs)))
    _check_scores(got, exp)
    # pageable buffers: the copying path
    pouts = [np.full(s_, np.nan) for s_ in ((n,), (n, 2), (n, 2))]
    h.call("fd_transform_host", P(pk.data), P(pk.offs), n, 1, P(pouts[0]), P(pouts[1]), P(pouts[2]))
    for a_, b_ in zip(outs, pouts):
        assert np.array_equal(a_.numpy(), b_)
    # rows [k, n) of the same buffers: offsets that do not start at zero, outputs written from their first element
    k = 7
    sub = [torch.full(s_, float("nan"), dtype=torch.float64).pin_memory() for s_ in ((n - k,), (n - k, 2), (n - k, 2))]
    h.call("fd_transform_host", P(data), P(offs[k:]), n - k, 1, P(sub[0]), P(sub[1]), P(sub[2]))
    for a_, b_ in zip(outs, sub):
        assert np.array_equal(a_.numpy()[k:], b_.numpy())


def test_sharded_transform_equals_whole(fd, shipped_model_dir):
    """Byte-balanced shards (one per rank in a multi-GPU job; no collective) give the same rows as one call."""
    model = fd.PipelineModel.load(shipped_model_dir)
    texts = messy_texts(900, seed=41, max_tokens=200)
    whole = model.transform(texts)
    pk = fd.PackedText.from_strings(texts)
    parts = [model.transform(pk.shard(r, 3)) for r in range(3)]
    for col in ("prediction", "probability", "rawPrediction"):
        assert np.array_equal(np.concatenate([p[col].cpu().numpy() for p in parts]), whole[col].cpu().numpy())


def test_empty_batch_and_errors(fd, shipped_model_dir):
    model = fd.PipelineModel.load(shipped_model_dir)
    assert model.transform([]).count() == 0
    rowptr, idx, val = model.featurize([])
    assert rowptr.tolist() == [0] and idx.numel() == 0
    res = model.transform([""]).first()
    assert res[0] == 0.0
    h = fd.Handle(0)
    with pytest.raises(fd.FdError):
        h.call("fd_transform", None, None, 1, 1, None, None, None, None)


def test_random_byte_strings_property(fd, shipped_model_dir):
    """Property test (hypothesis): arbitrary valid-UTF-8 dialogues -- control characters, every kind of white space,
    astral code points, lone combining marks -- give the oracle's indices, counts and margins bit for bit, in both
    the cleaning mode and the Tokenizer-only mode."""
    from hypothesis import HealthCheck, given, settings, strategies as st

    model = fd.PipelineModel.load(shipped_model_dir)
    ora = oracle_for(model)
    alphabet = st.one_of(st.sampled_from(list(" \t\n\x0b\x0c\r .,'!-aAbBzZiIkK09İKſß")), st.characters(blacklist_categories=("Cs",)))
    texts = st.lists(st.text(alphabet=alphabet, max_size=300), min_size=1, max_size=40)

    @settings(max_examples=25, deadline=None, suppress_health_check=list(HealthCheck))
    @given(texts)
    def run(batch):
        rowptr, idx, val = model.featurize(batch)
        for t, (ix, vl) in zip(batch, csr_rows(rowptr, idx, val)):
            eix, evl, _, _ = ora.featurize(t.encode("utf-8"))
            assert ix == eix and vl == evl, repr(t)
        exp = ora.transform([t.encode("utf-8") for t in batch])
        _check_scores(model.transform(batch), exp)
        ascii_only = [t for t in batch if t.isascii()]
        if ascii_only:  # Tokenizer-only mode (the kernel folds ASCII case; the host folds the rest)
            rowptr, idx, val = model.featurize(ascii_only, cleaned=True)
            for t, (ix, vl) in zip(ascii_only, csr_rows(rowptr, idx, val)):
                eix, evl, _, _ = ora.featurize(t.encode(), do_clean=False)
                assert ix == eix and vl == evl, repr(t)

    run()


def test_micro_batcher_and_batch_classification(fd, shipped_model_dir, golden):
    """Kafka-shaped loop (app_ui.py:195-226) and CSV batch loop (app_ui.py:144-150) as ONE transform per batch."""
    import json

    class Msg:
        def __init__(self, key, value, err=None):
            self._k, self._v, self._e = key, value, err

        def key(self):
            return self._k

        def value(self):
            return self._v

        def error(self):
            return self._e

    class Consumer:
        def __init__(self, msgs):
            self.q = list(msgs)

        def poll(self, timeout):
            return self.q.pop(0) if self.q else None

    class Producer:
        def __init__(self):
            self.sent, self.flushes = [], 0

        def produce(self, topic, key, value):
            self.sent.append((topic, key, json.loads(value)))

        def flush(self):
            self.flushes += 1

    model = fd.PipelineModel.load(shipped_model_dir)
    recs = golden["provisional_oracle_outputs"]
    msgs = [Msg(str(i).encode(), json.dumps({"text": r["text"]}).encode()) for i, r in enumerate(recs)]
    msgs.insert(3, Msg(b"bad", b"not json"))
    errors = []
    prod = Producer()
    mb = fd.MicroBatcher(model, Consumer(msgs), prod, "out-topic", max_batch=8, max_wait_s=1.0, on_error=errors.append)
    out = []
    while True:
        got = mb.step()
        if not got:
            break
        out += got
    assert len(out) == len(recs) and len(errors) == 1 and mb.batches == 2 and prod.flushes == 2
    for o, r, sent in zip(out, recs, prod.sent):
        assert o["prediction"] == r["prediction"] and abs(o["confidence"] - r["probability"][1]) <= 1e-9 * r["probability"][1]
        assert sent[0] == "out-topic" and sent[2]["original_text"] == r["text"] and sent[2]["prediction"] == r["prediction"]
    rows = fd.classify_batch(model, [r["text"] for r in recs])
    assert [x["prediction"] for x in rows] == [r["prediction"] for r in recs]
    assert rows[5]["predicted_label"] == "Potentially Fraudulent"


def test_many_deferred_dialogues_at_scale(fd):
    """70,000 dialogues that ALL overflow the in-shared-memory count table (about 190 distinct buckets each): every
    one goes through the wide-table kernel and its per-call deferred list.  Checked against the C port of the oracle
    (itself proven equal to the Python oracle in tests/test_oracle_c.py)."""
    from oracle import c_oracle
    model, corpus = _tree_model(fd, "rf", 12, 6)
    rng = np.random.default_rng(21)
    vocab = np.array([("w" + "".join(chr(97 + c) for c in rng.integers(0, 26, size=5))).encode() for _ in range(6000)], dtype=object)
    n = 70000
    rows = [b" ".join(vocab[rng.integers(0, 6000, size=200)]) for _ in range(n)]
    pk = fd.PackedText.from_strings(rows)
    res = model.transform(pk, cleaned=False)
    o = oracle_for(model)
    co = c_oracle.COracle(model._plan.stop.getOrDefault("stopWords"), False, o.num_features, o.binary, o.idf, o.classifier)
    pred, prob, raw = co.transform(pk.data, pk.offs, True)
    _check_scores(res, (pred, prob, raw))
    rowptr, idx, val = model.featurize(pk, cleaned=False)
    rp, ix, vl = co.featurize(pk.data, pk.offs, True)
    assert np.array_equal(rowptr.cpu().numpy(), rp) and np.array_equal(idx.cpu().numpy(), ix) and np.array_equal(val.cpu().numpy(), vl)
    assert (np.diff(rp) > 160).mean() > 0.99


@pytest.mark.parametrize("binary", [False, True])
def test_count_vectorizer_model_transform(fd, binary, tmp_path):
    """CountVectorizerModel (the reference training script's featuriser, fraud_detection_spark.py:52): dictionary
    lookup instead of hashing; unknown terms ignored; long terms, the empty term and 20,000 entries."""
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    rng = np.random.default_rng(6)
    corpus = synth.SynthCorpus(n_words=30000, seed=19)
    vocab = [w.decode() for w in corpus.words[corpus.n_stop:corpus.n_stop + 19990]]
    vocab = ["", "hello", "security", "antidisestablishmentarianism", "internationalisation", "social", "winner", "q", "im", "dont"] + vocab
    rng.shuffle(vocab)
    idf = np.round(rng.random(len(vocab)) * 4, 4)
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
              fd.CountVectorizerModel(vocab, inputCol="filtered_words", outputCol="raw_features", binary=binary),
              fd.IDFModel(idf, inputCol="raw_features", outputCol="features"),
              fd.LogisticRegressionModel(rng.normal(0, 1, len(vocab)), -0.3, featuresCol="features")]
    model = fd.PipelineModel(stages)
    pk, _ = corpus.sample(400, tokens=150, device="cpu", seed=2)
    d, o = pk.data.numpy(), pk.offs.numpy()
    rows = [bytes(d[o[i]:o[i + 1]]) for i in range(400)] + [t.encode() for t in messy_texts(300, seed=8)] + [t.encode() for t in _long_rows()[:8]]
    ora = oracle_for(model)
    rowptr, idx, val = model.featurize(rows)
    for t, (ix, vl) in zip(rows, csr_rows(rowptr, idx, val)):
        eix, evl, _, _ = ora.featurize(t)
        assert ix == eix and vl == evl, repr(t[:40])
    _check_scores(model.transform(rows), ora.transform(rows))
    model.save(str(tmp_path / "cv"))
    again = fd.PipelineModel.load(str(tmp_path / "cv"))
    assert again.stages[2].vocabulary == vocab
    _check_scores(again.transform(rows), ora.transform(rows))


@pytest.mark.parametrize("kind", ["rf", "gbt"])
@pytest.mark.parametrize("case", ["dense", "lookup-front", "too-many-features", "large-thresholds"])
def test_forest_walk_variants(fd, kind, case, monkeypatch):
    """The forest epilogue has two node-visit forms: a dense byte vector of counts over forest-local feature ids (at most
    2,048 distinct split features, count thresholds below 255) and the 16-bit lookup front over the dialogue's count table.
    Same dialogues, same oracle, for: the dense form, the front forced by FDB200_DENSE_WALK=0, a forest with more distinct
    split features than the dense form takes, and one whose thresholds ask for counts of several hundred (tiny idf)."""
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    rng = np.random.default_rng(31)
    nf = 1 << 16
    corpus = synth.SynthCorpus(n_words=6000, seed=5)
    n_feat_words = 5000 if case == "too-many-features" else 300
    feats = synth.hash_terms(corpus.words[corpus.n_stop:corpus.n_stop + n_feat_words] + [b""], nf)
    idf = np.log(1001.0 / (rng.integers(0, 1000, size=nf) + 1.0))
    if case == "large-thresholds":
        idf = idf * 1e-3          # thresholds stay as drawn, so the count they ask for grows a thousandfold
    if case == "lookup-front":
        monkeypatch.setenv("FDB200_DENSE_WALK", "0")
    depths = [7] * 40 if case == "too-many-features" else [6, 3, 8, 1, 5, 7, 2, 6, 4, 8, 5, 6]
    model = _forest_with_idf(fd, kind, idf, depths, nf, feats, seed=8)
    pk, _ = corpus.sample(600, tokens=120, device="cpu", seed=2)
    rows = [bytes(pk.data.numpy()[pk.offs[i]:pk.offs[i + 1]]) for i in range(pk.n_rows)]
    rows += [b"", b" ", (b"payment " * 400), (b"verify account " * 300)]      # repeated terms: counts in the hundreds
    res = model.transform([r.decode() for r in rows])
    _check_scores(res, oracle_for(model).transform(rows))


def test_feature_space_change_drops_idf_and_classifier(fd):
    """A call that changes the feature space (another numFeatures) must not leave the IDF vector and the classifier of the
    old space live: fd_transform answers FD_ERR_STATE until both are set again, and then scores in the new space."""
    import ctypes as C
    import torch
    lib = importlib.import_module("fraud-detection-spark-kafka-llm_b200._lib")
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    h = lib.Handle(0)
    dev = torch.device("cuda", 0)
    texts = [b"verify your account now", b"hello how are you", b"", b"account account payment"]
    pk = fd.PackedText.from_strings([t.decode() for t in texts]).to_device(dev)
    n = pk.n_rows
    outs = [torch.empty(s_, dtype=torch.float64, device=dev) for s_ in ((n,), (n, 2), (n, 2))]
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    def upload(nf, seed):
        rng = np.random.default_rng(seed)
        idf = rng.random(nf) + 0.5
        coef = rng.standard_normal(nf)
        h.call("fd_set_idf", lib.ptr(idf), nf)
        h.call("fd_set_linear", lib.ptr(coef), nf, C.c_double(0.1), C.c_double(0.5))
        return idf, coef

    def margins(nf, idf, coef):
        exp = []
        for t in texts:
            toks = t.decode().split(" ")
            while len(toks) > 1 and toks[-1] == "":          # Java's split drops trailing empties; "" itself stays [""]
                toks.pop()
            cnt = {}
            for w in toks:
                f = int(synth.hash_terms([w.encode()], nf)[0])
                cnt[f] = cnt.get(f, 0) + 1
            m = 0.0
            for f in sorted(cnt):
                m += (cnt[f] * idf[f]) * coef[f]
            exp.append(m + 0.1)
        return np.asarray(exp)

    def run():
        h.call("fd_transform", lib.ptr(pk.data), lib.ptr(pk.offs), n, 1, lib.ptr(outs[0]), lib.ptr(outs[1]), lib.ptr(outs[2]), st)
        torch.cuda.synchronize()
        return outs[2][:, 1].cpu().numpy()

    h.call("fd_set_hashing", 1000, 0)
    idf, coef = upload(1000, 1)
    assert np.array_equal(run(), margins(1000, idf, coef))
    h.call("fd_set_hashing", 50000, 0)                       # a larger space: the old vectors would be read out of bounds
    with pytest.raises(lib.FdError) as e:
        run()
    assert "-4" in str(e.value)                              # FD_ERR_STATE
    h.call("fd_set_idf", lib.ptr(np.ones(50000)), 50000)
    with pytest.raises(lib.FdError):
        run()                                                # the classifier is still missing
    idf, coef = upload(50000, 2)
    assert np.array_equal(run(), margins(50000, idf, coef))


@pytest.mark.parametrize("words", [[""], [], ["", ""]])
def test_degenerate_stop_word_lists(fd, shipped_model_dir, words):
    """A stop-word list that holds nothing but the empty string (or nothing at all) has no word table on the device: the
    kernels must not probe one.  Featurize and the fused LogisticRegression transform against the oracle."""
    base = fd.PipelineModel.load(shipped_model_dir)
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"),
              fd.StopWordsRemover(inputCol="words", outputCol="filtered_words", stopWords=words, caseSensitive=False)] + list(base.stages[2:])
    model = fd.PipelineModel(stages)
    texts = messy_texts(200, seed=4) + ["", " ", "a  b", "the the  the"]
    ora = oracle_for(model)
    rowptr, idx, val = model.featurize(texts)
    for t, (ix, vl) in zip(texts, csr_rows(rowptr, idx, val)):
        eix, evl, _, _ = ora.featurize(t.encode())
        assert ix == eix and vl == evl, repr(t[:50])
    _check_scores(model.transform(texts), ora.transform([t.encode() for t in texts]))


def test_featurize_finish_after_another_call_is_refused(fd, shipped_model_dir):
    """fd_featurize_begin stages its rows in the handle's workspace; a transform (or a document-frequency count) in between
    reuses that memory, so the pending fd_featurize_finish must fail with FD_ERR_STATE instead of emitting a corrupt CSR --
    and a fresh begin / finish pair works again."""
    import ctypes as C
    import torch
    lib = importlib.import_module("fraud-detection-spark-kafka-llm_b200._lib")
    model = fd.PipelineModel.load(shipped_model_dir)
    h = model.handle()
    dev = torch.device("cuda", h.device)
    texts = messy_texts(300, seed=6)
    pk = fd.PackedText.from_strings(texts).to_device(dev)
    n = pk.n_rows
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    nnz = C.c_int64(0)
    h.call("fd_featurize_begin", lib.ptr(pk.data), lib.ptr(pk.offs), n, 1, C.byref(nnz), st)
    assert nnz.value > 0
    model.transform(texts)                                   # reuses the workspace
    rowptr = torch.empty(n + 1, dtype=torch.int64, device=dev)
    idx = torch.empty(nnz.value, dtype=torch.int32, device=dev)
    val = torch.empty(nnz.value, dtype=torch.float64, device=dev)
    with pytest.raises(lib.FdError) as e:
        h.call("fd_featurize_finish", lib.ptr(rowptr), lib.ptr(idx), lib.ptr(val), nnz.value, st)
    assert "-4" in str(e.value)
    ora = oracle_for(model)
    r2, i2, v2 = model.featurize(texts)
    for t, (ix, vl) in zip(texts[:50], csr_rows(r2, i2, v2)):
        eix, evl, _, _ = ora.featurize(t.encode())
        assert ix == eix and vl == evl


def test_over_limit_status_survives_workspace_growth(fd):
    """fd_transform is asynchronous: a dialogue over the size limit is only recorded in the device-side status word.  A
    larger batch that makes the workspace grow (free + allocate) before anybody asked for the status must not lose it."""
    import torch
    lib = importlib.import_module("fraud-detection-spark-kafka-llm_b200._lib")
    rng = np.random.default_rng(3)
    nf = 1 << 18
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"),
              fd.HashingTF(inputCol="words", outputCol="features", numFeatures=nf),
              fd.LogisticRegressionModel(rng.normal(0.0, 1.0, size=nf), 0.1)]
    model = fd.PipelineModel(stages)
    letters = rng.integers(97, 123, size=(70000, 7), dtype=np.uint8)
    huge = b" ".join(bytes(r) for r in letters).decode()     # ~61k distinct buckets: over the limit of about 49k
    h = model.handle(0)
    before = int(h.lib.fd_workspace_bytes(h.h, 3))
    model.transform(["fine one", huge, "fine two"])          # (device output: no status read)
    big = ["row %d of a batch that needs a larger workspace" % i for i in range(300_000)]
    assert int(h.lib.fd_workspace_bytes(h.h, len(big))) > before
    res = model.transform(big)
    assert not torch.isnan(res["prediction"]).any()
    with pytest.raises(fd.FdError) as e:
        h.call("fd_status", None)
    assert e.value.code == lib.FD_ERR_LIMIT
    h.call("fd_status", None)                                # reported once, then clear
