"""GPU training parity: the fd_rf_* kernels (through the C-ABI) grow exactly the oracle's trees given the same
bins, bag weights and feature subsets (identical splits on identical bins -- BASELINE.json north_star), and
Pipeline.fit produces a model whose held-out AUC shows the planted signal."""
import ctypes as C
import importlib
import os

import numpy as np
import pytest
import torch

from test_training_cpu import _csr, _oracle_forest, _same_tree, _toy

pytestmark = pytest.mark.gpu
training = importlib.import_module("fraud-detection-spark-kafka-llm_b200.training")


def _fit_cuda(fd, rows_idx, rows_val, labels, f, n_trees, depth, bag, subset_fn, strategy):
    h = fd.Handle(0)
    tr = training.ForestTrainer(training.CudaTreeBackend(h), f, n_trees=n_trees, max_depth=depth, feature_subset_strategy=strategy, seed=1)
    rowptr, idx, val = (x.cuda() for x in _csr(rows_idx, rows_val))
    trees = tr.fit(rowptr, idx, val, torch.from_numpy(labels).cuda(), bag_weights=None if bag is None else torch.from_numpy(bag).cuda(),
                   subset_fn=subset_fn)
    return trees


@pytest.mark.parametrize("n,f,n_trees,depth", [(600, 40, 4, 5), (3000, 200, 6, 7)])
def test_forest_matches_oracle(fd, n, f, n_trees, depth):
    rows_idx, rows_val, labels, f = _toy(n, f, seed=n)
    rng = np.random.default_rng(2)
    bag = rng.poisson(1.0, size=(n_trees, n)).astype(np.uint8)
    k = int(np.ceil(np.sqrt(f))) + 3
    subsets = {}

    def subset_fn(t, node):
        return subsets.setdefault((t, node), np.random.default_rng([11, t, node]).permutation(f)[:k].astype(np.int32))

    trees = _fit_cuda(fd, rows_idx, rows_val, labels, f, n_trees, depth, bag, subset_fn, "sqrt")
    _, fi, fv = _csr(rows_idx, rows_val)
    thresholds = training.find_splits(fi.numpy(), fv.numpy(), f, float(n), min(32, n) - 1)
    exp = _oracle_forest(rows_idx, rows_val, labels, f, thresholds, bag, subset_fn, depth)
    assert max(t.n_nodes for t in trees) > 7
    for a, o in zip(trees, exp):
        _same_tree(a, o)


def test_decision_tree_all_features_matches_oracle(fd):
    rows_idx, rows_val, labels, f = _toy(1500, 60, seed=5)
    trees = _fit_cuda(fd, rows_idx, rows_val, labels, f, 1, 5, None, None, "auto")
    _, fi, fv = _csr(rows_idx, rows_val)
    thresholds = training.find_splits(fi.numpy(), fv.numpy(), f, 1500.0, 31)
    exp = _oracle_forest(rows_idx, rows_val, labels, f, thresholds, np.ones((1, 1500), np.uint8), lambda t, n: np.arange(f), 5)
    _same_tree(trees[0], exp[0])
    assert trees[0].n_nodes > 5


def test_long_columns_are_segmented(fd):
    """A feature present in nearly every row gives a column longer than one work unit (8192 entries)."""
    rng = np.random.default_rng(8)
    n, f = 30000, 12
    rows_idx, rows_val, labels = [], [], []
    for i in range(n):
        ix = [0] + sorted(rng.choice(np.arange(1, f), size=int(rng.integers(0, 4)), replace=False).tolist())
        c0 = int(rng.integers(1, 5))
        y = int(c0 >= 3) ^ int(rng.random() < 0.05)
        rows_idx.append(ix)
        rows_val.append([float(c0)] + [float(rng.integers(1, 3)) for _ in ix[1:]])
        labels.append(y)
    labels = np.asarray(labels, dtype=np.uint8)
    bag = rng.poisson(1.0, size=(2, n)).astype(np.uint8)
    sub = lambda t, node: np.arange(f, dtype=np.int32)
    trees = _fit_cuda(fd, rows_idx, rows_val, labels, f, 2, 3, bag, sub, "all")
    # (thresholds come from the trainer's own row sample here, so the root split is checked directly)
    assert trees[0].feature[0] == 0 and trees[1].feature[0] == 0
    assert 2.0 <= trees[0].threshold[0] <= 3.0
    stats = trees[0].stats
    assert stats[0].sum() == bag[0].sum()  # root counts = total bag weight (segmented histogram lost nothing)
    l, r = trees[0].left[0], trees[0].right[0]
    assert stats[l].sum() + stats[r].sum() == stats[0].sum()


def test_pipeline_fit_end_to_end(fd, tmp_path):
    from sklearn.metrics import roc_auc_score
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    corpus = synth.SynthCorpus(n_words=5000, seed=5)
    planted = {"ids": list(range(corpus.n_stop + 200, corpus.n_stop + 220)), "rate": 0.08}
    pk, lab = corpus.sample(6000, tokens=60, device="cuda", seed=3, planted=planted)
    n_train = 5000
    train = fd.PackedText(pk.data, pk.offs[:n_train + 1])
    test = fd.PackedText(pk.data, pk.offs[n_train:])
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
              fd.HashingTF(inputCol="filtered_words", outputCol="raw_features", numFeatures=1 << 14),
              fd.IDF(inputCol="raw_features", outputCol="features"),
              fd.RandomForestClassifier(featuresCol="features", labelCol="labels", numTrees=20, maxDepth=5, seed=42, featureSubsetStrategy="auto")]
    model = fd.Pipeline(stages=stages).fit(train, labels=lab[:n_train].cpu().numpy())
    assert [type(s).__name__ for s in model.stages][-2:] == ["IDFModel", "RandomForestClassificationModel"]
    out = model.transform(test)
    auc = roc_auc_score(lab[n_train:].cpu().numpy(), out["rawPrediction"][:, 1].cpu().numpy())
    assert auc > 0.9, auc
    model.save(str(tmp_path / "rf"))
    again = fd.PipelineModel.load(str(tmp_path / "rf"))
    assert torch.equal(again.transform(test)["prediction"], out["prediction"])
    imp = model.stages[-1].featureImportances
    assert abs(imp.sum() - 1.0) < 1e-9 and (imp > 0).sum() >= 5


def test_training_edge_cases(fd):
    """Degenerate inputs the level loop must survive: a pure node at the root, rows without any feature, fewer rows
    than a warp, maxDepth = 0, and a feature whose every value is identical (no threshold)."""
    h = fd.Handle(0)
    mk = lambda depth, trees=1: training.ForestTrainer(training.CudaTreeBackend(h), 8, n_trees=trees, max_depth=depth, feature_subset_strategy="all", seed=3)
    # all labels equal -> the root stays a leaf holding every row
    rows_idx = [[0, 1], [1], [], [2, 3]] * 3
    rows_val = [[1.0, 2.0], [1.0], [], [3.0, 1.0]] * 3
    rp, ix, vl = (x.cuda() for x in _csr(rows_idx, rows_val))
    t = mk(4).fit(rp, ix, vl, torch.zeros(12, dtype=torch.uint8).cuda())[0]
    assert t.n_nodes == 1 and t.stats[0].tolist() == [12.0, 0.0] and t.prediction[0] == 0.0
    # labels decided by the PRESENCE of feature 2 (rows without it are all-absent on that column)
    y = np.array([0, 0, 0, 1] * 3, dtype=np.uint8)
    t = mk(3).fit(rp, ix, vl, torch.from_numpy(y).cuda())[0]
    assert t.feature[0] in (2, 3) and t.n_nodes == 3 and sorted(t.stats[1:].sum(axis=1).tolist()) == [3.0, 9.0]
    assert {tuple(r) for r in t.stats[1:].tolist()} == {(9.0, 0.0), (0.0, 3.0)}
    # maxDepth = 0: a single leaf with the class totals
    t = mk(0).fit(rp, ix, vl, torch.from_numpy(y).cuda())[0]
    assert t.n_nodes == 1 and t.stats[0].tolist() == [9.0, 3.0]
    # constant feature (same value in every row) yields no threshold and is never chosen
    rows_idx = [[5]] * 10
    rows_val = [[2.5]] * 10
    rp, ix, vl = (x.cuda() for x in _csr(rows_idx, rows_val))
    y = np.array([0, 1] * 5, dtype=np.uint8)
    t = mk(3).fit(rp, ix, vl, torch.from_numpy(y).cuda())[0]
    assert t.n_nodes == 1 and t.stats[0].tolist() == [5.0, 5.0]
    # an empty shard (this rank holds no rows) must not break the kernels
    rp0 = torch.zeros(1, dtype=torch.int64).cuda()
    t = mk(2).fit(rp0, torch.zeros(0, dtype=torch.int32).cuda(), torch.zeros(0, dtype=torch.float64).cuda(), torch.zeros(0, dtype=torch.uint8).cuda())[0]
    assert t.n_nodes == 1 and t.stats[0].tolist() == [0.0, 0.0]


def test_decision_tree_pipeline_matches_oracle_end_to_end(fd):
    """Pipeline.fit with a DecisionTreeClassifier on text: featurise on the GPU, train on the GPU, and compare the
    tree with the oracle trained on the oracle's own features (thresholds from the same full-sample rule)."""
    from helpers import oracle_for
    from oracle import spark_oracle as so
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    corpus = synth.SynthCorpus(n_words=400, seed=8)
    planted = {"ids": list(range(corpus.n_stop + 30, corpus.n_stop + 36)), "rate": 0.15}
    pk, lab = corpus.sample(900, tokens=40, device="cpu", seed=4, planted=planted)
    data, offs = pk.data.numpy(), pk.offs.numpy()
    texts = [bytes(data[offs[i]:offs[i + 1]]).decode() for i in range(900)]
    nf = 512
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
              fd.HashingTF(inputCol="filtered_words", outputCol="raw_features", numFeatures=nf),
              fd.IDF(inputCol="raw_features", outputCol="features"),
              fd.DecisionTreeClassifier(featuresCol="features", labelCol="labels", maxDepth=5)]
    model = fd.Pipeline(stages=stages).fit(texts, labels=lab.numpy())
    tree = model.stages[-1].trees[0]
    # oracle side: features from the oracle featuriser with the fitted idf, then oracle training
    feat_only = fd.PipelineModel(model.stages[:4])
    ora = oracle_for(feat_only)
    rows = [ora.featurize(t.encode())[:2] for t in texts]
    rows_idx, rows_val = [r[0] for r in rows], [r[1] for r in rows]
    fi = np.asarray([x for r in rows_idx for x in r]); fv = np.asarray([x for r in rows_val for x in r])
    thresholds = training.find_splits(fi, fv, nf, 900.0, 31)
    exp = _oracle_forest(rows_idx, rows_val, lab.numpy().astype(np.uint8), nf, thresholds, np.ones((1, 900), np.uint8), lambda t, n: np.arange(nf), 5)
    _same_tree(tree, exp[0])
    assert tree.n_nodes >= 7
    # and the fitted idf equals the oracle's IDF.fit on the same rows
    tf_rows = [so.featurize_bytes(t.encode(), ora.stop_words, nf, None)[:2] for t in texts]
    idf, df, m = so.idf_fit(tf_rows, nf)
    assert np.array_equal(idf.view(np.uint64), model.stages[3].idf.view(np.uint64)) and m == 900


def test_count_vectorizer_fit_and_reference_training_pipeline(fd, tmp_path):
    """The reference's own training pipeline shape (fraud_detection_spark.py:47-65): Tokenizer -> StopWordsRemover ->
    CountVectorizer(vocabSize) -> IDF -> DecisionTreeClassifier, fitted end to end on the GPU.  The vocabulary (terms,
    total counts, document frequencies) is checked against the oracle's CountVectorizer.fit."""
    from helpers import oracle_for
    from oracle import spark_oracle as so
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    corpus = synth.SynthCorpus(n_words=3000, seed=23)
    planted = {"ids": list(range(corpus.n_stop + 50, corpus.n_stop + 58)), "rate": 0.1}
    pk, lab = corpus.sample(1500, tokens=80, device="cpu", seed=6, planted=planted)
    data, offs = pk.data.numpy(), pk.offs.numpy()
    texts = [bytes(data[offs[i]:offs[i + 1]]).decode() for i in range(1500)]
    texts[3] = "Hello, WORLD!! internationalisation antidisestablishmentarianism  hello the the"
    texts[4] = ""
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
              fd.CountVectorizer(inputCol="filtered_words", outputCol="raw_features", vocabSize=500),
              fd.IDF(inputCol="raw_features", outputCol="features"),
              fd.DecisionTreeClassifier(featuresCol="features", labelCol="labels", maxDepth=5)]
    model = fd.Pipeline(stages=stages).fit(texts, labels=lab.numpy())
    cv = model.stages[2]
    assert type(cv).__name__ == "CountVectorizerModel" and len(cv.vocabulary) == 500
    sw = [w.encode() for w in model.stages[1].getOrDefault("stopWords")]
    docs = [[t for t in so.java_split_ws(so.clean_bytes(x.encode())) if t not in set(sw)] for x in texts]
    vocab, tc, df = so.count_vectorizer_fit(docs, 500)
    assert [v.encode() for v in cv.vocabulary] == vocab
    for b, (c, d) in cv.term_counts.items():
        assert (c, d) == (tc[b], df[b]), b
    assert b"" in tc  # the empty token is a term like any other
    # the fitted pipeline scores like the oracle built from the same stages
    ora = oracle_for(model)
    res = model.transform(texts)
    exp = ora.transform([t.encode() for t in texts])
    assert np.array_equal(res["prediction"].cpu().numpy(), exp[0]) and np.array_equal(res["rawPrediction"].cpu().numpy(), exp[2])
    model.save(str(tmp_path / "ref_shape"))
    again = fd.PipelineModel.load(str(tmp_path / "ref_shape"))
    assert again.stages[2].vocabulary == cv.vocabulary and torch.equal(again.transform(texts)["prediction"], res["prediction"])
    # minDF filters by document frequency
    cv2 = fd.CountVectorizer(inputCol="filtered_words", outputCol="raw_features", vocabSize=100000, minDF=3.0)
    m2 = fd.Pipeline(stages=[stages[0], stages[1], cv2]).fit(texts)
    exp2, _, _ = so.count_vectorizer_fit(docs, 100000, min_df=3.0)
    assert [v.encode() for v in m2.stages[2].vocabulary] == exp2


def test_reference_trainer_flow_example(fd, tmp_path):
    """examples/fraud_detection_b200.py = the reference's fraud_detection_spark.py main flow on this stack."""
    import runpy
    mod = runpy.run_path(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples", "fraud_detection_b200.py"))
    report, models = mod["main"](["--rows", "1600", "--trees", "20", "--save", str(tmp_path / "fraud_detection_model")])
    for name in ("DecisionTree", "RandomForest"):
        m = report[name]["Test"]["metrics"]
        lo_auc, lo_acc = (0.9, 0.85) if name == "RandomForest" else (0.75, 0.7)  # one depth-5 tree on 20 planted words
        assert m["AUC"] > lo_auc and m["Accuracy"] > lo_acc, (name, m)
        assert report[name]["Test"]["confusion_matrix"].sum() > 200
    saved = fd.PipelineModel.load(str(tmp_path / "fraud_detection_model"))
    assert [type(s).__name__ for s in saved.stages] == ["Tokenizer", "StopWordsRemover", "CountVectorizerModel", "IDFModel", "DecisionTreeClassificationModel"]


def test_bag_poisson_kernel(fd):
    """fd_rf_bag_poisson: Poisson(rate) statistics per tree, reproducible for a seed, independent of launch shape (a
    prefix of the rows gets the same draws), root class totals equal to the weights summed by label."""
    h = fd.Handle(0)
    b = training.CudaTreeBackend(h)
    n, T = 200_003, 7
    lab = (torch.rand(n, device="cuda") < 0.3).to(torch.uint8)
    for rate in (1.0, 0.5, 3.0):
        bagw, root = b.bag_poisson(1234, rate, T, lab)
        again, _ = b.bag_poisson(1234, rate, T, lab)
        other, _ = b.bag_poisson(1235, rate, T, lab)
        assert torch.equal(bagw, again) and not torch.equal(bagw, other)
        w = bagw.to(torch.float64)
        assert abs(float(w.mean()) - rate) < 0.02 * max(1.0, rate) and abs(float(w.var()) - rate) < 0.05 * max(1.0, rate)
        p0 = float((bagw == 0).double().mean())
        assert abs(p0 - np.exp(-rate)) < 0.01
        # trees are independent streams
        assert abs(float(torch.corrcoef(w[:2])[0, 1])) < 0.02
        w1 = (bagw * lab.bool()).sum(dim=1, dtype=torch.int64)
        exp_root = torch.stack([bagw.sum(dim=1, dtype=torch.int64) - w1, w1], 1)
        assert torch.equal(root, exp_root)
        short, _ = b.bag_poisson(1234, rate, T, lab[:1001])
        assert torch.equal(short, bagw[:, :1001])


def test_root_histogram_at_scale_equals_bincount(fd):
    """A C5-sized self-check that needs no oracle: the root-level histogram of fd_rf_histogram over 6M rows / 120M entries
    (columns of up to several hundred segments, u32 cells in the tens of millions) equals torch.bincount over the same
    (item, bin, label) triples weighted by the bag weights."""
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev)
    g.manual_seed(12)
    n, per_row, F = 6_000_000, 20, 4096
    # feature ids skewed towards the low ones (long columns), distinct inside a row by construction: base + stride * j
    base = (torch.rand(n, generator=g, device=dev) ** 3 * 64).long()
    stride = torch.randint(1, F // per_row, (n,), generator=g, device=dev)
    idx = (base[:, None] + stride[:, None] * torch.arange(per_row, device=dev)[None, :]).to(torch.int32).reshape(-1)
    assert int(idx.max()) < F
    nnz = idx.numel()
    bins = torch.randint(0, 32, (nnz,), generator=g, device=dev, dtype=torch.uint8)
    bins[torch.rand(nnz, generator=g, device=dev) < 0.9] = 3  # most of a column in one bin: cells in the millions
    labels = (torch.rand(n, generator=g, device=dev) < 0.4).to(torch.uint8)
    bagw = torch.randint(0, 6, (2, n), generator=g, device=dev, dtype=torch.uint8)
    # CSC exactly as ForestTrainer.fit lays it out
    row_of = torch.arange(n, dtype=torch.int32, device=dev).repeat_interleave(per_row)
    order = torch.argsort(idx, stable=True)
    row_lab = row_of | (labels.to(torch.int32) << 31).repeat_interleave(per_row)
    colptr = torch.zeros(F + 1, dtype=torch.int64, device=dev)
    torch.cumsum(torch.bincount(idx, minlength=F), 0, out=colptr[1:])
    csc = (colptr, row_lab[order].contiguous(), bins[order].contiguous())
    state = bagw.to(torch.int32) | (1 << 8)
    feats = torch.cat([torch.arange(0, 40, device=dev), torch.randint(40, F, (260,), generator=g, device=dev)]).to(torch.int32)
    tree = 1
    items = (torch.full_like(feats, tree), torch.ones_like(feats), feats)
    backend = training.CudaTreeBackend(fd.Handle(0))
    units = training._units_device(colptr[1:] - colptr[:-1], feats.long())
    assert int(units[0].numel()) > feats.numel() + 500, "the test is meant to cover segmented columns"
    hist = backend.histogram(csc, state, items, units).to(torch.int64)
    # reference: every entry whose feature is in the item list (duplicates in the list each get the full count)
    w = bagw[tree].to(torch.int64)[row_of.long()]
    key = idx.long() * 64 + bins.long() * 2 + labels.long().repeat_interleave(per_row)
    per_feature = torch.bincount(key, weights=w.double(), minlength=F * 64).view(F, 32, 2).round().to(torch.int64)
    assert torch.equal(hist, per_feature[feats.long()])
    assert int(hist.max()) > 1_000_000 and int(hist.sum()) > 0


def test_bag_weight_overflow_guard(fd):
    """Histogram cells are int32 after the all-reduce: a forest whose total bag weight per tree reaches 2^31 is refused."""
    h = fd.Handle(0)
    tr = training.ForestTrainer(training.CudaTreeBackend(h), 4, n_trees=2, max_depth=2, feature_subset_strategy="all", seed=1)
    n = 9_000_000
    rowptr = torch.arange(n + 1, dtype=torch.int64, device="cuda")
    idx = torch.zeros(n, dtype=torch.int32, device="cuda")
    val = torch.ones(n, dtype=torch.float64, device="cuda")
    labels = torch.zeros(n, dtype=torch.uint8, device="cuda")
    bag = torch.full((2, n), 255, dtype=torch.uint8, device="cuda")
    with pytest.raises(ValueError, match="2\\^31"):
        tr.fit(rowptr, idx, val, labels, bag_weights=bag, subset_fn=lambda t, node: np.arange(4, dtype=np.int32))


def test_doc_freq_private_vectors_at_scale(fd):
    """fd_doc_freq on a matrix large enough for its private count vectors (>= 2^22 entries): equals a bincount of the
    entries with a positive value, with a heavily skewed feature distribution and some non-positive values."""
    import torch
    lib = importlib.import_module("fraud-detection-spark-kafka-llm_b200._lib")
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev)
    g.manual_seed(5)
    nnz, nf = 6_000_000, 5000
    idx = (torch.rand(nnz, generator=g, device=dev) ** 4 * nf).to(torch.int32).clamp_(max=nf - 1)   # a few very hot features
    val = torch.rand(nnz, generator=g, device=dev, dtype=torch.float64) - 0.1                   # ~10% not positive
    val[::1000] = 0.0
    h = lib.Handle(0)
    df = torch.empty(nf, dtype=torch.int64, device=dev)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    h.call("fd_doc_freq", lib.ptr(idx), lib.ptr(val), nnz, nf, lib.ptr(df), st)
    exp = torch.bincount(idx[val > 0].long(), minlength=nf)
    assert torch.equal(df, exp)
    assert int(df.max()) > 100_000
    # and the scaling kernel: one exact multiply per entry
    idf = torch.rand(nf, generator=g, device=dev, dtype=torch.float64)
    want = val * idf[idx.long()]
    h.call("fd_idf_scale", lib.ptr(idx), lib.ptr(val), nnz, lib.ptr(idf), nf, st)
    assert torch.equal(val, want)


@pytest.mark.parametrize("num_features", [300, 1 << 18, (1 << 25) + 3])
def test_csc_build_equals_stable_sort(fd, num_features):
    """fd_csc_build (a radix sort over just the bits a feature id can have) against a stable argsort by feature id:
    columns in feature order, rows ascending inside a column, label in bit 31 of the row word."""
    lib = importlib.import_module("fraud-detection-spark-kafka-llm_b200._lib")
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev)
    g.manual_seed(3)
    n_rows, per_row = 20_000, 12
    base = torch.randint(0, num_features, (n_rows, 1), generator=g, device=dev)
    idx = ((base + torch.arange(per_row, device=dev) * 7919) % num_features).sort(dim=1).values.to(torch.int32).reshape(-1)
    rowptr = torch.arange(n_rows + 1, device=dev, dtype=torch.int64) * per_row
    bins = torch.randint(0, 32, (idx.numel(),), generator=g, device=dev).to(torch.uint8)
    labels = (torch.rand(n_rows, generator=g, device=dev) < 0.3).to(torch.uint8)
    backend = training.CudaTreeBackend(lib.Handle(0))
    rowidx, cbins = backend.csc_build(rowptr, idx, bins, labels, num_features)
    order = torch.argsort(idx, stable=True)
    row_of = torch.arange(n_rows, device=dev).repeat_interleave(per_row)
    want = (row_of | (labels.long() << 31).repeat_interleave(per_row))[order]
    assert torch.equal(rowidx.view(torch.int32).long() & 0xFFFFFFFF, want)
    assert torch.equal(cbins, bins[order])
