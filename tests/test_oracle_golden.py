"""Pin the CPU oracle against the Spark-written numbers in the reference's saved model (G1/G2/G3,
SURVEY.md section 8c) and against its own committed provisional outputs."""
import math

import numpy as np

from oracle import spark_oracle as so


def _shipped(fd, shipped_model_dir):
    return fd.PipelineModel.load(shipped_model_dir)


def test_g1_idf_formula_bit_exact(fd, shipped_model_dir, golden):
    m = _shipped(fd, shipped_model_dir)
    idf_stage = m.stages[3]
    assert idf_stage.numDocs == golden["G1"]["num_docs"] == 1150
    mine = so.idf_from_doc_freq(idf_stage.docFreq, idf_stage.numDocs, 0)
    assert mine.size == golden["G1"]["n"] == 10000
    assert np.array_equal(mine.view(np.uint64), idf_stage.idf.view(np.uint64))  # all 10,000 doubles, bit for bit
    for df, v in golden["G1"]["examples"].items():
        assert math.log((1150 + 1.0) / (int(df) + 1.0)) == v
    # the product-side IDF.fit formula gives the same bits
    fitted = fd.IDF(minDocFreq=0).fit_from_doc_freq(idf_stage.docFreq, idf_stage.numDocs)
    assert np.array_equal(fitted.idf.view(np.uint64), idf_stage.idf.view(np.uint64))


def test_g2_hash_and_tokenizer_known_answers(fd, shipped_model_dir, golden):
    m = _shipped(fd, shipped_model_dir)
    df, nd = m.stages[3].docFreq, m.stages[3].numDocs
    full = [int(i) for i in np.nonzero(df == nd)[0]]
    assert full == golden["G2"]["full_df_buckets"] == [2833, 3372, 9168, 9889]
    # every training dialogue starts "Innocent: Hello. Suspect: Hello..." => these four tokens hit every document
    expect = {"innocent": (660822833, 2833), "": (142593372, 3372), "suspect": (-780940832, 9168), "hello": (-488910111, 9889)}
    for w, (h, b) in expect.items():
        assert so.murmur3_x86_32(w.encode(), 42) == h
        assert so.non_negative_mod(h, 10000) == b
        assert df[b] == nd
    for w, rec in golden["G2"]["words"].items():
        h = so.murmur3_x86_32(w.encode(), 42)
        assert h == rec["hash"] and so.non_negative_mod(h, 10000) == rec["bucket"]
        assert df[rec["bucket"]] == rec["doc_freq"] > 0
    # the pre-3.0 hash variant lands on never-seen buckets: rejected by the artefact
    for w in ("suspect", "hello"):
        assert df[so.non_negative_mod(so.murmur3_x86_32_legacy(w.encode(), 42), 10000)] == 0
    # "Innocent: Hello." cleaned + tokenised gives an empty token between ". " halves
    toks = so.java_split_ws(so.clean_bytes(b"Innocent: Hello.  Suspect: Hello"))
    assert toks == [b"innocent", b"hello", b"", b"suspect", b"hello"]


def test_g3_lr_support(fd, shipped_model_dir, golden):
    m = _shipped(fd, shipped_model_dir)
    coef, df, nd = m.stages[4].coefficients, m.stages[3].docFreq, m.stages[3].numDocs
    nz = set(np.nonzero(coef)[0].tolist())
    assert len(nz) == golden["G3"]["coef_nnz"] == 4081
    assert nz == set(np.nonzero((df > 0) & (df < nd))[0].tolist())
    assert m.stages[4].intercept == golden["G3"]["intercept"] == -7.218662911169931


def test_java_split_examples():
    assert so.java_split_ws(b"a  b") == [b"a", b"", b"b"]
    assert so.java_split_ws(b" a") == [b"", b"a"]
    assert so.java_split_ws(b"a ") == [b"a"]
    assert so.java_split_ws(b"  a b  ") == [b"", b"", b"a", b"b"]
    assert so.java_split_ws(b" ") == []
    assert so.java_split_ws(b"") == [b""]
    assert so.java_split_ws(b"a\tb\nc\x0bd\x0ce\rf") == [b"a", b"b", b"c", b"d", b"e", b"f"]


def test_clean_bytes_matches_unicode_rule_on_every_code_point():
    keep = []
    for cp in range(0x110000):
        if 0xD800 <= cp <= 0xDFFF:
            continue
        ch = chr(cp)
        if so.clean_text(ch) != so.clean_bytes(ch.encode("utf-8")).decode("ascii"):
            raise AssertionError(hex(cp))
        if so.clean_text(ch) and cp >= 0x80:
            keep.append(cp)
    assert keep == [0x130, 0x212A]
    s = "İK test K\tx\ny Ünï-cödé 123 ŉ ǰ"
    assert so.clean_text(s) == so.clean_bytes(s.encode()).decode()


def test_non_negative_mod():
    for x in (-2**31, -1, 0, 1, 2**31 - 1, -780940832):
        for m in (1, 7, 10000, 262144, 2**31 - 1):
            assert so.non_negative_mod(x, m) == x % m


def test_provisional_outputs_stable(fd, shipped_model_dir, golden):
    m = _shipped(fd, shipped_model_dir)
    pipe = so.OraclePipeline(m.stages[1].getOrDefault("stopWords"), False, 10000, False, m.stages[3].idf,
                             {"kind": "lr", "coef": m.stages[4].coefficients, "intercept": m.stages[4].intercept, "threshold": 0.5})
    for rec in golden["provisional_oracle_outputs"]:
        idx, val, ntok, nkept = pipe.featurize(rec["text"].encode("utf-8"))
        assert (ntok, nkept, idx) == (rec["n_tokens"], rec["n_kept"], rec["idx"])
        assert val == rec["val"]
        pred, prob, raw = pipe.score(idx, val)
        assert raw[1] == rec["margin"] and prob == rec["probability"] and pred == rec["prediction"]
    long = [r for r in golden["provisional_oracle_outputs"] if r["n_tokens"] == 375][0]
    assert (long["n_kept"], len(long["idx"]), long["prediction"]) == (192, 112, 1.0)


def test_tree_predictors_small_case():
    t = so.Tree([3, -1, -1], [1.5, 0, 0], [1, -1, -1], [2, -1, -1], [[0, 0], [8, 2], [1, 9]])
    assert so.dt_predict([3], [1.0], t) == (0.0, [0.8, 0.2], [8.0, 2.0])
    assert so.dt_predict([3], [2.0], t)[0] == 1.0
    assert so.dt_predict([], [], t)[0] == 0.0  # absent feature reads 0.0 <= 1.5 -> left
    p, prob, raw = so.rf_predict([3], [2.0], [t, t])
    assert raw == [0.2, 1.8] and p == 1.0
    r = so.Tree([3, -1, -1], [1.5, 0, 0], [1, -1, -1], [2, -1, -1], [0.0, -0.5, 0.25])
    p, prob, raw = so.gbt_predict([3], [2.0], [r, r], [1.0, 0.1])
    assert raw[1] == 0.25 + 0.25 * 0.1 and p == 1.0 and prob[0] == 1.0 / (1.0 + math.exp(2 * raw[1]))
