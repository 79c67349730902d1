"""CPU oracle: a plain-Python/numpy restatement of the Spark 3.5.5 semantics on the hot path.

THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import it.  The product path
(fraud-detection-spark-kafka-llm_b200/) never imports anything from oracle/ and fails loudly
when its CUDA library is missing.

What it restates
----------------
The reference (/root/reference) holds *no* arithmetic of its own for this path: every step runs
inside the un-vendored third-party dependency ``pyspark==3.5.5`` (reference ``requirements.txt:1``;
every saved stage is stamped ``"sparkVersion":"3.5.5"``).  The reference's call sites are
``utils/agent_api.py:129`` (PipelineModel.load), ``utils/agent_api.py:143-144`` (cleaning),
``utils/agent_api.py:158-167`` (transform + result extraction), ``fraud_detection_spark.py:42-45``
(cleaning, training side), ``fraud_detection_spark.py:50-53`` (feature stages),
``fraud_detection_spark.py:59-74`` (tree classifiers) and ``fraud_detection_spark.py:85-91`` (fit).
Each function below names the Spark class whose published algorithm it restates and the reference
call site / artefact that anchors it.

Parity status
-------------
* Hash function, seed, tail handling, negative modulo, lower-casing, punctuation removal, hashing of
  empty tokens and the IDF formula are PINNED by Spark-written numbers inside the reference's saved
  model (golden vectors G1/G2/G3 in tests/golden/, checked by tests/test_oracle_golden.py).
* LogisticRegression scoring is anchored on the saved coefficient matrix/intercept (format pinned),
  its arithmetic order is restated from Spark's BLAS.dot(sparse, sparse) -- "parity unpinned" beyond
  the format.
* DecisionTree / RandomForest / GBT scoring, the tree on-disk format and RandomForest training are
  restated from the published Spark sources with NO Spark-written fixture in the reference:
  **parity unpinned** for those parts.

Everything is fp64 (Python float / numpy float64); integer work is exact.
"""
from __future__ import annotations

import math
import re
from typing import Iterable, List, Optional, Sequence, Tuple

import numpy as np

# ---------------------------------------------------------------------------------------------
# A.1  caller-side cleaning            reference: utils/agent_api.py:143-144,
#                                                 fraud_detection_spark.py:42-45
# ---------------------------------------------------------------------------------------------
_CLEAN_RE = re.compile(r"[^a-zA-Z ]")


def clean_text(s: str) -> str:
    """``regexp_replace(lower(col), "[^a-zA-Z ]", "")`` on one string.

    Spark's ``lower`` is ``UTF8String.toLowerCase`` (Java ``String.toLowerCase`` off the ASCII fast
    path); Python's ``str.lower`` follows the same Unicode case mapping.  The regex then keeps only
    ASCII letters and U+0020.
    """
    return _CLEAN_RE.sub("", s.lower())


def clean_bytes(b: bytes) -> bytes:
    """Byte-level statement of :func:`clean_text` for valid UTF-8 input (what the kernel implements).

    ASCII letters are lower-cased, 0x20 is kept, every other ASCII byte is dropped, every non-ASCII
    sequence is dropped except U+0130 (C4 B0 -> 'i', its lower case is 'i' + U+0307) and U+212A
    (E2 84 AA -> 'k', KELVIN SIGN).  tests/test_oracle_golden.py checks this against
    :func:`clean_text` over every Unicode code point.
    """
    out = bytearray()
    n = len(b)
    for p in range(n):
        c = b[p]
        if 0x41 <= c <= 0x5A:
            out.append(c | 0x20)
        elif 0x61 <= c <= 0x7A or c == 0x20:
            out.append(c)
        elif c == 0xB0 and p >= 1 and b[p - 1] == 0xC4:
            out.append(0x69)  # 'i'
        elif c == 0xAA and p >= 2 and b[p - 1] == 0x84 and b[p - 2] == 0xE2:
            out.append(0x6B)  # 'k'
    return bytes(out)


# ---------------------------------------------------------------------------------------------
# A.2  Tokenizer                       reference: fraud_detection_spark.py:50,
#                                      stages/0_Tokenizer_38303107ee91/metadata/part-00000:1
# ---------------------------------------------------------------------------------------------
_JAVA_WS = b" \t\n\x0b\x0c\r"  # Java regex \s


def java_split_ws(b: bytes) -> List[bytes]:
    """Java ``String.split("\\\\s")`` (limit 0) on the UTF-8 bytes of a string.

    Every single white-space char is a boundary (runs are NOT collapsed), leading/interior empty
    strings are kept, trailing empty strings are removed, and when nothing matches the whole input
    is returned -- so ``"" -> [""]`` while ``" " -> []``.
    """
    parts: List[bytes] = []
    start = 0
    matched = False
    for p, c in enumerate(b):
        if c in _JAVA_WS:
            matched = True
            parts.append(b[start:p])
            start = p + 1
    if not matched:
        return [b]
    parts.append(b[start:])
    while parts and parts[-1] == b"":
        parts.pop()
    return parts


def tokenize(s: str) -> List[str]:
    """``org.apache.spark.ml.feature.Tokenizer``: ``_.toLowerCase.split("\\\\s")``."""
    return [t.decode("utf-8") for t in java_split_ws(s.lower().encode("utf-8"))]


# ---------------------------------------------------------------------------------------------
# A.3  StopWordsRemover                reference: fraud_detection_spark.py:51,
#                                      stages/1_StopWordsRemover_8c0b00b256b3/metadata/part-00000:1
# ---------------------------------------------------------------------------------------------
def remove_stopwords(tokens: Sequence[str], stop_words: Iterable[str], case_sensitive: bool = False) -> List[str]:
    """Keep ``t`` iff it is not in the stop-word set (compared lower-cased unless caseSensitive)."""
    if case_sensitive:
        sw = set(stop_words)
        return [t for t in tokens if t not in sw]
    sw = {w.lower() for w in stop_words}
    return [t for t in tokens if t.lower() not in sw]


# ---------------------------------------------------------------------------------------------
# A.4  HashingTF                       reference artefact:
#                                      stages/2_HashingTF_e7eba1072633/metadata/part-00000:1
#      org.apache.spark.unsafe.hash.Murmur3_x86_32.hashUnsafeBytes2, Utils.nonNegativeMod
# ---------------------------------------------------------------------------------------------
_M32 = 0xFFFFFFFF


def _rotl32(x: int, r: int) -> int:
    return ((x << r) | (x >> (32 - r))) & _M32


def murmur3_x86_32(data: bytes, seed: int = 42) -> int:
    """Standard MurmurHash3_x86_32 (little-endian blocks, LE-assembled tail); result as signed int32."""
    c1, c2 = 0xCC9E2D51, 0x1B873593
    h = seed & _M32
    n = len(data)
    nblocks = n // 4
    for i in range(nblocks):
        k = int.from_bytes(data[4 * i:4 * i + 4], "little")
        k = (k * c1) & _M32
        k = _rotl32(k, 15)
        k = (k * c2) & _M32
        h ^= k
        h = _rotl32(h, 13)
        h = (h * 5 + 0xE6546B64) & _M32
    tail = data[4 * nblocks:]
    if tail:
        k = int.from_bytes(tail, "little")
        k = (k * c1) & _M32
        k = _rotl32(k, 15)
        k = (k * c2) & _M32
        h ^= k
    h ^= n
    h ^= h >> 16
    h = (h * 0x85EBCA6B) & _M32
    h ^= h >> 13
    h = (h * 0xC2B2AE35) & _M32
    h ^= h >> 16
    return h - (1 << 32) if h & 0x80000000 else h


def murmur3_x86_32_legacy(data: bytes, seed: int = 42) -> int:
    """Pre-3.0 Spark variant (``hashUnsafeBytes``: tail bytes mixed one at a time as signed bytes).

    Only used by the golden test to show that the saved artefact REJECTS it (G2)."""
    c1, c2 = 0xCC9E2D51, 0x1B873593

    def mix_k1(k):
        k = (k * c1) & _M32
        k = _rotl32(k, 15)
        return (k * c2) & _M32

    def mix_h1(h, k):
        h ^= k
        h = _rotl32(h, 13)
        return (h * 5 + 0xE6546B64) & _M32

    h = seed & _M32
    n = len(data)
    aligned = n - n % 4
    for i in range(0, aligned, 4):
        h = mix_h1(h, mix_k1(int.from_bytes(data[i:i + 4], "little")))
    for i in range(aligned, n):
        byte = data[i] - 256 if data[i] >= 128 else data[i]
        h = mix_h1(h, mix_k1(byte & _M32))
    h ^= n
    h ^= h >> 16
    h = (h * 0x85EBCA6B) & _M32
    h ^= h >> 13
    h = (h * 0xC2B2AE35) & _M32
    h ^= h >> 16
    return h - (1 << 32) if h & 0x80000000 else h


def non_negative_mod(x: int, mod: int) -> int:
    """``Utils.nonNegativeMod``: ``((x % mod) + mod) % mod`` with Java's sign-following ``%``."""
    raw = int(math.fmod(x, mod))  # sign follows the dividend, like Java
    return raw + (mod if raw < 0 else 0)


def hashing_tf_index(term: bytes, num_features: int) -> int:
    return non_negative_mod(murmur3_x86_32(term, 42), num_features)


def hashing_tf(tokens: Sequence[bytes], num_features: int, binary: bool = False) -> Tuple[List[int], List[float]]:
    """Term-frequency sparse vector: indices ascending (``Vectors.sparse`` sorts), f64 counts."""
    counts = {}
    for t in tokens:
        i = hashing_tf_index(t, num_features)
        counts[i] = 1.0 if binary else counts.get(i, 0.0) + 1.0
    idx = sorted(counts)
    return idx, [counts[i] for i in idx]


# ---------------------------------------------------------------------------------------------
# CountVectorizer / CountVectorizerModel        reference: fraud_detection_spark.py:52      -- parity unpinned
# ---------------------------------------------------------------------------------------------
def count_vectorize(tokens: Sequence[bytes], vocab_index: dict, binary: bool = False, min_tf: float = 1.0):
    """``CountVectorizerModel.transform`` on one document: dictionary lookup, unknown terms ignored, counts filtered
    by minTF (absolute when >= 1, else a fraction of the document's token count), indices ascending."""
    counts = {}
    for t in tokens:
        i = vocab_index.get(t)
        if i is not None:
            counts[i] = counts.get(i, 0.0) + 1.0
    eff = min_tf if min_tf >= 1.0 else len(tokens) * min_tf
    idx = sorted(i for i, c in counts.items() if c >= eff)
    return idx, [1.0 if binary else counts[i] for i in idx]


def count_vectorizer_fit(docs: Sequence[Sequence[bytes]], vocab_size: int, min_df: float = 1.0, max_df: float = 2.0 ** 63):
    """``CountVectorizer.fit``: total term counts and document frequencies, optional df filter, then the
    ``vocab_size`` most frequent terms.  Spark breaks count ties by partition order (unspecified); here ties go to the
    smaller term bytes, and so does the product code."""
    tc, df = {}, {}
    for d in docs:
        seen = set()
        for t in d:
            tc[t] = tc.get(t, 0) + 1
            if t not in seen:
                seen.add(t)
                df[t] = df.get(t, 0) + 1
    n = len(docs)
    lo = min_df if min_df >= 1.0 else min_df * n
    hi = max_df if max_df >= 1.0 else max_df * n
    terms = [t for t in tc if lo <= df[t] <= hi]
    terms.sort(key=lambda t: (-tc[t], t))
    return terms[:vocab_size], tc, df


# ---------------------------------------------------------------------------------------------
# A.5  IDF / IDFModel                  reference: fraud_detection_spark.py:53,
#                                      stages/3_IDF_58bd96296a82/{metadata,data}
# ---------------------------------------------------------------------------------------------
def idf_from_doc_freq(doc_freq: Sequence[int], num_docs: int, min_doc_freq: int = 0) -> np.ndarray:
    """``idf[j] = log((m + 1.0) / (df[j] + 1.0))`` (0.0 below minDocFreq); pinned bit-exactly by G1."""
    out = np.zeros(len(doc_freq), dtype=np.float64)
    for j, df in enumerate(doc_freq):
        if df >= min_doc_freq:
            out[j] = math.log((num_docs + 1.0) / (df + 1.0))
    return out


def idf_fit(rows: Sequence[Tuple[Sequence[int], Sequence[float]]], num_features: int, min_doc_freq: int = 0):
    """``IDF.fit``: document frequency counts entries with value > 0."""
    df = np.zeros(num_features, dtype=np.int64)
    for idx, val in rows:
        for i, v in zip(idx, val):
            if v > 0.0:
                df[i] += 1
    m = len(rows)
    return idf_from_doc_freq(df, m, min_doc_freq), df, m


def idf_transform(idx: Sequence[int], val: Sequence[float], idf: np.ndarray) -> List[float]:
    """``newValues[k] = values[k] * idf(indices[k])`` -- one exact fp64 multiply per entry."""
    return [float(v) * float(idf[i]) for i, v in zip(idx, val)]


# ---------------------------------------------------------------------------------------------
# featurise one dialogue end to end (a1..a5)
# ---------------------------------------------------------------------------------------------
def featurize_bytes(raw: bytes, stop_words: Sequence[bytes], num_features: int, idf: Optional[np.ndarray],
                    do_clean: bool = True, binary: bool = False, vocab_index: Optional[dict] = None):
    """raw UTF-8 bytes -> (idx ascending, tf-idf values, n_tokens, n_kept).

    ``stop_words`` must already be case-folded by the caller when caseSensitive=false.
    With ``do_clean=False`` the input is taken as the ``clean_text`` column and only ASCII
    lower-casing is applied before splitting (the Tokenizer's toLowerCase on ASCII text).
    """
    if do_clean:
        b = clean_bytes(raw)
    else:
        b = bytes((c | 0x20) if 0x41 <= c <= 0x5A else c for c in raw)
    toks = java_split_ws(b)
    sw = set(stop_words)
    kept = [t for t in toks if t not in sw]
    idx, tf = count_vectorize(kept, vocab_index, binary) if vocab_index is not None else hashing_tf(kept, num_features, binary)
    val = idf_transform(idx, tf, idf) if idf is not None else tf
    return idx, val, len(toks), len(kept)


# ---------------------------------------------------------------------------------------------
# A.6  classifiers
# ---------------------------------------------------------------------------------------------
def _exp(x: float) -> float:
    """``math.exp`` with Java's overflow behaviour (+Infinity instead of an exception)."""
    try:
        return math.exp(x)
    except OverflowError:
        return math.inf


def lr_predict(idx: Sequence[int], val: Sequence[float], coef_dense: np.ndarray, intercept: float,
               threshold: float = 0.5):
    """Binomial ``LogisticRegressionModel``: sequential ascending-index dot, intercept last.

    reference artefact: stages/4_LogisticRegression_65ef69d6dfd9/{metadata,data}.
    Returns (prediction, [p0, p1], [raw0, raw1]).
    """
    s = 0.0
    for i, v in zip(idx, val):
        c = float(coef_dense[i])
        if c != 0.0:  # BLAS.dot(sparse, sparse) only visits indices present on both sides
            s = s + float(v) * c
    m = s + intercept
    raw = [-m, m]
    p0 = 1.0 / (1.0 + _exp(-raw[0]))
    p1 = 1.0 - p0
    if threshold <= 0.0:
        raw_thr = -math.inf
    elif threshold >= 1.0:
        raw_thr = math.inf
    else:
        raw_thr = math.log(threshold / (1.0 - threshold))
    pred = 1.0 if m > raw_thr else 0.0
    return pred, [p0, p1], raw


class Tree:
    """One decision tree as parallel node arrays (ids as in the model file: pre-order, root 0).

    feature[n] < 0 marks a leaf.  ``leaf_value`` is ``[n_nodes, n_classes]`` class-weight stats for
    classification trees or ``[n_nodes]`` predictions for regression (GBT) trees.
    """

    def __init__(self, feature, threshold, left, right, leaf_value):
        self.feature = np.asarray(feature, dtype=np.int32)
        self.threshold = np.asarray(threshold, dtype=np.float64)
        self.left = np.asarray(left, dtype=np.int32)
        self.right = np.asarray(right, dtype=np.int32)
        self.leaf_value = np.asarray(leaf_value, dtype=np.float64)

    def leaf_of(self, idx: Sequence[int], val: Sequence[float]) -> int:
        """``features(featureIndex) <= threshold`` goes left; absent index reads 0.0."""
        look = dict(zip(idx, val))
        n = 0
        while self.feature[n] >= 0:
            x = look.get(int(self.feature[n]), 0.0)
            n = int(self.left[n]) if x <= self.threshold[n] else int(self.right[n])
        return n


def _argmax_first(v: Sequence[float]) -> int:
    best, bi = -math.inf, 0
    for i, x in enumerate(v):
        if x > best:
            best, bi = x, i
    return bi


def dt_predict(idx, val, tree: Tree):
    """``DecisionTreeClassificationModel``: raw = leaf stats, prob = raw / sum(raw), argmax (first)."""
    raw = [float(x) for x in tree.leaf_value[tree.leaf_of(idx, val)]]
    tot = 0.0
    for x in raw:
        tot += x
    prob = [x / tot for x in raw] if tot != 0.0 else list(raw)
    return float(_argmax_first(raw)), prob, raw


def rf_predict(idx, val, trees: Sequence[Tree]):
    """``RandomForestClassificationModel``: per tree IN ORDER votes[i] += c[i] / sum(c)."""
    k = trees[0].leaf_value.shape[1]
    votes = [0.0] * k
    for t in trees:
        c = t.leaf_value[t.leaf_of(idx, val)]
        tot = 0.0
        for x in c:
            tot += float(x)
        if tot != 0.0:
            for i in range(k):
                votes[i] += float(c[i]) / tot
    s = 0.0
    for x in votes:
        s += x
    prob = [x / s for x in votes] if s != 0.0 else list(votes)
    return float(_argmax_first(votes)), prob, list(votes)


def gbt_predict(idx, val, trees: Sequence[Tree], weights: Sequence[float]):
    """``GBTClassificationModel`` (LogLoss): margin = sum_t pred_t * w_t (sequential), raw = [-m, m],
    p0 = 1/(1+exp(-2*raw0)), p1 = 1 - p0, prediction = argmax(raw) (m <= 0 -> 0.0).

    The reference order of this dot product depends on the netlib BLAS backend Spark picked up at run
    time (f2j / Java8 / FMA / native); the sequential order is used here -- parity unpinned.
    """
    m = 0.0
    for t, w in zip(trees, weights):
        m = m + float(t.leaf_value[t.leaf_of(idx, val)]) * float(w)
    raw = [-m, m]
    p0 = 1.0 / (1.0 + _exp(-2.0 * raw[0]))
    p1 = 1.0 - p0
    return float(_argmax_first(raw)), [p0, p1], raw


# ---------------------------------------------------------------------------------------------
# A.7  RandomForest training  (org.apache.spark.ml.tree.impl.RandomForest)  -- parity unpinned
#      reference call site: fraud_detection_spark.py:59-74, 85-91
# ---------------------------------------------------------------------------------------------
EPSILON = 2.220446049250313e-16  # org.apache.spark.ml.impl.Utils.EPSILON


def find_splits_for_feature(nz_values: Sequence[float], weighted_num_samples: float, num_splits: int) -> List[float]:
    """``RandomForest.findSplitsForContinuousFeature`` for unit instance weights.

    ``nz_values``: the feature's NON-ZERO values among the sampled rows (``findSplitsBySorting``
    filters ``value != 0.0``); the implied zero mass is ``weighted_num_samples - len(nz_values)``
    (``weighted_num_samples`` = sampling fraction x number of examples -- the EXPECTED sample size, it
    equals the row count when the fraction is 1).  ``num_splits`` = ``min(maxBins, numExamples) - 1``.
    Returns ascending thresholds (empty for a constant or never-seen feature).
    """
    if len(nz_values) == 0:
        return []
    counts = {}
    for v in nz_values:
        counts[float(v)] = counts.get(float(v), 0.0) + 1.0
    part = float(len(nz_values))
    tolerance = EPSILON * len(nz_values) * 100
    if weighted_num_samples - part > tolerance:
        counts[0.0] = weighted_num_samples - part
    values = sorted(counts)
    if len(values) <= 1:
        return []
    possible = len(values) - 1
    if possible <= num_splits:
        return [(values[i - 1] + values[i]) / 2.0 for i in range(1, len(values))]
    stride = weighted_num_samples / (num_splits + 1)
    out: List[float] = []
    index = 1
    current = counts[values[0]]
    target = stride
    while index < len(values):
        previous = current
        current += counts[values[index]]
        if abs(previous - target) < abs(current - target):
            out.append((values[index - 1] + values[index]) / 2.0)
            target += stride
        index += 1
    return out


def bin_value(thresholds: Sequence[float], v: float) -> int:
    """``TreePoint.findBin``: Java binarySearch -> exact hit keeps its index, miss -> insertion point."""
    lo, hi = 0, len(thresholds)
    while lo < hi:
        mid = (lo + hi) // 2
        if thresholds[mid] < v:
            lo = mid + 1
        else:
            hi = mid
    return lo


def gini(counts: Sequence[float], total: float) -> float:
    """``Gini.calculate``: 1 - sum (c/total)^2, accumulated as ``impurity -= f*f`` in class order."""
    if total == 0.0:
        return 0.0
    imp = 1.0
    for c in counts:
        f = c / total
        imp -= f * f
    return imp


_MIN_GAIN = -1.7976931348623157e308  # Double.MinValue marks an invalid split


def best_split_for_node(hist: np.ndarray, n_splits: Sequence[int], feature_ids: Sequence[int],
                        min_instances: int = 1, min_info_gain: float = 0.0):
    """``RandomForest.binsToBestSplit`` for one node.

    ``hist[j, b, c]`` = weighted count of class ``c`` in bin ``b`` of the j-th feature of the node's
    subset (subset order = ``feature_ids`` order); ``n_splits[j]`` = number of thresholds of that
    feature.  Unit instance weights, so rawCount == sum of class counts.
    Returns (gain, j_best, split_idx, left_counts, right_counts, parent_counts); j_best = -1 if no
    valid split.  First maximum wins over splits, then over features, as Scala's ``maxBy``.
    """
    n_feat, n_bins, n_cls = hist.shape
    parent = [0.0] * n_cls
    # parent stats: total of any feature's bins (all identical); Spark reads them off the first one
    for b in range(n_bins):
        for c in range(n_cls):
            parent[c] += float(hist[0, b, c])
    ptotal = 0.0
    for c in parent:
        ptotal += c
    pimp = gini(parent, ptotal)
    best = (_MIN_GAIN, -1, -1, None, None)
    for j in range(n_feat):
        ns = int(n_splits[j])
        if ns == 0:
            continue
        cum = [[0.0] * n_cls for _ in range(n_bins)]
        run = [0.0] * n_cls
        for b in range(n_bins):
            for c in range(n_cls):
                run[c] += float(hist[j, b, c])
                cum[b][c] = run[c]
        total_stats = cum[ns]  # bins 0..ns cover everything (values land in bins 0..ns)
        fbest = (_MIN_GAIN, -1, None, None)
        for s in range(ns):
            left = cum[s]
            right = [total_stats[c] - left[c] for c in range(n_cls)]
            lt = 0.0
            for c in left:
                lt += c
            rt = 0.0
            for c in right:
                rt += c
            if lt < min_instances or rt < min_instances:
                g = _MIN_GAIN
            else:
                tc = lt + rt
                g = pimp - (lt / tc) * gini(left, lt) - (rt / tc) * gini(right, rt)
                if g < min_info_gain:
                    g = _MIN_GAIN
            if g > fbest[0]:
                fbest = (g, s, list(left), right)
        if fbest[1] >= 0 and fbest[0] > best[0]:
            best = (fbest[0], j, fbest[1], fbest[2], fbest[3])
    return best + (parent,)


def train_forest(rows_idx: Sequence[Sequence[int]], rows_bin: Sequence[Sequence[int]], labels: Sequence[int],
                 thresholds: Sequence[Sequence[float]], bin_of_zero: Sequence[int], bag_weights: np.ndarray,
                 subset_fn, max_depth: int, n_classes: int = 2, max_bins: int = 32, prune: bool = True):
    """Level-wise RandomForest growth on pre-binned sparse rows, given explicit bag weights and
    feature subsets (the parity contract: identical splits on identical bins / bags / subsets).

    ``rows_idx[i]`` / ``rows_bin[i]``: ascending feature ids of row i and their bins;
    absent features sit in ``bin_of_zero[f]``.  ``bag_weights[t, i]``: integer bootstrap count.
    ``subset_fn(tree, node_id) -> sequence of feature ids`` (node ids heap-style: root 1, children
    2n / 2n+1, as Spark's ``LearningNode`` indices).  Returns a list of :class:`Tree`.
    """
    n_trees = bag_weights.shape[0]
    n = len(labels)
    forest = []
    for t in range(n_trees):
        # heap-id -> dict(node)
        nodes = {1: dict(rows=[i for i in range(n) if bag_weights[t, i] > 0], depth=0)}
        frontier = [1]
        while frontier:
            nxt = []
            for nid in frontier:
                nd = nodes[nid]
                stats = [0.0] * n_classes
                for i in nd["rows"]:
                    stats[labels[i]] += float(bag_weights[t, i])
                nd["stats"] = stats
                tot = 0.0
                for c in stats:
                    tot += c
                nd["impurity"] = gini(stats, tot)
                nd["leaf"] = True
                if nd["depth"] >= max_depth or (nid != 1 and abs(nd["impurity"]) < EPSILON):
                    continue
                feats = list(subset_fn(t, nid))
                hist = np.zeros((len(feats), max_bins, n_classes), dtype=np.float64)
                pos = {f: j for j, f in enumerate(feats)}
                for i in nd["rows"]:
                    w = float(bag_weights[t, i])
                    seen = set()
                    for f, b in zip(rows_idx[i], rows_bin[i]):
                        j = pos.get(f)
                        if j is not None:
                            hist[j, b, labels[i]] += w
                            seen.add(j)
                    for j, f in enumerate(feats):
                        if j not in seen:
                            hist[j, bin_of_zero[f], labels[i]] += w
                ns = [len(thresholds[f]) for f in feats]
                gain, j, s, lc, rc, _ = best_split_for_node(hist, ns, feats)
                if j < 0 or gain <= 0.0:
                    continue
                f = feats[j]
                nd.update(leaf=False, feature=f, split=s, threshold=thresholds[f][s], gain=gain)
                lrows, rrows = [], []
                for i in nd["rows"]:
                    b = bin_of_zero[f]
                    for ff, bb in zip(rows_idx[i], rows_bin[i]):
                        if ff == f:
                            b = bb
                            break
                    (lrows if b <= s else rrows).append(i)
                nodes[2 * nid] = dict(rows=lrows, depth=nd["depth"] + 1)
                nodes[2 * nid + 1] = dict(rows=rrows, depth=nd["depth"] + 1)
                nxt += [2 * nid, 2 * nid + 1]
            frontier = nxt
        forest.append(_heap_to_tree(nodes, n_classes, prune))
    return forest


def _heap_to_tree(nodes: dict, n_classes: int, prune: bool) -> Tree:
    """Convert heap-indexed learning nodes into pre-order arrays (``LearningNode.toNode``), pruning
    internal nodes whose two children are leaves with the same prediction."""

    def build(nid):
        nd = nodes[nid]
        pred = _argmax_first(nd["stats"])
        if nd["leaf"]:
            return dict(leaf=True, stats=nd["stats"], pred=pred)
        l = build(2 * nid)
        r = build(2 * nid + 1)
        if prune and l["leaf"] and r["leaf"] and l["pred"] == r["pred"]:
            return dict(leaf=True, stats=nd["stats"], pred=pred)
        return dict(leaf=False, stats=nd["stats"], pred=pred, feature=nd["feature"], threshold=nd["threshold"], l=l, r=r)

    feature, threshold, left, right, leafv = [], [], [], [], []

    def emit(nd):
        me = len(feature)
        feature.append(-1 if nd["leaf"] else nd["feature"])
        threshold.append(0.0 if nd["leaf"] else nd["threshold"])
        left.append(-1)
        right.append(-1)
        leafv.append(list(nd["stats"]))
        if not nd["leaf"]:
            left[me] = emit(nd["l"])
            right[me] = emit(nd["r"])
        return me

    emit(build(1))
    return Tree(feature, threshold, left, right, np.asarray(leafv, dtype=np.float64).reshape(-1, n_classes))


# ---------------------------------------------------------------------------------------------
# evaluators   (mllib BinaryClassificationMetrics / MulticlassMetrics)   reference: fraud_detection_spark.py:101-119
# ---------------------------------------------------------------------------------------------
def area_under_roc(scores: Sequence[float], labels: Sequence[float], num_bins: int = 1000) -> float:
    """``BinaryClassificationMetrics(scoreAndLabels, numBins).areaUnderROC`` on ONE partition: counts per distinct
    score (descending), optional down-sampling in chunks of ``distinct // numBins``, ROC points with (0,0) and
    (1,1) added, trapezoid rule summed left to right."""
    # scores are keys of a JVM map and are sorted with java.lang.Double.compare: equality and order by bit pattern, so
    # -0.0 and 0.0 are two distinct scores with 0.0 > -0.0
    import struct

    def jkey(v):
        b = struct.unpack("<Q", struct.pack("<d", float(v)))[0]
        return (b ^ 0xFFFFFFFFFFFFFFFF) if b >> 63 else (b | 0x8000000000000000)

    counts = {}
    for s, y in zip(scores, labels):
        c = counts.setdefault(jkey(s), [0.0, 0.0])
        c[0 if y > 0.5 else 1] += 1.0
    items = [(s, counts[s]) for s in sorted(counts, reverse=True)]
    grouping = len(items) // num_bins if num_bins > 0 else 0
    if grouping >= 2:
        binned = []
        for i in range(0, len(items), grouping):
            chunk = items[i:i + grouping]
            binned.append((chunk[-1][0], [sum(c[1][0] for c in chunk), sum(c[1][1] for c in chunk)]))
        items = binned
    tot_p = sum(c[0] for _, c in items)
    tot_n = sum(c[1] for _, c in items)
    pts = [(0.0, 0.0)]
    cp = cn = 0.0
    for _, c in items:
        cp += c[0]
        cn += c[1]
        pts.append((cn / tot_n if tot_n > 0 else 0.0, cp / tot_p if tot_p > 0 else 0.0))
    pts.append((1.0, 1.0))
    area = 0.0
    for (x0, y0), (x1, y1) in zip(pts[:-1], pts[1:]):
        area += (x1 - x0) * (y1 + y0) / 2.0
    return area


def multiclass_metrics(labels: Sequence[float], predictions: Sequence[float]) -> dict:
    """``MulticlassMetrics``: accuracy, weightedPrecision, weightedRecall, weightedFMeasure(beta = 1)."""
    classes = sorted(set(float(y) for y in labels))
    n = float(len(labels))
    out = {"accuracy": sum(1.0 for y, p in zip(labels, predictions) if y == p) / n, "weightedPrecision": 0.0,
           "weightedRecall": 0.0, "f1": 0.0}
    for c in classes:
        tp = sum(1.0 for y, p in zip(labels, predictions) if y == c and p == c)
        fp = sum(1.0 for y, p in zip(labels, predictions) if y != c and p == c)
        cnt = sum(1.0 for y in labels if y == c)
        prec = tp / (tp + fp) if tp + fp > 0 else 0.0
        rec = tp / cnt if cnt > 0 else 0.0
        f = 2.0 * prec * rec / (prec + rec) if prec + rec > 0 else 0.0
        w = cnt / n
        out["weightedPrecision"] += prec * w
        out["weightedRecall"] += rec * w
        out["f1"] += f * w
    return out


# ---------------------------------------------------------------------------------------------
# convenience: a whole pipeline over a batch (used by tests and the golden generator)
# ---------------------------------------------------------------------------------------------
class OraclePipeline:
    """Feature stages + one classifier, evaluated row by row with the functions above."""

    def __init__(self, stop_words: Sequence[str], case_sensitive: bool, num_features: int, binary: bool,
                 idf: Optional[np.ndarray], classifier: dict, vocabulary: Optional[Sequence[str]] = None):
        sw = list(stop_words) if case_sensitive else [w.lower() for w in stop_words]
        self.stop_words = [w.encode("utf-8") for w in sw]
        self.num_features = int(num_features)
        self.binary = bool(binary)
        self.idf = None if idf is None else np.asarray(idf, dtype=np.float64)
        self.classifier = classifier  # {"kind": "lr"|"dt"|"rf"|"gbt", ...}
        self.vocab_index = None if vocabulary is None else {w.encode("utf-8"): i for i, w in enumerate(vocabulary)}

    def featurize(self, raw: bytes, do_clean: bool = True):
        return featurize_bytes(raw, self.stop_words, self.num_features, self.idf, do_clean, self.binary, self.vocab_index)

    def score(self, idx, val):
        c = self.classifier
        k = c["kind"]
        if k == "lr":
            return lr_predict(idx, val, c["coef"], c["intercept"], c.get("threshold", 0.5))
        if k == "dt":
            return dt_predict(idx, val, c["trees"][0])
        if k == "rf":
            return rf_predict(idx, val, c["trees"])
        if k == "gbt":
            return gbt_predict(idx, val, c["trees"], c["weights"])
        raise ValueError(k)

    def transform(self, rows: Sequence[bytes], do_clean: bool = True):
        pred, prob, raw = [], [], []
        for r in rows:
            idx, val, _, _ = self.featurize(r, do_clean)
            p, pr, rw = self.score(idx, val)
            pred.append(p)
            prob.append(pr)
            raw.append(rw)
        return np.asarray(pred), np.asarray(prob), np.asarray(raw)
