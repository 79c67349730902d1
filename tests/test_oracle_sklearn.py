"""Independent cross-checks of the oracle's tree rules against scikit-learn's CART (no Spark-written tree
exists in the reference and Spark cannot run here, so sklearn is the only second implementation available offline).

What sklearn CAN pin: the Gini gain of every split Spark's rules choose, the row partition of each split when no two
candidate splits tie, the predict walk (left iff x <= threshold, absent feature = 0.0), class probabilities of a
single tree (leaf counts / total) and of a forest (mean of the trees' leaf distributions).  What it cannot: first-max
tie-breaking, the pruning of equal-prediction siblings (switched off here), GBT's margin order, the Parquet layout."""
import importlib

import numpy as np
import pytest

from oracle import spark_oracle as so

training = importlib.import_module("fraud-detection-spark-kafka-llm_b200.training")


def _dense_rows(n, f, seed):
    """Dense integer-valued rows (values 0..4, so every midpoint between distinct values is a threshold) with labels that
    depend on several features plus noise, which keeps gains distinct at most nodes."""
    rng = np.random.default_rng(seed)
    x = rng.integers(0, 5, size=(n, f)).astype(np.float64)
    score = 1.3 * (x[:, 0] > 2) + 0.9 * (x[:, 1] <= 1) + 0.7 * (x[:, 2] > 0) * (x[:, 3] > 3) + 0.31 * x[:, 4] + rng.normal(0, 0.8, n)
    y = (score > np.median(score)).astype(np.int64)
    return x, y


def _sparse(x):
    rows_idx = [np.nonzero(r)[0].tolist() for r in x]
    rows_val = [r[np.nonzero(r)[0]].tolist() for r in x]
    return rows_idx, rows_val


def _oracle_tree(x, y, depth, prune=False):
    n, f = x.shape
    rows_idx, rows_val = _sparse(x)
    feat = np.asarray([j for r in rows_idx for j in r], dtype=np.int64)
    val = np.asarray([v for r in rows_val for v in r], dtype=np.float64)
    thr_ptr, thr, n_splits, boz = training.find_splits(feat, val, f, float(n), 31)
    tl = [thr[thr_ptr[j]:thr_ptr[j + 1]].tolist() for j in range(f)]
    rows_bin = [[so.bin_value(tl[j], v) for j, v in zip(ix, vl)] for ix, vl in zip(rows_idx, rows_val)]
    bag = np.ones((1, n), dtype=np.int64)
    return so.train_forest(rows_idx, rows_bin, y.tolist(), tl, boz.tolist(), bag, lambda t, nid: range(f), depth, prune=prune)[0]


def _gain(y, left):
    def g(v):
        if len(v) == 0:
            return 0.0
        p = np.bincount(v, minlength=2) / len(v)
        return 1.0 - float((p * p).sum())
    n = len(y)
    return g(y) - left.sum() / n * g(y[left]) - (~left).sum() / n * g(y[~left])


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
@pytest.mark.parametrize("depth", [3, 5])
def test_every_split_of_the_oracle_tree_is_a_cart_optimum(seed, depth):
    """Walk the oracle's tree and sklearn's tree together from the root.  At every node both must reach the same Gini gain
    (both maximise it over the same candidate thresholds); where they also cut the rows the same way the walk continues
    into both children, where they tie on gain but cut differently (tie-breaking is not comparable) the subtree is left.
    Most of the tree must be walked, and no node may show a gain difference."""
    from sklearn.tree import DecisionTreeClassifier
    x, y = _dense_rows(700, 9, seed)
    tree = _oracle_tree(x, y, depth)
    sk = DecisionTreeClassifier(criterion="gini", max_depth=depth, random_state=0).fit(x, y).tree_
    compared = ties = 0
    stack = [(0, 0, np.ones(len(y), dtype=bool))]
    while stack:
        on, sn, rows = stack.pop()
        o_leaf, s_leaf = tree.feature[on] < 0, sk.children_left[sn] < 0
        yy = y[rows]
        if o_leaf or s_leaf:
            # Spark stops at gain <= 0; sklearn may still split a node with zero improvement
            if o_leaf and not s_leaf:
                assert _gain(yy, x[rows][:, sk.feature[sn]] <= sk.threshold[sn]) < 1e-12
            if s_leaf and not o_leaf:
                pytest.fail("sklearn stopped where the oracle found a positive gain")
            continue
        lo = x[rows][:, int(tree.feature[on])] <= float(tree.threshold[on])
        ls = x[rows][:, sk.feature[sn]] <= sk.threshold[sn]
        assert abs(_gain(yy, lo) - _gain(yy, ls)) < 1e-12, (on, sn)
        compared += 1
        if np.array_equal(lo, ls):
            idx = np.nonzero(rows)[0]
            left_rows, right_rows = np.zeros_like(rows), np.zeros_like(rows)
            left_rows[idx[lo]] = True
            right_rows[idx[~lo]] = True
            stack.append((int(tree.left[on]), int(sk.children_left[sn]), left_rows))
            stack.append((int(tree.right[on]), int(sk.children_right[sn]), right_rows))
        else:
            ties += 1
    assert compared >= (5 if depth == 3 else 10) and ties <= compared // 3, (compared, ties)


def _tree_from_sklearn(sk):
    """sklearn's arrays as an oracle Tree: same node numbering (sklearn is pre-order too), leaf value = class counts."""
    counts = sk.value[:, 0, :] * sk.weighted_n_node_samples[:, None]
    if np.allclose(sk.value[:, 0, :].sum(axis=1), sk.weighted_n_node_samples):  # older sklearn stores counts, not fractions
        counts = sk.value[:, 0, :]
    return so.Tree(np.where(sk.children_left < 0, -1, sk.feature), np.where(sk.children_left < 0, 0.0, sk.threshold),
                   sk.children_left, sk.children_right, np.round(counts))


@pytest.mark.parametrize("seed", [0, 1])
def test_predict_rules_against_sklearn(seed):
    """dt_predict / rf_predict on sklearn-grown trees must give sklearn's own predict_proba: pins the walk (x <= threshold
    goes left, an absent sparse feature reads 0.0), probability = leaf counts / total, and the forest's mean of the trees'
    leaf distributions (Spark sums c_i / total per tree and normalises: the same number up to rounding)."""
    from sklearn.ensemble import RandomForestClassifier
    from sklearn.tree import DecisionTreeClassifier
    x, y = _dense_rows(500, 8, seed)
    xt, _ = _dense_rows(300, 8, seed + 50)
    rows_idx, rows_val = _sparse(xt)
    dt = DecisionTreeClassifier(criterion="gini", max_depth=6, random_state=seed).fit(x, y)
    t = _tree_from_sklearn(dt.tree_)
    exp = dt.predict_proba(xt)
    for i, (ix, vl) in enumerate(zip(rows_idx, rows_val)):
        pred, prob, raw = so.dt_predict(ix, vl, t)
        assert np.allclose(prob, exp[i], rtol=1e-12, atol=1e-15)
        if abs(exp[i, 0] - exp[i, 1]) > 1e-12:
            assert pred == float(np.argmax(exp[i]))
        else:
            assert pred == 0.0  # first maximum
    rf = RandomForestClassifier(n_estimators=15, max_depth=5, bootstrap=True, random_state=seed).fit(x, y)
    trees = [_tree_from_sklearn(e.tree_) for e in rf.estimators_]
    exp = rf.predict_proba(xt)
    for i, (ix, vl) in enumerate(zip(rows_idx, rows_val)):
        _, prob, raw = so.rf_predict(ix, vl, trees)
        assert np.allclose(prob, exp[i], rtol=1e-12, atol=1e-15)
        assert abs(sum(raw) - len(trees)) < 1e-9  # every tree casts one normalised vote
