"""Evaluators (areaUnderROC, accuracy / weighted precision / recall / f1, crosstab) against the oracle's
restatement of mllib's BinaryClassificationMetrics / MulticlassMetrics, and against sklearn where they coincide."""
import numpy as np
import pytest

from oracle import spark_oracle as so


def _data(n, seed, distinct=None):
    rng = np.random.default_rng(seed)
    y = (rng.random(n) < 0.4).astype(np.float64)
    s = rng.normal(0, 1, n) + 1.2 * y
    if distinct:
        s = np.round(s * distinct) / distinct + 0.0  # (+ 0.0: no negative zeros here, see test_signed_zero_scores_are_two_scores)
    pred = (s > 0.6).astype(np.float64)
    return s, y, pred


@pytest.mark.parametrize("n,distinct", [(300, None), (5000, None), (5000, 3), (40, 2)])
def test_area_under_roc_matches_oracle(fd, n, distinct):
    s, y, pred = _data(n, n, distinct)
    raw = np.stack([-s, s], 1)
    ev = fd.BinaryClassificationEvaluator(labelCol="labels", rawPredictionCol="rawPrediction")
    got = ev.evaluate({"rawPrediction": raw, "labels": y})
    assert abs(got - so.area_under_roc(s, y, 1000)) < 1e-12
    if n <= 1000 or distinct:  # no down-sampling: the exact ROC area
        from sklearn.metrics import roc_auc_score
        assert abs(got - roc_auc_score(y, s)) < 1e-12


def test_signed_zero_scores_are_two_scores(fd):
    """mllib keys its counts by java.lang.Double and sorts with Double.compare: -0.0 and 0.0 are two distinct scores with
    0.0 ranked above -0.0 (sklearn, like IEEE comparison, would tie them).  The oracle and the evaluator restate that."""
    s = np.array([0.0, -0.0, -0.0, 0.0, 1.0, -1.0])
    y = np.array([0.0, 1.0, 1.0, 0.0, 1.0, 0.0])
    got = fd.BinaryClassificationEvaluator(labelCol="y").evaluate({"rawPrediction": s, "y": y})
    assert got == so.area_under_roc(s, y, 1000)
    tied = so.area_under_roc(np.abs(s) * np.sign(s + 0.0), y, 1000)  # the same scores with the zeros merged
    assert got < tied  # the two positives at -0.0 rank below the two negatives at 0.0


def test_degenerate_auc_and_pr(fd):
    ev = fd.BinaryClassificationEvaluator(labelCol="y")
    # no negatives at all: mllib's curve jumps from (0, 1) to (1, 1) and reports 1.0 -- restated, not "fixed"
    one_class = ev.evaluate({"rawPrediction": np.array([[0.0, 1.0], [0.0, 2.0]]), "y": np.array([1.0, 1.0])})
    assert one_class == so.area_under_roc([1.0, 2.0], [1.0, 1.0]) == 1.0
    pr = fd.BinaryClassificationEvaluator(labelCol="y", metricName="areaUnderPR")
    assert 0.0 <= pr.evaluate({"rawPrediction": np.array([0.9, 0.1, 0.8, 0.3]), "y": np.array([1.0, 0.0, 1.0, 0.0])}) <= 1.0


def test_multiclass_metrics_and_crosstab(fd):
    s, y, pred = _data(2000, 5)
    exp = so.multiclass_metrics(y.tolist(), pred.tolist())
    ev = fd.MulticlassClassificationEvaluator(labelCol="labels")
    cols = {"prediction": pred, "labels": y}
    for name in ("accuracy", "weightedPrecision", "weightedRecall", "f1"):
        assert abs(ev.setMetricName(name).evaluate(cols) - exp[name]) < 1e-12, name
    cm = fd.confusion_matrix(y, pred)
    assert cm["labels"].tolist() == [0.0, 1.0] and cm["matrix"].sum() == 2000
    assert cm["matrix"][1, 1] == int(((y == 1) & (pred == 1)).sum()) and cm["matrix"][0, 1] == int(((y == 0) & (pred == 1)).sum())
    # a class that is predicted but never true carries no weight
    m = so.multiclass_metrics([0.0, 0.0, 1.0], [0.0, 2.0, 1.0])
    got = fd.MulticlassClassificationEvaluator(labelCol="l", metricName="weightedPrecision").evaluate({"prediction": np.array([0.0, 2.0, 1.0]), "l": np.array([0.0, 0.0, 1.0])})
    assert abs(got - m["weightedPrecision"]) < 1e-12


@pytest.mark.gpu
def test_evaluators_on_device_columns(fd):
    import torch
    s, y, pred = _data(20000, 9)
    raw = torch.from_numpy(np.stack([-s, s], 1)).cuda()
    got = fd.BinaryClassificationEvaluator(labelCol="labels").evaluate({"rawPrediction": raw, "labels": torch.from_numpy(y).cuda()})
    assert abs(got - so.area_under_roc(s, y, 1000)) < 1e-12
    acc = fd.MulticlassClassificationEvaluator(labelCol="labels", metricName="accuracy").evaluate(
        {"prediction": torch.from_numpy(pred).cuda(), "labels": torch.from_numpy(y).cuda()})
    assert abs(acc - so.multiclass_metrics(y.tolist(), pred.tolist())["accuracy"]) < 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("n,distinct", [(300, None), (50000, None), (50000, 3), (40, 2), (1, None)])
def test_device_kernels_against_the_oracle(fd, n, distinct):
    """fd_eval_auc_roc / fd_eval_confusion2 (through the evaluator classes, columns on the GPU): down-sampled and exact
    curves, heavily tied scores, a single row, a strided score column (rawPrediction[:, 1]), a class that never occurs."""
    import torch
    s, y, pred = _data(n, 100 + n, distinct)
    raw = torch.from_numpy(np.stack([-s, s], 1)).cuda()
    ev = fd.BinaryClassificationEvaluator(labelCol="labels")
    got = ev.evaluate({"rawPrediction": raw, "labels": torch.from_numpy(y).cuda()})
    assert abs(got - so.area_under_roc(s, y, 1000)) < 1e-12
    ones = torch.ones(n, dtype=torch.float64, device="cuda")
    assert ev.evaluate({"rawPrediction": raw, "labels": ones}) == so.area_under_roc(s, np.ones(n), 1000)
    exp = so.multiclass_metrics(y.tolist(), pred.tolist())
    cols = {"prediction": torch.from_numpy(pred).cuda(), "labels": torch.from_numpy(y).cuda()}
    mc = fd.MulticlassClassificationEvaluator(labelCol="labels")
    for name in ("accuracy", "weightedPrecision", "weightedRecall", "f1"):
        assert abs(mc.setMetricName(name).evaluate(cols) - exp[name]) < 1e-12, name
    cm = fd.confusion_matrix(cols["labels"], cols["prediction"])
    host = fd.confusion_matrix(y, pred)
    assert cm["labels"].tolist() == host["labels"].tolist() and np.array_equal(cm["matrix"], host["matrix"])
    # values outside {0, 1}: the general tabulation takes over
    three = fd.confusion_matrix(torch.tensor([0.0, 2.0, 1.0]).cuda(), torch.tensor([0.0, 2.0, 2.0]).cuda())
    assert three["labels"].tolist() == [0.0, 1.0, 2.0] and three["matrix"].sum() == 3
