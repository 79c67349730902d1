"""Evaluators the reference applies to the transform output (``fraud_detection_spark.py:101-119``):
``BinaryClassificationEvaluator`` (areaUnderROC on ``rawPrediction[1]``), ``MulticlassClassificationEvaluator``
(accuracy, weightedPrecision, weightedRecall, f1) and ``crosstab(labels, prediction)``.

Spark semantics restated (mllib ``BinaryClassificationMetrics`` with numBins = 1000 on a single partition,
``MulticlassMetrics``).  Columns on a CUDA device go through the library's kernels (``fd_eval_auc_roc``: radix sort of the
scores, group heads, cumulative positives, chunk points, trapezoid sum; ``fd_eval_confusion2``: the four counts of a binary
problem), so they never leave the device; host columns, areaUnderPR and labels outside {0, 1} take the tensor-op
restatement below, which is also what the kernels are tested against."""
from __future__ import annotations

from typing import Dict

import numpy as np


def _t(x, like=None):
    import torch
    if torch.is_tensor(x):
        return x
    return torch.as_tensor(np.asarray(x), device=None if like is None else like.device)


def _col(data, name):
    return data[name] if not hasattr(data, "_cols") else data[name]


_HANDLES = {}


def _handle(device_index: int):
    """One library handle per device for the evaluators (they need no model)."""
    from . import _lib
    if device_index not in _HANDLES:
        _HANDLES[device_index] = _lib.Handle(device_index)
    return _HANDLES[device_index]


def _stream(t):
    import ctypes as C
    import torch
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


class BinaryClassificationEvaluator:
    """``pyspark.ml.evaluation.BinaryClassificationEvaluator`` (metricName areaUnderROC / areaUnderPR)."""

    def __init__(self, labelCol: str = "label", rawPredictionCol: str = "rawPrediction", metricName: str = "areaUnderROC",
                 numBins: int = 1000):
        self.labelCol, self.rawPredictionCol, self.metricName, self.numBins = labelCol, rawPredictionCol, metricName, int(numBins)

    def evaluate(self, predictions, labels=None) -> float:
        """``predictions``: TransformResult / dict holding ``rawPredictionCol`` ([N, 2] or [N]); ``labels``: the label
        column (or taken from ``predictions[labelCol]``)."""
        import torch
        raw = _t(_col(predictions, self.rawPredictionCol))
        score = raw[:, 1] if raw.ndim == 2 else raw
        y = _t(labels if labels is not None else _col(predictions, self.labelCol), like=score).to(score.device).to(torch.float64)
        if score.is_cuda and score.dtype == torch.float64 and self.metricName == "areaUnderROC":
            from . import _lib
            y = y.contiguous()
            out = torch.zeros(3, dtype=torch.float64, device=score.device)
            _handle(score.device.index).call("fd_eval_auc_roc", _lib.ptr(score), int(score.stride(0)) if score.numel() else 1, _lib.ptr(y),
                                             int(score.numel()), int(self.numBins), _lib.ptr(out), _stream(score))
            area, tp, tn = (float(v) for v in out.cpu().tolist())
            x_last, y_last = (1.0 if tn > 0 else 0.0), (1.0 if tp > 0 else 0.0)
            return area + (1.0 - x_last) * (1.0 + y_last) / 2.0  # the closing segment to (1, 1)
        # scores compare as java.lang.Double does (by bit pattern: 0.0 > -0.0, two distinct scores): order-preserving int64 keys
        bits = score.to(torch.float64).contiguous().view(torch.int64)
        score = torch.where(bits < 0, ~bits ^ torch.iinfo(torch.int64).min, bits)
        order = torch.argsort(score, descending=True, stable=True)
        s, y = score[order], y[order]
        # counts per distinct score, in descending score order
        _, inv, _ = torch.unique_consecutive(s, return_inverse=True, return_counts=True)
        n_distinct = int(inv[-1].item()) + 1 if s.numel() else 0
        pos = torch.zeros(n_distinct, dtype=torch.float64, device=s.device).index_add_(0, inv, (y > 0.5).to(torch.float64))
        neg = torch.zeros(n_distinct, dtype=torch.float64, device=s.device).index_add_(0, inv, (y <= 0.5).to(torch.float64))
        # down-sampling to about numBins points: chunks of `grouping` consecutive distinct scores
        if self.numBins > 0 and n_distinct // self.numBins >= 2:
            grouping = n_distinct // self.numBins
            gid = torch.arange(n_distinct, device=s.device) // grouping
            n_groups = int(gid[-1].item()) + 1
            pos = torch.zeros(n_groups, dtype=torch.float64, device=s.device).index_add_(0, gid, pos)
            neg = torch.zeros(n_groups, dtype=torch.float64, device=s.device).index_add_(0, gid, neg)
        cp, cn = torch.cumsum(pos, 0), torch.cumsum(neg, 0)
        tp_total, fp_total = float(cp[-1].item()) if cp.numel() else 0.0, float(cn[-1].item()) if cn.numel() else 0.0
        z, one = torch.zeros(1, dtype=torch.float64, device=s.device), torch.ones(1, dtype=torch.float64, device=s.device)
        recall = cp / tp_total if tp_total > 0 else torch.zeros_like(cp)
        if self.metricName == "areaUnderROC":
            fpr = cn / fp_total if fp_total > 0 else torch.zeros_like(cn)
            x = torch.cat([z, fpr, one])
            yv = torch.cat([z, recall, one])
        elif self.metricName == "areaUnderPR":
            prec = torch.where(cp + cn > 0, cp / (cp + cn), torch.ones_like(cp))
            first = prec[:1] if prec.numel() else one
            x = torch.cat([z, recall])
            yv = torch.cat([first, prec])
        else:
            raise ValueError(f"unsupported metricName '{self.metricName}'")
        return float(((x[1:] - x[:-1]) * (yv[1:] + yv[:-1]) / 2.0).sum().item())


def confusion_matrix(labels, prediction) -> Dict[str, np.ndarray]:
    """Counts by (true label, prediction): ``{"labels": sorted distinct values, "matrix": int64 [L, L]}`` --
    what ``preds.crosstab("labels", "prediction")`` (fraud_detection_spark.py:118) tabulates."""
    import torch
    y = _t(labels)
    p = _t(prediction, like=y).to(y.device)
    if y.is_cuda and y.numel():
        from . import _lib
        yd, pd = y.to(torch.float64).contiguous(), p.to(torch.float64).contiguous()
        counts = torch.zeros(5, dtype=torch.int64, device=y.device)
        _handle(y.device.index).call("fd_eval_confusion2", _lib.ptr(yd), _lib.ptr(pd), int(yd.numel()), _lib.ptr(counts), _stream(yd))
        c = counts.cpu().numpy()
        if c[4] == 0:  # binary labels and predictions: the kernel's four counts are the whole table
            full = c[:4].reshape(2, 2)
            seen = [v for v in (0, 1) if full[v, :].sum() + full[:, v].sum() > 0]
            return {"labels": np.asarray(seen, dtype=np.float64), "matrix": full[np.ix_(seen, seen)].astype(np.int64)}
    vals = torch.unique(torch.cat([y.to(torch.float64), p.to(torch.float64)]))
    yi = torch.searchsorted(vals, y.to(torch.float64))
    pi = torch.searchsorted(vals, p.to(torch.float64))
    L = vals.numel()
    m = torch.bincount(yi * L + pi, minlength=L * L).view(L, L)
    return {"labels": vals.cpu().numpy(), "matrix": m.cpu().numpy()}


class MulticlassClassificationEvaluator:
    """``pyspark.ml.evaluation.MulticlassClassificationEvaluator``: f1 (default), accuracy, weightedPrecision,
    weightedRecall -- the four the reference reads (fraud_detection_spark.py:108-114)."""

    def __init__(self, labelCol: str = "label", predictionCol: str = "prediction", metricName: str = "f1"):
        self.labelCol, self.predictionCol, self.metricName = labelCol, predictionCol, metricName

    def setMetricName(self, name: str):
        self.metricName = name
        return self

    def evaluate(self, predictions, labels=None) -> float:
        cm = confusion_matrix(labels if labels is not None else _col(predictions, self.labelCol), _col(predictions, self.predictionCol))
        m = cm["matrix"].astype(np.float64)
        total = m.sum()
        label_count = m.sum(axis=1)            # by true label
        pred_count = m.sum(axis=0)
        tp = np.diag(m)
        present = label_count > 0              # MulticlassMetrics.labels = the TRUE labels
        precision = np.where(pred_count > 0, tp / np.where(pred_count > 0, pred_count, 1.0), 0.0)
        recall = np.where(label_count > 0, tp / np.where(label_count > 0, label_count, 1.0), 0.0)
        f1 = np.where(precision + recall > 0, 2.0 * precision * recall / np.where(precision + recall > 0, precision + recall, 1.0), 0.0)
        w = np.where(present, label_count / total, 0.0)
        if self.metricName == "accuracy":
            return float(tp.sum() / total)
        if self.metricName == "weightedPrecision":
            return float((precision * w).sum())
        if self.metricName == "weightedRecall":
            return float((recall * w).sum())
        if self.metricName == "f1":
            return float((f1 * w).sum())
        raise ValueError(f"unsupported metricName '{self.metricName}'")
