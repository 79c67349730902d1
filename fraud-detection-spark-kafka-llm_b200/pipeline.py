"""``Pipeline`` / ``PipelineModel`` with the pyspark.ml surface the reference uses, executing on one B200.

Reference: ``PipelineModel.load(model_path)`` (utils/agent_api.py:129), ``model.transform(df)``
(utils/agent_api.py:158, fraud_detection_spark.py:98), ``model.stages[i]`` (fraud_detection_spark.py:227-241),
``Pipeline(stages=...).fit(df)`` (fraud_detection_spark.py:85-91), ``model.save(path)`` (:393) and the result
access ``.select(col).first()[0]`` / ``probability[1]`` (utils/agent_api.py:161-167).

Host code only decides WHAT to run; every per-dialogue byte and flop runs in csrc/libfdb200.so through the
ctypes C-ABI (``_lib``), with torch tensors as device buffers.  No CPU fallback exists.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
from typing import Any, Dict, List, Optional, Sequence

import numpy as np

from . import _lib
from . import spark_format as sf
from . import stages as S
from .packing import PackedText


def _torch():
    import torch
    return torch


def _stream_ptr(device):
    torch = _torch()
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class Row(tuple):
    """Result row: positional like pyspark's Row, with the column names kept."""

    def __new__(cls, names, values):
        r = super().__new__(cls, values)
        r.__fields__ = tuple(names)
        return r

    def __getattr__(self, item):
        try:
            return self[self.__fields__.index(item)]
        except ValueError:
            raise AttributeError(item)

    def asDict(self):
        return dict(zip(self.__fields__, self))


class TransformResult:
    """Columns produced by ``PipelineModel.transform`` (device tensors), with the few DataFrame verbs the
    reference calls on the result: ``select``, ``first``, ``collect``, ``count``, ``toPandas``."""

    def __init__(self, columns: Dict[str, Any], n_rows: int):
        self._cols = columns
        self._n = int(n_rows)

    @property
    def columns(self) -> List[str]:
        return list(self._cols)

    def __getitem__(self, name: str):
        return self._cols[name]

    def __contains__(self, name):
        return name in self._cols

    def count(self) -> int:
        return self._n

    def select(self, *names) -> "TransformResult":
        if len(names) == 1 and isinstance(names[0], (list, tuple)):
            names = tuple(names[0])
        missing = [n for n in names if n not in self._cols]
        if missing:
            raise KeyError(f"cannot resolve column(s) {missing} among {self.columns}")
        return TransformResult({n: self._cols[n] for n in names}, self._n)

    def _host(self, name):
        v = self._cols[name]
        if hasattr(v, "detach"):
            return v.detach().cpu().numpy()
        return np.asarray(v)

    def _row(self, host, i):
        vals = []
        for name in self._cols:
            a = host[name][i]
            vals.append(float(a) if np.ndim(a) == 0 and isinstance(a, (np.floating, float)) else a)
        return Row(list(self._cols), vals)

    def first(self) -> Optional[Row]:
        if self._n == 0:
            return None
        host = {}
        for name, v in self._cols.items():
            part = v[:1]
            host[name] = part.detach().cpu().numpy() if hasattr(part, "detach") else np.asarray(part)
        return self._row(host, 0)

    def collect(self) -> List[Row]:
        host = {n: self._host(n) for n in self._cols}
        return [self._row(host, i) for i in range(self._n)]

    def toPandas(self):
        import pandas as pd
        host = {n: self._host(n) for n in self._cols}
        return pd.DataFrame({n: (list(a) if a.ndim > 1 else a) for n, a in host.items()})


class _Plan:
    """What a list of stages means for the kernels, after checking the column wiring."""

    def __init__(self, stages: Sequence[S.Stage]):
        self.tokenizer = self.stop = self.hashing = self.idf = self.classifier = None
        col = None
        for st in stages:
            if isinstance(st, S.Tokenizer):
                if self.tokenizer or self.hashing:
                    raise ValueError("unexpected Tokenizer position")
                self.tokenizer = st
                self.text_col = st.getOrDefault("inputCol")
            elif isinstance(st, S.StopWordsRemover):
                if not self.tokenizer or self.hashing:
                    raise ValueError("StopWordsRemover must follow the Tokenizer")
                self._wire(st, col)
                self.stop = st
            elif isinstance(st, (S.HashingTF, S.CountVectorizerModel)):
                if not self.tokenizer or self.hashing:
                    raise ValueError(f"{type(st).__name__} must follow the Tokenizer")
                self._wire(st, col)
                self.hashing = st  # the term -> index stage: hashing trick or vocabulary lookup
            elif isinstance(st, S.IDFModel):
                if not self.hashing or self.idf:
                    raise ValueError("IDFModel must follow HashingTF")
                self._wire(st, col)
                self.idf = st
            elif isinstance(st, S.ClassificationModel):
                if not self.hashing or self.classifier:
                    raise ValueError("the classifier must be the last stage after HashingTF")
                if st.getOrDefault("featuresCol") != col:
                    raise ValueError(f"classifier featuresCol '{st.getOrDefault('featuresCol')}' is not produced (have '{col}')")
                self.classifier = st
                continue
            else:
                raise ValueError(f"stage {type(st).__name__} is not supported on the B200 path")
            col = st.getOrDefault("outputCol")
        if not self.hashing:
            raise ValueError("pipeline needs Tokenizer -> [StopWordsRemover] -> HashingTF | CountVectorizerModel")
        self.features_col = col
        nf = self.num_features
        if self.idf is not None and self.idf.idf.size != nf:
            raise ValueError("IDF vector size differs from HashingTF.numFeatures")
        if self.classifier is not None and self.classifier.numFeatures not in (0, nf):
            raise ValueError("classifier numFeatures differs from HashingTF.numFeatures")

    @property
    def num_features(self) -> int:
        if isinstance(self.hashing, S.CountVectorizerModel):
            return len(self.hashing.vocabulary)
        return int(self.hashing.getOrDefault("numFeatures"))

    @staticmethod
    def _wire(st, col):
        if st.getOrDefault("inputCol") != col:
            raise ValueError(f"{type(st).__name__}.inputCol '{st.getOrDefault('inputCol')}' does not match upstream column '{col}'")


class PipelineModel:
    """Fitted pipeline.  ``transform`` returns ``prediction`` (f64[N]), ``probability`` (f64[N,2]) and
    ``rawPrediction`` (f64[N,2]) exactly as the saved Spark stages would."""

    spark_class = "org.apache.spark.ml.PipelineModel"

    def __init__(self, stages: Sequence[S.Stage], uid: Optional[str] = None):
        self.stages = list(stages)
        self.uid = uid or S._uid("PipelineModel")
        self._plan = _Plan(self.stages)
        self._handles: Dict[int, _lib.Handle] = {}

    # ---------------------------------------------------------------- persistence
    @classmethod
    def load(cls, path: str) -> "PipelineModel":
        meta = sf.read_metadata(path)
        if meta["class"] != cls.spark_class:
            raise ValueError(f"{path} holds a {meta['class']}, not a PipelineModel")
        uids = meta["paramMap"]["stageUids"]
        stages = [S.load_stage(os.path.join(path, "stages", sf.stage_dir_name(i, len(uids), u))) for i, u in enumerate(uids)]
        return cls(stages, uid=meta["uid"])

    def save(self, path: str, overwrite: bool = False) -> None:
        if os.path.exists(path):
            if not overwrite:
                raise FileExistsError(f"{path} exists (pass overwrite=True, as `write().overwrite()`)")
            shutil.rmtree(path)
        os.makedirs(path)
        sf.write_metadata(path, self.spark_class, self.uid, {"stageUids": [s.uid for s in self.stages]}, {})
        for i, s in enumerate(self.stages):
            s.save(os.path.join(path, "stages", sf.stage_dir_name(i, len(self.stages), s.uid)))

    def write(self):
        model = self

        class _W:
            def overwrite(self_w):
                self_w._ow = True
                return self_w

            def save(self_w, path):
                model.save(path, overwrite=getattr(self_w, "_ow", False))

        return _W()

    # ---------------------------------------------------------------- device model
    def handle(self, device: Optional[int] = None) -> _lib.Handle:
        torch = _torch()
        if not torch.cuda.is_available():
            raise RuntimeError("no CUDA device: this pipeline only runs on a GPU (there is no CPU fallback)")
        if device is None:
            device = torch.cuda.current_device()
        device = int(device)
        if device in self._handles:
            return self._handles[device]
        h = _lib.Handle(device)
        p = self._plan
        if p.stop is not None:
            cs = bool(p.stop.getOrDefault("caseSensitive"))
            words = [w if cs else w.lower() for w in p.stop.getOrDefault("stopWords")]
            enc = [w.encode("utf-8") for w in words]
            offs = np.zeros(len(enc) + 1, dtype=np.int32)
            np.cumsum([len(e) for e in enc], out=offs[1:])
            blob = np.frombuffer(b"".join(enc) + b"\0", dtype=np.uint8).copy()
            h.call("fd_set_stopwords", _lib.ptr(blob), _lib.ptr(offs), len(enc), int(cs))
        nf = p.num_features
        if isinstance(p.hashing, S.CountVectorizerModel):
            if float(p.hashing.getOrDefault("minTF")) != 1.0:
                raise ValueError("CountVectorizerModel: only minTF = 1.0 is implemented on the device path")
            enc = [w.encode("utf-8") for w in p.hashing.vocabulary]
            offs = np.zeros(len(enc) + 1, dtype=np.int32)
            np.cumsum([len(e) for e in enc], out=offs[1:])
            blob = np.frombuffer(b"".join(enc) + b"\0", dtype=np.uint8).copy()
            h.call("fd_set_vocabulary", _lib.ptr(blob), _lib.ptr(offs), len(enc), int(bool(p.hashing.getOrDefault("binary"))))
        else:
            h.call("fd_set_hashing", nf, int(bool(p.hashing.getOrDefault("binary"))))
        if p.idf is not None:
            h.call("fd_set_idf", _lib.ptr(p.idf.idf), nf)
        c = p.classifier
        if isinstance(c, S.LogisticRegressionModel):
            h.call("fd_set_linear", _lib.ptr(c.coefficients), nf, C.c_double(c.intercept), C.c_double(float(c.getOrDefault("threshold"))))
        elif isinstance(c, S._TreeEnsembleModel):
            kind = {"dt": _lib.FD_KIND_DT, "rf": _lib.FD_KIND_RF, "gbt": _lib.FD_KIND_GBT}[c.kind]
            offs, feat, thr, left, right, leaf = c._forest_arrays()
            tw = np.ascontiguousarray(c.treeWeights, dtype=np.float64)
            h.call("fd_set_forest", kind, len(c.trees), _lib.ptr(offs), _lib.ptr(feat), _lib.ptr(thr), _lib.ptr(left),
                   _lib.ptr(right), _lib.ptr(leaf), _lib.ptr(tw), int(c.numClasses))
        self._handles[device] = h
        return h

    # ---------------------------------------------------------------- input handling
    def _packed(self, batch, cleaned: Optional[bool]):
        """-> (PackedText, flags).  A list/Series of strings is taken as raw ``dialogue`` text unless
        ``cleaned=True``; a DataFrame/dict is read from the Tokenizer's inputCol (already cleaned, as in the
        reference) or, failing that, from a ``dialogue`` column (cleaning fused into the kernel)."""
        text_col = self._plan.text_col
        if isinstance(batch, PackedText):
            pk = batch
            cleaned = False if cleaned is None else cleaned
        else:
            if hasattr(batch, "columns") and hasattr(batch, "__getitem__") and not isinstance(batch, dict):  # pandas DataFrame
                cols = list(batch.columns)
                batch = {c: batch[c].tolist() for c in cols if c in (text_col, "dialogue")}
            if isinstance(batch, dict):
                if text_col in batch:
                    texts, cleaned = batch[text_col], (True if cleaned is None else cleaned)
                elif "dialogue" in batch:
                    texts, cleaned = batch["dialogue"], (False if cleaned is None else cleaned)
                else:
                    raise KeyError(f"input has neither '{text_col}' nor 'dialogue' column")
            else:
                if isinstance(batch, (str, bytes)):
                    batch = [batch]
                texts = batch.tolist() if hasattr(batch, "tolist") else list(batch)
                cleaned = False if cleaned is None else cleaned
            if cleaned:
                # Tokenizer applies String.toLowerCase to the whole string: the kernel folds ASCII, non-ASCII
                # letters (rare after the reference's cleaning, which leaves only [a-z ]) are folded here
                texts = [t.lower() if isinstance(t, str) and not t.isascii() else t for t in texts]
            pk = PackedText.from_strings(texts)
        return pk, (0 if cleaned else _lib.FD_FLAG_CLEAN)

    # ---------------------------------------------------------------- compute
    def transform(self, batch, cleaned: Optional[bool] = None, device: Optional[int] = None, output: str = "device") -> TransformResult:
        """Run the whole pipeline.  ``output="device"`` returns torch CUDA tensors; ``output="host"`` goes
        through ``fd_transform_host`` (chunked, H2D / kernel / D2H overlapped) and returns numpy arrays."""
        if self._plan.classifier is None:
            raise ValueError("pipeline has no classifier stage; use featurize()")
        torch = _torch()
        h = self.handle(device)
        pk, flags = self._packed(batch, cleaned)
        n = pk.n_rows
        names = [self._plan.classifier.getOrDefault(k) for k in ("predictionCol", "probabilityCol", "rawPredictionCol")]
        if output == "host" and not pk.is_device:
            data = pk.data if isinstance(pk.data, np.ndarray) else pk.data.numpy()
            offs = pk.offs if isinstance(pk.offs, np.ndarray) else pk.offs.numpy()
            if torch.is_tensor(pk.data) and pk.data.is_pinned() and torch.is_tensor(pk.offs) and pk.offs.is_pinned():
                # pinned input: pinned outputs too (from torch's caching host allocator), so that a micro-batch is read
                # and written in place by the kernels
                pred, prob, raw = (torch.empty(s_, dtype=torch.float64, pin_memory=True).numpy() for s_ in ((n,), (n, 2), (n, 2)))
            else:
                pred = np.empty(n, dtype=np.float64)
                prob = np.empty((n, 2), dtype=np.float64)
                raw = np.empty((n, 2), dtype=np.float64)
            h.call("fd_transform_host", _lib.ptr(data), _lib.ptr(offs), n, flags, _lib.ptr(pred), _lib.ptr(prob), _lib.ptr(raw))
            return TransformResult(dict(zip(names, (pred, prob, raw))), n)
        dev = torch.device("cuda", h.device)
        pk = pk.to_device(dev)
        pred = torch.empty(n, dtype=torch.float64, device=dev)
        prob = torch.empty((n, 2), dtype=torch.float64, device=dev)
        raw = torch.empty((n, 2), dtype=torch.float64, device=dev)
        h.call("fd_transform", _lib.ptr(pk.data), _lib.ptr(pk.offs), n, flags, _lib.ptr(pred), _lib.ptr(prob), _lib.ptr(raw),
               _stream_ptr(dev))
        res = TransformResult(dict(zip(names, (pred, prob, raw))), n)
        if output == "host":
            h.call("fd_status", _stream_ptr(dev))
            res = TransformResult({k: v.cpu().numpy() for k, v in res._cols.items()}, n)
        return res

    def featurize(self, batch, cleaned: Optional[bool] = None, device: Optional[int] = None):
        """Feature stages only -> CSR ``(rowptr int64[N+1], idx int32[nnz], val f64[nnz])`` device tensors:
        the ``features`` column (or ``raw_features`` when the pipeline has no IDF stage)."""
        torch = _torch()
        h = self.handle(device)
        pk, flags = self._packed(batch, cleaned)
        dev = torch.device("cuda", h.device)
        pk = pk.to_device(dev)
        n = pk.n_rows
        nnz = C.c_int64(0)
        st = _stream_ptr(dev)
        h.call("fd_featurize_begin", _lib.ptr(pk.data), _lib.ptr(pk.offs), n, flags, C.byref(nnz), st)
        rowptr = torch.empty(n + 1, dtype=torch.int64, device=dev)
        idx = torch.empty(max(nnz.value, 1), dtype=torch.int32, device=dev)
        val = torch.empty(max(nnz.value, 1), dtype=torch.float64, device=dev)
        h.call("fd_featurize_finish", _lib.ptr(rowptr), _lib.ptr(idx), _lib.ptr(val), nnz.value, st)
        return rowptr, idx[:nnz.value], val[:nnz.value]

    def score(self, rowptr, idx, val, device: Optional[int] = None) -> TransformResult:
        """Classifier stage on CSR features (device tensors)."""
        torch = _torch()
        h = self.handle(device)
        dev = torch.device("cuda", h.device)
        n = int(rowptr.shape[0]) - 1
        pred = torch.empty(n, dtype=torch.float64, device=dev)
        prob = torch.empty((n, 2), dtype=torch.float64, device=dev)
        raw = torch.empty((n, 2), dtype=torch.float64, device=dev)
        h.call("fd_score", _lib.ptr(rowptr), _lib.ptr(idx), _lib.ptr(val), n, _lib.ptr(pred), _lib.ptr(prob), _lib.ptr(raw),
               _stream_ptr(dev))
        names = [self._plan.classifier.getOrDefault(k) for k in ("predictionCol", "probabilityCol", "rawPredictionCol")]
        return TransformResult(dict(zip(names, (pred, prob, raw))), n)


class Pipeline:
    """``pyspark.ml.Pipeline``: ``Pipeline(stages=[...]).fit(batch)`` (reference fraud_detection_spark.py:85-91).

    Supported shape: Tokenizer -> [StopWordsRemover] -> HashingTF -> [IDF | IDFModel] -> tree estimator / model.
    Feature stages run through ``fd_featurize``; ``IDF.fit`` counts document frequencies on the device (summed
    over ranks); the tree estimator trains with the fd_rf_* kernels.  Each rank passes ITS shard of rows."""

    def __init__(self, stages: Sequence[S.Stage]):
        self.stages = list(stages)

    def getStages(self):
        return self.stages

    def fit(self, batch, labels=None, cleaned: Optional[bool] = None, device: Optional[int] = None, **explicit) -> PipelineModel:
        torch = _torch()
        import torch.distributed as dist
        feats = [s for s in self.stages if isinstance(s, (S.Tokenizer, S.StopWordsRemover, S.HashingTF, S.CountVectorizer, S.CountVectorizerModel))]
        idf = [s for s in self.stages if isinstance(s, (S.IDF, S.IDFModel))]
        est = [s for s in self.stages if isinstance(s, (S._TreeEstimator, S.ClassificationModel))]
        if len(idf) > 1 or len(est) > 1 or len(feats) + len(idf) + len(est) != len(self.stages):
            raise ValueError("unsupported pipeline shape for the B200 path")
        if isinstance(batch, dict) and labels is None and est:
            labels = batch[est[0].getOrDefault("labelCol")]
        for i, st in enumerate(feats):
            if isinstance(st, S.CountVectorizer):  # fit the vocabulary on the device, from the stages before it
                pre = PipelineModel(feats[:i] + [S.HashingTF(inputCol=st.getOrDefault("inputCol"), outputCol=st.getOrDefault("outputCol"))])
                pk, flags = pre._packed(batch, cleaned)
                hnd = pre.handle(device)
                pk = pk.to_device(torch.device("cuda", hnd.device))
                feats[i] = st.fit_packed(hnd, pk, flags)
                batch, cleaned = pk, not bool(flags & _lib.FD_FLAG_CLEAN)
        import time
        from . import training as _training
        prof = _training.PROFILE_DEFAULT  # phase timers (they synchronise the device)
        ph: Dict[str, float] = {}

        def tick(name, t0):
            if prof:
                torch.cuda.synchronize()
                ph[name] = ph.get(name, 0.0) + time.perf_counter() - t0
            return time.perf_counter()

        tt = time.perf_counter()
        tf_model = PipelineModel(feats)
        rowptr, idx, val = tf_model.featurize(batch, cleaned=cleaned, device=device)
        tt = tick("featurize", tt)
        dev = rowptr.device
        nf = tf_model._plan.num_features
        fitted: List[S.Stage] = list(feats)
        if idf:
            st = idf[0]
            if isinstance(st, S.IDF):
                # docFreq[j] = documents with value > 0 at j: counted by the library's kernel, summed over ranks below
                df = torch.empty(nf, dtype=torch.int64, device=dev)
                tf_model.handle(device).call("fd_doc_freq", _lib.ptr(idx), _lib.ptr(val), int(idx.numel()), int(nf), _lib.ptr(df), _stream_ptr(dev))
                nd = torch.tensor([rowptr.numel() - 1], dtype=torch.int64, device=dev)
                if dist.is_available() and dist.is_initialized():
                    dist.all_reduce(df)
                    dist.all_reduce(nd)
                st = st.fit_from_doc_freq(df.cpu().numpy(), int(nd.item()))
            fitted.append(st)
            d_idf = torch.from_numpy(np.ascontiguousarray(st.idf, dtype=np.float64)).to(dev)
            tf_model.handle(device).call("fd_idf_scale", _lib.ptr(idx), _lib.ptr(val), int(idx.numel()), _lib.ptr(d_idf), int(nf),
                                         _stream_ptr(dev))  # IDFModel.transform: one exact fp64 multiply per entry, in place
            tt = tick("idf fit+transform", tt)
        if est:
            c = est[0]
            if isinstance(c, S._TreeEstimator):
                lab = labels if torch.is_tensor(labels) else torch.as_tensor(np.asarray(labels))
                lab = lab.to(dev).to(torch.uint8)
                c = c.fit_csr(tf_model.handle(device), rowptr, idx, val, lab, nf, **explicit)
                c.paramMap.setdefault("featuresCol", fitted[-1].getOrDefault("outputCol"))
                tt = tick("tree estimator", tt)
                if prof and hasattr(c, "trainer_stats"):
                    c.trainer_stats["pipeline_phases_s"] = dict(ph)
            fitted.append(c)
        return PipelineModel(fitted)


def predict_and_get_label(model: PipelineModel, text: str) -> Dict[str, Any]:
    """The reference's ``DeepSeekClassificationAgent.predict_and_get_label`` (utils/agent_api.py:155-175) on
    this pipeline: raw dialogue in, ``{"prediction", "confidence"}`` out (confidence = probability[1])."""
    out = model.transform([text])
    pred_label = out.select("prediction").first()[0]
    try:
        confidence = float(out.select("probability").first()[0][1])
    except Exception as e:  # same contract as the reference: failure to read the vector -> None
        print(f"Error getting confidence: {e}")
        confidence = None
    return {"prediction": pred_label, "confidence": confidence}
