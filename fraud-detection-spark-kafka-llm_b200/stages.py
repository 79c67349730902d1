"""Host-side mirrors of the pyspark.ml stages on the hot path (same class names, params and on-disk
format), holding only parameters and model data; all per-dialogue arithmetic runs in the CUDA library.

Reference call sites: ``fraud_detection_spark.py:50-53`` (Tokenizer, StopWordsRemover, IDF),
``fraud_detection_spark.py:59-74`` (DecisionTreeClassifier, RandomForestClassifier) and the shipped
``dialogue_classification_model/stages/*`` (HashingTF, IDFModel, LogisticRegressionModel).
"""
from __future__ import annotations

import os
import uuid
from typing import Any, Dict, List, Optional, Sequence

import numpy as np
import pyarrow as pa

from . import spark_format as sf


def _uid(prefix: str) -> str:
    return f"{prefix}_{uuid.uuid4().hex[:12]}"


class Stage:
    """Common param plumbing (``Params``): explicit params override defaults, as in Spark."""

    spark_class = ""
    uid_prefix = "Stage"
    defaults: Dict[str, Any] = {}

    def __init__(self, uid: Optional[str] = None, **params):
        self.uid = uid or _uid(self.uid_prefix)
        self.paramMap: Dict[str, Any] = {k: v for k, v in params.items() if v is not None}
        self.defaultParamMap: Dict[str, Any] = dict(self.defaults)

    def getOrDefault(self, name: str):
        if name in self.paramMap:
            return self.paramMap[name]
        if name in self.defaultParamMap:
            return self.defaultParamMap[name]
        raise KeyError(f"{type(self).__name__} has no param '{name}'")

    def set(self, name: str, value):
        self.paramMap[name] = value
        return self

    def __getattr__(self, item):
        # getInputCol() / getNumFeatures() ... style accessors
        if item.startswith("get") and len(item) > 3:
            name = item[3].lower() + item[4:]
            return lambda: self.getOrDefault(name)
        raise AttributeError(item)

    # ---- persistence
    def _extra_metadata(self) -> Dict[str, Any]:
        return {}

    def _write_data(self, path: str) -> None:
        pass

    def save(self, path: str) -> None:
        os.makedirs(path, exist_ok=True)
        sf.write_metadata(path, self.spark_class, self.uid, self.paramMap, self.defaultParamMap, self._extra_metadata())
        self._write_data(path)

    @classmethod
    def _from_metadata(cls, meta: Dict[str, Any]):
        obj = cls.__new__(cls)
        obj.uid = meta["uid"]
        obj.paramMap = dict(meta.get("paramMap", {}))
        obj.defaultParamMap = dict(meta.get("defaultParamMap", {}))
        return obj

    @classmethod
    def _load(cls, path: str, meta: Dict[str, Any]):
        return cls._from_metadata(meta)


# ------------------------------------------------------------------------------------ feature stages
class Tokenizer(Stage):
    """``org.apache.spark.ml.feature.Tokenizer`` (reference fraud_detection_spark.py:50)."""
    spark_class = "org.apache.spark.ml.feature.Tokenizer"
    uid_prefix = "Tokenizer"

    def __init__(self, inputCol=None, outputCol=None, uid=None):
        super().__init__(uid, inputCol=inputCol, outputCol=outputCol)
        self.defaultParamMap = {"outputCol": self.uid + "__output"}


ENGLISH_STOP_WORDS: List[str] = []  # filled lazily from the bundled list


def load_default_stop_words(language: str = "english") -> List[str]:
    """``StopWordsRemover.loadDefaultStopWords("english")`` -- Spark's bundled 181-word list."""
    if language != "english":
        raise ValueError("only the english default stop-word list is bundled")
    global ENGLISH_STOP_WORDS
    if not ENGLISH_STOP_WORDS:
        here = os.path.dirname(os.path.abspath(__file__))
        with open(os.path.join(here, "english_stop_words.txt"), encoding="utf-8") as f:
            ENGLISH_STOP_WORDS = [w.rstrip("\n") for w in f if w.rstrip("\n")]
    return list(ENGLISH_STOP_WORDS)


class StopWordsRemover(Stage):
    """``org.apache.spark.ml.feature.StopWordsRemover`` (reference fraud_detection_spark.py:51)."""
    spark_class = "org.apache.spark.ml.feature.StopWordsRemover"
    uid_prefix = "StopWordsRemover"

    def __init__(self, inputCol=None, outputCol=None, stopWords=None, caseSensitive=None, locale=None, uid=None):
        super().__init__(uid, inputCol=inputCol, outputCol=outputCol, stopWords=stopWords, caseSensitive=caseSensitive,
                         locale=locale)
        self.defaultParamMap = {"caseSensitive": False, "locale": "en", "stopWords": load_default_stop_words(),
                                "outputCol": self.uid + "__output"}


class HashingTF(Stage):
    """``org.apache.spark.ml.feature.HashingTF`` (shipped stage 2; murmur3_x86_32 seed 42)."""
    spark_class = "org.apache.spark.ml.feature.HashingTF"
    uid_prefix = "HashingTF"

    def __init__(self, inputCol=None, outputCol=None, numFeatures=None, binary=None, uid=None):
        super().__init__(uid, inputCol=inputCol, outputCol=outputCol, numFeatures=numFeatures, binary=binary)
        self.defaultParamMap = {"outputCol": self.uid + "__output", "numFeatures": 262144, "binary": False}


class CountVectorizerModel(Stage):
    """``org.apache.spark.ml.feature.CountVectorizerModel`` -- the featuriser the reference's training script builds
    (``CountVectorizer(..., vocabSize=20000)``, fraud_detection_spark.py:52).  Data: ``vocabulary: array<string>``;
    term -> index by dictionary, terms outside the vocabulary are ignored.  Only the default ``minTF = 1.0`` is
    implemented on the device path."""
    spark_class = "org.apache.spark.ml.feature.CountVectorizerModel"
    uid_prefix = "CountVectorizer"

    def __init__(self, vocabulary: Sequence[str], inputCol=None, outputCol=None, binary=None, minTF=None, uid=None):
        super().__init__(uid, inputCol=inputCol, outputCol=outputCol, binary=binary, minTF=minTF)
        self.defaultParamMap = {"outputCol": self.uid + "__output", "vocabSize": 262144, "minDF": 1.0,
                                "maxDF": 9.223372036854776e18, "minTF": 1.0, "binary": False}
        self.vocabulary = list(vocabulary)

    @classmethod
    def _load(cls, path, meta):
        obj = cls._from_metadata(meta)
        t = sf.read_parquet_dir(os.path.join(path, "data"))
        obj.vocabulary = list(t.column("vocabulary")[0].as_py())
        return obj

    def _write_data(self, path):
        schema = pa.schema([pa.field("vocabulary", pa.list_(pa.field("element", pa.string(), nullable=True)))])
        sf.write_parquet_dir(os.path.join(path, "data"), pa.Table.from_pylist([{"vocabulary": self.vocabulary}], schema=schema))


class CountVectorizer(Stage):
    """``org.apache.spark.ml.feature.CountVectorizer`` estimator (fraud_detection_spark.py:52); fitted on the device
    by ``Pipeline.fit`` (term counts -> top ``vocabSize`` terms)."""
    spark_class = "org.apache.spark.ml.feature.CountVectorizer"
    uid_prefix = "CountVectorizer"

    def __init__(self, inputCol=None, outputCol=None, vocabSize=None, minDF=None, maxDF=None, minTF=None, binary=None, uid=None):
        super().__init__(uid, inputCol=inputCol, outputCol=outputCol, vocabSize=vocabSize, minDF=minDF, maxDF=maxDF, minTF=minTF, binary=binary)
        self.defaultParamMap = {"outputCol": self.uid + "__output", "vocabSize": 262144, "minDF": 1.0,
                                "maxDF": 9.223372036854776e18, "minTF": 1.0, "binary": False}

    def fit_packed(self, handle, packed, flags: int) -> CountVectorizerModel:
        """Count terms over this rank's packed dialogues on the device (``fd_vocab_count``), merge the per-rank
        tables when a process group is initialised, and keep the ``vocabSize`` most frequent terms that pass
        minDF / maxDF.  Count ties go to the smaller term bytes (Spark leaves them to partition order)."""
        import ctypes as C

        import torch
        import torch.distributed as dist

        from . import _lib
        dev = packed.data.device
        n = packed.n_rows
        total_bytes = int((packed.offs[-1] - packed.offs[0]).item()) if n else 0
        log2 = max(12, min(28, int(np.ceil(np.log2(max(4096, total_bytes // 16))))))
        pool_bytes = max(1 << 16, total_bytes // 8)
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        while True:
            keys = torch.full((1 << log2,), -1, dtype=torch.int64, device=dev)
            info = torch.zeros((1 << log2, 4), dtype=torch.int32, device=dev)
            pool = torch.zeros(pool_bytes, dtype=torch.uint8, device=dev)
            counters = torch.zeros(3, dtype=torch.int64, device=dev)
            handle.call("fd_vocab_count", _lib.ptr(packed.data), _lib.ptr(packed.offs), n, int(flags), _lib.ptr(keys), _lib.ptr(info), log2,
                        _lib.ptr(pool), pool_bytes, _lib.ptr(counters), st)
            handle.call("fd_status", st)
            entries, used, overflow = (int(x) for x in counters.tolist())
            if not overflow and entries <= 0.6 * (1 << log2):
                break
            if used > pool_bytes:
                pool_bytes = used + (1 << 16)
            if overflow and used <= pool_bytes or entries > 0.6 * (1 << log2):
                log2 += 2
            if log2 > 30:
                raise RuntimeError("too many distinct terms for the term table")
        occ = torch.nonzero(keys != -1).flatten()
        k = keys[occ].cpu().numpy().view(np.uint64)
        inf = info[occ].cpu().numpy().view(np.uint32)
        pool_h = pool[:max(used, 1)].cpu().numpy()
        terms = {}
        for key, (cnt, df, off, ln) in zip(k.tolist(), inf.tolist()):
            if ln <= 7:
                b = int(key & ((1 << 56) - 1)).to_bytes(7, "little")[:ln]
            elif off == 0xFFFFFFFF:
                continue  # a term longer than a staging window: not representable (documented limit)
            else:
                b = bytes(pool_h[off:off + ln])
            terms[b] = (cnt, df)
        n_docs = n
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            parts = [None] * dist.get_world_size()
            dist.all_gather_object(parts, (terms, n))
            terms, n_docs = {}, 0
            for t, nd in parts:
                n_docs += nd
                for b, (c, d) in t.items():
                    c0, d0 = terms.get(b, (0, 0))
                    terms[b] = (c0 + c, d0 + d)
        g = self.getOrDefault
        min_df, max_df = float(g("minDF")), float(g("maxDF"))
        lo = min_df if min_df >= 1.0 else min_df * n_docs
        hi = max_df if max_df >= 1.0 else max_df * n_docs
        keep = [b for b, (c, d) in terms.items() if lo <= d <= hi] if ("minDF" in self.paramMap or "maxDF" in self.paramMap) else list(terms)
        keep.sort(key=lambda b: (-terms[b][0], b))
        vocab = [b.decode("utf-8", "replace") for b in keep[:int(g("vocabSize"))]]
        if not vocab:
            raise ValueError("The vocabulary size should be > 0. Lower minDF as necessary.")
        model = self.model_from_vocabulary(vocab)
        model.term_counts = {b: terms[b] for b in keep[:int(g("vocabSize"))]}
        return model

    def model_from_vocabulary(self, vocabulary: Sequence[str]) -> CountVectorizerModel:
        m = CountVectorizerModel(vocabulary, uid=self.uid)
        m.paramMap = dict(self.paramMap)
        m.defaultParamMap = dict(self.defaultParamMap)
        return m


class IDF(Stage):
    """``org.apache.spark.ml.feature.IDF`` estimator (reference fraud_detection_spark.py:53)."""
    spark_class = "org.apache.spark.ml.feature.IDF"
    uid_prefix = "IDF"

    def __init__(self, inputCol=None, outputCol=None, minDocFreq=None, uid=None):
        super().__init__(uid, inputCol=inputCol, outputCol=outputCol, minDocFreq=minDocFreq)
        self.defaultParamMap = {"outputCol": self.uid + "__output", "minDocFreq": 0}

    def fit_from_doc_freq(self, doc_freq: np.ndarray, num_docs: int) -> "IDFModel":
        """``idf[j] = log((m + 1) / (df[j] + 1))``, 0 below minDocFreq (IDF.DocumentFrequencyAggregator.idf).
        Evaluated on the host with the platform's correctly-rounded double ``log`` (bit-exact against
        the 10,000 Spark-stored doubles of the shipped model, golden vector G1)."""
        df = np.asarray(doc_freq, dtype=np.int64)
        idf = np.log((float(num_docs) + 1.0) / (df.astype(np.float64) + 1.0))
        idf[df < int(self.getOrDefault("minDocFreq"))] = 0.0
        m = IDFModel(idf, df, int(num_docs), uid=self.uid)
        m.paramMap = dict(self.paramMap)
        m.defaultParamMap = dict(self.defaultParamMap)
        return m


class IDFModel(Stage):
    """``org.apache.spark.ml.feature.IDFModel`` (shipped stage 3)."""
    spark_class = "org.apache.spark.ml.feature.IDFModel"
    uid_prefix = "IDF"

    def __init__(self, idf, docFreq=None, numDocs=0, uid=None, inputCol=None, outputCol=None):
        super().__init__(uid, inputCol=inputCol, outputCol=outputCol)
        self.defaultParamMap = {"outputCol": self.uid + "__output", "minDocFreq": 0}
        self.idf = np.ascontiguousarray(idf, dtype=np.float64)
        self.docFreq = np.zeros(self.idf.size, np.int64) if docFreq is None else np.ascontiguousarray(docFreq, dtype=np.int64)
        self.numDocs = int(numDocs)

    @classmethod
    def _load(cls, path, meta):
        obj = cls._from_metadata(meta)
        t = sf.read_parquet_dir(os.path.join(path, "data"))
        obj.idf = sf.decode_vector(t.column("idf")[0].as_py())
        if "docFreq" in t.column_names:
            obj.docFreq = np.asarray(t.column("docFreq")[0].as_py(), dtype=np.int64)
            obj.numDocs = int(t.column("numDocs")[0].as_py())
        else:  # written by Spark < 3.0
            obj.docFreq = np.zeros(obj.idf.size, np.int64)
            obj.numDocs = 0
        return obj

    def _write_data(self, path):
        schema = pa.schema([pa.field("idf", sf.VECTOR_TYPE), pa.field("docFreq", pa.list_(pa.field("element", pa.int64(), nullable=False))),
                            pa.field("numDocs", pa.int64(), nullable=False)])
        t = pa.Table.from_pylist([{"idf": sf.encode_dense_vector(self.idf), "docFreq": [int(x) for x in self.docFreq],
                                   "numDocs": self.numDocs}], schema=schema)
        sf.write_parquet_dir(os.path.join(path, "data"), t)


# ------------------------------------------------------------------------------------ classifiers
_CLS_DEFAULTS = {"featuresCol": "features", "labelCol": "label", "predictionCol": "prediction",
                 "probabilityCol": "probability", "rawPredictionCol": "rawPrediction"}


class ClassificationModel(Stage):
    numClasses = 2
    numFeatures = 0


class LogisticRegressionModel(ClassificationModel):
    """Binomial ``org.apache.spark.ml.classification.LogisticRegressionModel`` (shipped stage 4)."""
    spark_class = "org.apache.spark.ml.classification.LogisticRegressionModel"
    uid_prefix = "LogisticRegression"
    defaults = dict(_CLS_DEFAULTS, threshold=0.5)

    def __init__(self, coefficients, intercept, uid=None, **params):
        super().__init__(uid, **params)
        self.coefficients = np.ascontiguousarray(coefficients, dtype=np.float64)
        self.intercept = float(intercept)
        self.numFeatures = int(self.coefficients.size)

    @classmethod
    def _load(cls, path, meta):
        obj = cls._from_metadata(meta)
        t = sf.read_parquet_dir(os.path.join(path, "data"))
        if bool(t.column("isMultinomial")[0].as_py()) or int(t.column("numClasses")[0].as_py()) != 2:
            raise ValueError("only binomial LogisticRegressionModel is supported on this path")
        mat = sf.decode_matrix(t.column("coefficientMatrix")[0].as_py())
        obj.coefficients = np.ascontiguousarray(mat[0], dtype=np.float64)
        obj.intercept = float(sf.decode_vector(t.column("interceptVector")[0].as_py())[0])
        obj.numFeatures = int(t.column("numFeatures")[0].as_py())
        obj.numClasses = 2
        return obj

    def _write_data(self, path):
        nz = np.nonzero(self.coefficients)[0]
        schema = pa.schema([pa.field("numClasses", pa.int32(), nullable=False), pa.field("numFeatures", pa.int32(), nullable=False),
                            pa.field("interceptVector", sf.VECTOR_TYPE), pa.field("coefficientMatrix", sf.MATRIX_TYPE),
                            pa.field("isMultinomial", pa.bool_(), nullable=False)])
        row = {"numClasses": 2, "numFeatures": self.numFeatures, "interceptVector": sf.encode_dense_vector([self.intercept]),
               "coefficientMatrix": {"type": 0, "numRows": 1, "numCols": self.numFeatures, "colPtrs": [0, int(nz.size)],
                                     "rowIndices": [int(i) for i in nz], "values": [float(v) for v in self.coefficients[nz]],
                                     "isTransposed": True},
               "isMultinomial": False}
        sf.write_parquet_dir(os.path.join(path, "data"), pa.Table.from_pylist([row], schema=schema))


class _TreeEnsembleModel(ClassificationModel):
    kind = ""
    trees: List[sf.TreeArrays]
    treeWeights: np.ndarray

    def _forest_arrays(self):
        """Concatenate the trees into the SoA arrays fd_set_forest takes."""
        offs = np.zeros(len(self.trees) + 1, dtype=np.int32)
        np.cumsum([t.n_nodes for t in self.trees], out=offs[1:])
        cat = lambda name: np.ascontiguousarray(np.concatenate([getattr(t, name) for t in self.trees]))
        if self.kind == "gbt":
            leaf = np.ascontiguousarray(np.concatenate([t.prediction for t in self.trees]), dtype=np.float64)
        else:
            leaf = np.ascontiguousarray(np.concatenate([t.stats[:, :self.numClasses] for t in self.trees]), dtype=np.float64)
        return offs, cat("feature"), cat("threshold"), cat("left"), cat("right"), leaf

    @property
    def featureImportances(self) -> np.ndarray:
        """``TreeEnsembleModel.featureImportances`` (reference fraud_detection_spark.py:233, 241): per tree
        sum of gain x node instance count per split feature, normalised per tree (not for GBT, whose
        ``perTreeNormalization`` is false in Spark), summed over the trees, normalised."""
        total = np.zeros(self.numFeatures, dtype=np.float64)
        for t in self.trees:
            imp = np.zeros(self.numFeatures, dtype=np.float64)
            cnt = t.stats.sum(axis=1) if self.kind != "gbt" else t.stats[:, 0]
            for n in range(t.n_nodes):
                if t.feature[n] >= 0:
                    imp[t.feature[n]] += t.gain[n] * cnt[n]
            s = imp.sum()
            if self.kind == "gbt":
                total += imp
            elif s > 0:
                total += imp / s
        s = total.sum()
        return total / s if s > 0 else total


class DecisionTreeClassificationModel(_TreeEnsembleModel):
    """``org.apache.spark.ml.classification.DecisionTreeClassificationModel`` (reference :59-65)."""
    spark_class = "org.apache.spark.ml.classification.DecisionTreeClassificationModel"
    uid_prefix = "DecisionTreeClassifier"
    kind = "dt"
    defaults = dict(_CLS_DEFAULTS, maxDepth=5, maxBins=32, minInstancesPerNode=1, minInfoGain=0.0, impurity="gini", seed=0,
                    maxMemoryInMB=256, cacheNodeIds=False, checkpointInterval=10, minWeightFractionPerNode=0.0, leafCol="")

    def __init__(self, tree: sf.TreeArrays, numFeatures: int, numClasses: int = 2, uid=None, **params):
        super().__init__(uid, **params)
        self.trees = [tree]
        self.treeWeights = np.ones(1)
        self.numFeatures, self.numClasses = int(numFeatures), int(numClasses)

    def _extra_metadata(self):
        return {"numFeatures": self.numFeatures, "numClasses": self.numClasses}

    @classmethod
    def _load(cls, path, meta):
        obj = cls._from_metadata(meta)
        t = sf.read_parquet_dir(os.path.join(path, "data"))
        obj.trees = [sf.nodes_to_tree(t.to_pylist())]
        obj.treeWeights = np.ones(1)
        obj.numFeatures, obj.numClasses = int(meta["numFeatures"]), int(meta["numClasses"])
        return obj

    def _write_data(self, path):
        t = pa.Table.from_pylist(sf.tree_to_nodes(self.trees[0]), schema=pa.schema(sf.NODE_FIELDS))
        sf.write_parquet_dir(os.path.join(path, "data"), t)


class _EnsembleIO(_TreeEnsembleModel):
    tree_class = "org.apache.spark.ml.classification.DecisionTreeClassificationModel"
    tree_uid_prefix = "dtc"

    def _extra_metadata(self):
        d = {"numFeatures": self.numFeatures, "numTrees": len(self.trees)}
        if self.kind == "rf":
            d["numClasses"] = self.numClasses
        return d

    @classmethod
    def _load(cls, path, meta):
        obj = cls._from_metadata(meta)
        rows = sf.read_parquet_dir(os.path.join(path, "data")).to_pylist()
        by_tree: Dict[int, list] = {}
        for r in rows:
            by_tree.setdefault(int(r["treeID"]), []).append(r["nodeData"])
        n_trees = int(meta["numTrees"])
        if sorted(by_tree) != list(range(n_trees)):
            raise ValueError("ensemble data does not hold trees 0..numTrees-1")
        obj.trees = [sf.nodes_to_tree(by_tree[i]) for i in range(n_trees)]
        tm = sf.read_parquet_dir(os.path.join(path, "treesMetadata")).to_pylist()
        w = np.ones(n_trees)
        for r in tm:
            w[int(r["treeID"])] = float(r["weights"])
        obj.treeWeights = w
        obj.numFeatures = int(meta["numFeatures"])
        obj.numClasses = int(meta.get("numClasses", 2))
        return obj

    def _write_data(self, path):
        data_schema = pa.schema([pa.field("treeID", pa.int32(), nullable=False), pa.field("nodeData", sf.NODE_TYPE)])
        rows = [{"treeID": i, "nodeData": nd} for i, t in enumerate(self.trees) for nd in sf.tree_to_nodes(t)]
        sf.write_parquet_dir(os.path.join(path, "data"), pa.Table.from_pylist(rows, schema=data_schema))
        tm_schema = pa.schema([pa.field("treeID", pa.int32(), nullable=False), pa.field("metadata", pa.string()),
                               pa.field("weights", pa.float64(), nullable=False)])
        tm = [{"treeID": i, "metadata": sf.metadata_json(self.tree_class, _uid(self.tree_uid_prefix), {}, {},
                                                         {"numFeatures": self.numFeatures, **({"numClasses": self.numClasses} if self.kind == "rf" else {})}),
               "weights": float(self.treeWeights[i])} for i in range(len(self.trees))]
        sf.write_parquet_dir(os.path.join(path, "treesMetadata"), pa.Table.from_pylist(tm, schema=tm_schema))


class RandomForestClassificationModel(_EnsembleIO):
    """``org.apache.spark.ml.classification.RandomForestClassificationModel`` (reference :67-74)."""
    spark_class = "org.apache.spark.ml.classification.RandomForestClassificationModel"
    uid_prefix = "RandomForestClassifier"
    kind = "rf"
    defaults = dict(_CLS_DEFAULTS, maxDepth=5, maxBins=32, numTrees=20, featureSubsetStrategy="auto", impurity="gini",
                    subsamplingRate=1.0, bootstrap=True, minInstancesPerNode=1, minInfoGain=0.0, seed=0)

    def __init__(self, trees: Sequence[sf.TreeArrays], numFeatures: int, numClasses: int = 2, uid=None, **params):
        super().__init__(uid, **params)
        self.trees = list(trees)
        self.treeWeights = np.ones(len(self.trees))
        self.numFeatures, self.numClasses = int(numFeatures), int(numClasses)


class GBTClassificationModel(_EnsembleIO):
    """``org.apache.spark.ml.classification.GBTClassificationModel`` (BASELINE.json config 4)."""
    spark_class = "org.apache.spark.ml.classification.GBTClassificationModel"
    uid_prefix = "GBTClassifier"
    kind = "gbt"
    tree_class = "org.apache.spark.ml.regression.DecisionTreeRegressionModel"
    tree_uid_prefix = "dtr"
    defaults = dict(_CLS_DEFAULTS, maxDepth=5, maxBins=32, maxIter=20, stepSize=0.1, lossType="logistic", impurity="variance",
                    subsamplingRate=1.0, featureSubsetStrategy="all", minInstancesPerNode=1, minInfoGain=0.0, seed=0)

    def __init__(self, trees: Sequence[sf.TreeArrays], treeWeights: Sequence[float], numFeatures: int, uid=None, **params):
        super().__init__(uid, **params)
        self.trees = list(trees)
        self.treeWeights = np.ascontiguousarray(treeWeights, dtype=np.float64)
        self.numFeatures, self.numClasses = int(numFeatures), 2


# ------------------------------------------------------------------------------------ estimators
class _TreeEstimator(Stage):
    """Shared ``fit_csr``: trains on this rank's CSR feature rows with the CUDA backend (all-reducing the
    per-node histograms over the default process group when one is initialised)."""
    n_trees_param = None

    def _trainer(self, handle, n_features):
        from .training import CudaTreeBackend, ForestTrainer
        g = self.getOrDefault
        n_trees = int(g("numTrees")) if self.n_trees_param else 1
        return ForestTrainer(CudaTreeBackend(handle), n_features, n_trees=n_trees, max_depth=int(g("maxDepth")),
                             max_bins=int(g("maxBins")), min_instances_per_node=int(g("minInstancesPerNode")),
                             min_info_gain=float(g("minInfoGain")),
                             feature_subset_strategy=str(g("featureSubsetStrategy")) if self.n_trees_param else "all",
                             subsampling_rate=float(g("subsamplingRate")) if self.n_trees_param else 1.0, seed=int(g("seed")))

    def fit_csr(self, handle, rowptr, idx, val, labels, n_features, **explicit):
        if str(self.getOrDefault("impurity")).lower() != "gini":
            raise ValueError("only impurity='gini' is implemented on this path")
        trainer = self._trainer(handle, n_features)
        trees = trainer.fit(rowptr, idx, val, labels, **explicit)
        model = self._make_model(trees, n_features)
        model.paramMap = {k: v for k, v in self.paramMap.items()}
        model.trainer_stats = trainer.stats
        return model


class DecisionTreeClassifier(_TreeEstimator):
    """``org.apache.spark.ml.classification.DecisionTreeClassifier`` (reference fraud_detection_spark.py:59-65)."""
    spark_class = "org.apache.spark.ml.classification.DecisionTreeClassifier"
    uid_prefix = "DecisionTreeClassifier"
    defaults = dict(DecisionTreeClassificationModel.defaults)

    def __init__(self, uid=None, **params):
        super().__init__(uid, **params)

    def _make_model(self, trees, n_features):
        return DecisionTreeClassificationModel(trees[0], n_features, 2, uid=self.uid)


class RandomForestClassifier(_TreeEstimator):
    """``org.apache.spark.ml.classification.RandomForestClassifier`` (reference fraud_detection_spark.py:67-74)."""
    spark_class = "org.apache.spark.ml.classification.RandomForestClassifier"
    uid_prefix = "RandomForestClassifier"
    n_trees_param = "numTrees"
    defaults = dict(RandomForestClassificationModel.defaults)

    def __init__(self, uid=None, **params):
        super().__init__(uid, **params)

    def _make_model(self, trees, n_features):
        return RandomForestClassificationModel(trees, n_features, 2, uid=self.uid)


STAGE_CLASSES = {c.spark_class: c for c in (Tokenizer, StopWordsRemover, HashingTF, CountVectorizerModel, IDF, IDFModel, LogisticRegressionModel,
                                             DecisionTreeClassificationModel, RandomForestClassificationModel,
                                             GBTClassificationModel)}


def load_stage(path: str) -> Stage:
    meta = sf.read_metadata(path)
    cls = STAGE_CLASSES.get(meta["class"])
    if cls is None:
        raise ValueError(f"unsupported Spark ML stage class '{meta['class']}' at {path}")
    return cls._load(path, meta)
