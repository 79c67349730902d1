"""ctypes wrapper of oracle/liboracle.so (oracle.c) -- TEST / BASELINE INFRASTRUCTURE ONLY.

Used by tests/ (agreement with spark_oracle.py, and as the fast checker at sizes the Python oracle
cannot finish) and by bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")


class _Model(C.Structure):
    _fields_ = [("sw_bytes", C.c_void_p), ("sw_offs", C.c_void_p), ("n_sw", C.c_int32), ("sw_table", C.c_void_p),
                ("sw_cap", C.c_int32), ("num_features", C.c_int32), ("binary", C.c_int32), ("idf", C.c_void_p),
                ("kind", C.c_int32), ("coef", C.c_void_p), ("intercept", C.c_double), ("raw_threshold", C.c_double),
                ("n_trees", C.c_int32), ("tree_offs", C.c_void_p), ("feature", C.c_void_p), ("threshold", C.c_void_p),
                ("left", C.c_void_p), ("right", C.c_void_p), ("leaf", C.c_void_p), ("tree_w", C.c_void_p)]


def load():
    if not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(os.path.join(_HERE, "oracle.c")):
        subprocess.run(["make", "-s", "-C", _HERE], check=True)
    lib = C.CDLL(_LIB)
    lib.oracle_murmur3.restype = C.c_uint32
    return lib


def _p(a):
    return None if a is None else a.ctypes.data


class COracle:
    """Same construction inputs as spark_oracle.OraclePipeline (classifier dict may be None)."""

    def __init__(self, stop_words, case_sensitive, num_features, binary, idf, classifier):
        self.lib = load()
        sw = list(stop_words) if case_sensitive else [w.lower() for w in stop_words]
        enc = [w.encode("utf-8") for w in sw]
        self._sw_bytes = np.frombuffer(b"".join(enc) + b"\0", dtype=np.uint8).copy()
        self._sw_offs = np.zeros(len(enc) + 1, dtype=np.int32)
        np.cumsum([len(e) for e in enc], out=self._sw_offs[1:])
        cap = 16
        while cap < 2 * max(1, len(enc)):
            cap *= 2
        self._sw_table = np.full(cap, -1, dtype=np.int32)
        self._keep = []
        m = _Model()
        m.sw_bytes, m.sw_offs, m.n_sw, m.sw_table, m.sw_cap = _p(self._sw_bytes), _p(self._sw_offs), len(enc), _p(self._sw_table), cap
        m.num_features, m.binary = int(num_features), int(bool(binary))
        self._idf = None if idf is None else np.ascontiguousarray(idf, dtype=np.float64)
        m.idf = _p(self._idf)
        m.kind = 0
        if classifier is not None:
            k = classifier["kind"]
            if k == "lr":
                m.kind = 1
                self._coef = np.ascontiguousarray(classifier["coef"], dtype=np.float64)
                m.coef = _p(self._coef)
                m.intercept = float(classifier["intercept"])
                t = float(classifier.get("threshold", 0.5))
                m.raw_threshold = -np.inf if t <= 0 else (np.inf if t >= 1 else float(np.log(t / (1.0 - t))))
            else:
                m.kind = {"dt": 2, "rf": 3, "gbt": 4}[k]
                trees = classifier["trees"]
                offs = np.zeros(len(trees) + 1, dtype=np.int32)
                np.cumsum([t.feature.size for t in trees], out=offs[1:])
                self._t = [offs] + [np.ascontiguousarray(np.concatenate([getattr(t, n) for t in trees])) for n in
                                    ("feature", "threshold", "left", "right")]
                self._leaf = np.ascontiguousarray(np.concatenate([t.leaf_value for t in trees]), dtype=np.float64)
                self._tw = np.ascontiguousarray(classifier.get("weights", np.ones(len(trees))), dtype=np.float64)
                m.n_trees = len(trees)
                m.tree_offs, m.feature, m.threshold, m.left, m.right = (_p(a) for a in self._t)
                m.leaf, m.tree_w = _p(self._leaf), _p(self._tw)
        self.m = m
        self.lib.oracle_build_stopwords(C.byref(m))

    def transform(self, data: np.ndarray, offs: np.ndarray, do_clean=True, n_threads=None):
        n = offs.size - 1
        pred = np.empty(n)
        prob = np.empty((n, 2))
        raw = np.empty((n, 2))
        self.lib.oracle_transform(C.byref(self.m), C.c_void_p(data.ctypes.data), C.c_void_p(offs.ctypes.data), C.c_int64(n),
                                  int(do_clean), C.c_void_p(pred.ctypes.data), C.c_void_p(prob.ctypes.data),
                                  C.c_void_p(raw.ctypes.data), int(n_threads or os.cpu_count() or 1))
        return pred, prob, raw

    def featurize(self, data: np.ndarray, offs: np.ndarray, do_clean=True, n_threads=None):
        n = offs.size - 1
        nt = int(n_threads or os.cpu_count() or 1)
        nnz = np.zeros(n, dtype=np.int32)
        args = (C.byref(self.m), C.c_void_p(data.ctypes.data), C.c_void_p(offs.ctypes.data), C.c_int64(n), int(do_clean))
        self.lib.oracle_featurize(*args, C.c_void_p(nnz.ctypes.data), None, None, None, nt)
        rowptr = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(nnz, out=rowptr[1:])
        idx = np.empty(max(1, int(rowptr[-1])), dtype=np.int32)
        val = np.empty(max(1, int(rowptr[-1])), dtype=np.float64)
        self.lib.oracle_featurize(*args, None, C.c_void_p(rowptr.ctypes.data), C.c_void_p(idx.ctypes.data),
                                  C.c_void_p(val.ctypes.data), nt)
        return rowptr, idx[:rowptr[-1]], val[:rowptr[-1]]
