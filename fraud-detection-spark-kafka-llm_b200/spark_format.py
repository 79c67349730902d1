"""Read and write Spark ML's on-disk model format (JSON metadata + snappy Parquet data).

The reference loads its model with ``PipelineModel.load(model_path)`` (reference
``utils/agent_api.py:129``) and saves one with ``dt_model.save(path)``
(``fraud_detection_spark.py:393``).  The layout restated here is the one the reference's shipped
``dialogue_classification_model/`` directory exhibits:

* ``<dir>/metadata/part-00000`` -- ONE line of JSON ``{class, timestamp, sparkVersion, uid, paramMap,
  defaultParamMap, ...extras}`` next to an empty ``_SUCCESS``;
* a pipeline keeps its stages under ``stages/{idx}_{uid}/`` with the index zero-padded to the digit
  count of the number of stages;
* model data sits in ``<stage>/data/*.parquet`` (possibly several part files).

Hadoop ``.crc`` sidecars are ignored on read and not written (a wrong checksum would be rejected by a
real Hadoop LocalFileSystem, an absent one is not).  Pure host code: json + pyarrow, no device work.
"""
from __future__ import annotations

import json
import os
import time
import uuid
from typing import Any, Dict, List, Optional, Sequence

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq

SPARK_VERSION = "3.5.5"


# ------------------------------------------------------------------ metadata
def read_metadata(path: str) -> Dict[str, Any]:
    """``DefaultParamsReader.loadMetadata``: first line of ``<path>/metadata/part-*``."""
    md = os.path.join(path, "metadata")
    if not os.path.isdir(md):
        raise FileNotFoundError(f"no Spark ML metadata directory at {md}")
    parts = sorted(f for f in os.listdir(md) if f.startswith("part-"))
    if not parts:
        raise FileNotFoundError(f"no part file in {md}")
    with open(os.path.join(md, parts[0]), "r", encoding="utf-8") as f:
        line = f.readline()
    meta = json.loads(line)
    for key in ("class", "uid", "paramMap"):
        if key not in meta:
            raise ValueError(f"{md}: metadata lacks '{key}'")
    meta.setdefault("defaultParamMap", {})
    return meta


def resolve_params(meta: Dict[str, Any]) -> Dict[str, Any]:
    """Explicit ``paramMap`` entries win over ``defaultParamMap`` ones, as ``getAndSetParams`` does."""
    out = dict(meta.get("defaultParamMap", {}))
    out.update(meta.get("paramMap", {}))
    return out


def write_metadata(path: str, cls: str, uid: str, param_map: Dict[str, Any], default_param_map: Dict[str, Any],
                   extra: Optional[Dict[str, Any]] = None) -> None:
    md = os.path.join(path, "metadata")
    os.makedirs(md, exist_ok=True)
    meta = {"class": cls, "timestamp": int(time.time() * 1000), "sparkVersion": SPARK_VERSION, "uid": uid,
            "paramMap": param_map, "defaultParamMap": default_param_map}
    if extra:
        meta.update(extra)
    with open(os.path.join(md, "part-00000"), "w", encoding="utf-8") as f:
        f.write(json.dumps(meta, separators=(",", ":")) + "\n")
    open(os.path.join(md, "_SUCCESS"), "w").close()


def metadata_json(cls: str, uid: str, param_map: Dict[str, Any], default_param_map: Dict[str, Any],
                  extra: Optional[Dict[str, Any]] = None) -> str:
    meta = {"class": cls, "timestamp": int(time.time() * 1000), "sparkVersion": SPARK_VERSION, "uid": uid,
            "paramMap": param_map, "defaultParamMap": default_param_map}
    if extra:
        meta.update(extra)
    return json.dumps(meta, separators=(",", ":"))


# ------------------------------------------------------------------ parquet
def read_parquet_dir(path: str) -> pa.Table:
    files = sorted(os.path.join(path, f) for f in os.listdir(path) if f.endswith(".parquet"))
    if not files:
        raise FileNotFoundError(f"no parquet part file under {path}")
    return pa.concat_tables([pq.read_table(f) for f in files])


def write_parquet_dir(path: str, table: pa.Table) -> None:
    os.makedirs(path, exist_ok=True)
    name = f"part-00000-{uuid.uuid4()}-c000.snappy.parquet"
    pq.write_table(table, os.path.join(path, name), compression="snappy")
    open(os.path.join(path, "_SUCCESS"), "w").close()


# ------------------------------------------------------------------ VectorUDT / MatrixUDT
VECTOR_TYPE = pa.struct([
    pa.field("type", pa.int8(), nullable=False),
    pa.field("size", pa.int32()),
    pa.field("indices", pa.list_(pa.field("element", pa.int32(), nullable=False))),
    pa.field("values", pa.list_(pa.field("element", pa.float64(), nullable=False))),
])

MATRIX_TYPE = pa.struct([
    pa.field("type", pa.int8(), nullable=False),
    pa.field("numRows", pa.int32(), nullable=False),
    pa.field("numCols", pa.int32(), nullable=False),
    pa.field("colPtrs", pa.list_(pa.field("element", pa.int32(), nullable=False))),
    pa.field("rowIndices", pa.list_(pa.field("element", pa.int32(), nullable=False))),
    pa.field("values", pa.list_(pa.field("element", pa.float64(), nullable=False))),
    pa.field("isTransposed", pa.bool_(), nullable=False),
])


def decode_vector(v: Dict[str, Any]) -> np.ndarray:
    """VectorUDT struct -> dense f64 array (type 1 = dense, 0 = sparse)."""
    if v["type"] == 1:
        return np.asarray(v["values"], dtype=np.float64)
    out = np.zeros(int(v["size"]), dtype=np.float64)
    out[np.asarray(v["indices"], dtype=np.int64)] = np.asarray(v["values"], dtype=np.float64)
    return out


def encode_dense_vector(a: Sequence[float]) -> Dict[str, Any]:
    return {"type": 1, "size": None, "indices": None, "values": [float(x) for x in a]}


def decode_matrix(m: Dict[str, Any]) -> np.ndarray:
    """MatrixUDT struct -> dense [numRows, numCols] f64 (type 0 = sparse CSC, or CSR when isTransposed)."""
    nr, nc = int(m["numRows"]), int(m["numCols"])
    vals = np.asarray(m["values"], dtype=np.float64)
    if m["type"] == 1:
        return vals.reshape(nr, nc) if m["isTransposed"] else vals.reshape(nc, nr).T.copy()
    out = np.zeros((nr, nc), dtype=np.float64)
    ptrs = np.asarray(m["colPtrs"], dtype=np.int64)
    inner = np.asarray(m["rowIndices"], dtype=np.int64)
    if m["isTransposed"]:  # CSR: colPtrs are row pointers, rowIndices are column ids
        for r in range(nr):
            out[r, inner[ptrs[r]:ptrs[r + 1]]] = vals[ptrs[r]:ptrs[r + 1]]
    else:
        for c in range(nc):
            out[inner[ptrs[c]:ptrs[c + 1]], c] = vals[ptrs[c]:ptrs[c + 1]]
    return out


# ------------------------------------------------------------------ tree node data (ml.tree.NodeData)
SPLIT_TYPE = pa.struct([
    pa.field("featureIndex", pa.int32(), nullable=False),
    pa.field("leftCategoriesOrThreshold", pa.list_(pa.field("element", pa.float64(), nullable=False))),
    pa.field("numCategories", pa.int32(), nullable=False),
])

NODE_FIELDS = [
    pa.field("id", pa.int32(), nullable=False),
    pa.field("prediction", pa.float64(), nullable=False),
    pa.field("impurity", pa.float64(), nullable=False),
    pa.field("impurityStats", pa.list_(pa.field("element", pa.float64(), nullable=False))),
    pa.field("rawCount", pa.int64(), nullable=False),
    pa.field("gain", pa.float64(), nullable=False),
    pa.field("leftChild", pa.int32(), nullable=False),
    pa.field("rightChild", pa.int32(), nullable=False),
    pa.field("split", SPLIT_TYPE),
]
NODE_TYPE = pa.struct(NODE_FIELDS)


class TreeArrays:
    """One tree as parallel arrays indexed by Spark's pre-order node id (root 0).

    ``feature[n] < 0`` marks a leaf.  ``stats[n]`` is the node's ``impurityStats`` (class counts for a
    classification tree; ``[count, sum, sumSq]`` for a regression tree), ``prediction[n]`` its
    prediction.
    """

    def __init__(self, feature, threshold, left, right, prediction, stats, impurity=None, gain=None, raw_count=None):
        self.feature = np.ascontiguousarray(feature, dtype=np.int32)
        self.threshold = np.ascontiguousarray(threshold, dtype=np.float64)
        self.left = np.ascontiguousarray(left, dtype=np.int32)
        self.right = np.ascontiguousarray(right, dtype=np.int32)
        self.prediction = np.ascontiguousarray(prediction, dtype=np.float64)
        self.stats = np.ascontiguousarray(stats, dtype=np.float64)
        n = self.feature.size
        self.impurity = np.zeros(n) if impurity is None else np.ascontiguousarray(impurity, dtype=np.float64)
        self.gain = np.where(self.feature < 0, -1.0, 0.0) if gain is None else np.ascontiguousarray(gain, dtype=np.float64)
        if raw_count is None:
            raw_count = self.stats.sum(axis=1) if self.stats.ndim == 2 else np.zeros(n)
        self.raw_count = np.ascontiguousarray(raw_count, dtype=np.int64)

    @property
    def n_nodes(self) -> int:
        return int(self.feature.size)

    def depth(self) -> int:
        d = np.zeros(self.n_nodes, dtype=np.int32)
        for n in range(self.n_nodes):  # pre-order: parents precede children
            if self.feature[n] >= 0:
                d[self.left[n]] = d[n] + 1
                d[self.right[n]] = d[n] + 1
        return int(d.max()) if self.n_nodes else 0


def nodes_to_tree(rows: List[Dict[str, Any]]) -> TreeArrays:
    """NodeData rows (any order) -> :class:`TreeArrays`; rejects categorical splits (never produced on
    this path: every TF-IDF feature is continuous)."""
    rows = sorted(rows, key=lambda r: r["id"])
    n = len(rows)
    if [r["id"] for r in rows] != list(range(n)):
        raise ValueError("tree node ids are not 0..n-1")
    feature = np.full(n, -1, np.int32)
    threshold = np.zeros(n)
    left = np.full(n, -1, np.int32)
    right = np.full(n, -1, np.int32)
    pred = np.zeros(n)
    width = max(len(r["impurityStats"]) for r in rows)
    stats = np.zeros((n, width))
    imp = np.zeros(n)
    gain = np.zeros(n)
    raw = np.zeros(n, np.int64)
    for i, r in enumerate(rows):
        pred[i] = r["prediction"]
        imp[i] = r["impurity"]
        gain[i] = r["gain"]
        raw[i] = r["rawCount"]
        stats[i, :len(r["impurityStats"])] = r["impurityStats"]
        if r["leftChild"] >= 0:
            sp = r["split"]
            if sp["numCategories"] != -1:
                raise ValueError("categorical tree splits are not supported on this path")
            feature[i] = sp["featureIndex"]
            threshold[i] = sp["leftCategoriesOrThreshold"][0]
            left[i], right[i] = r["leftChild"], r["rightChild"]
    return TreeArrays(feature, threshold, left, right, pred, stats, imp, gain, raw)


def tree_to_nodes(t: TreeArrays) -> List[Dict[str, Any]]:
    out = []
    for i in range(t.n_nodes):
        leaf = t.feature[i] < 0
        out.append({
            "id": i, "prediction": float(t.prediction[i]), "impurity": float(t.impurity[i]),
            "impurityStats": [float(x) for x in np.atleast_1d(t.stats[i])], "rawCount": int(t.raw_count[i]),
            "gain": -1.0 if leaf else float(t.gain[i]),
            "leftChild": -1 if leaf else int(t.left[i]), "rightChild": -1 if leaf else int(t.right[i]),
            "split": {"featureIndex": -1, "leftCategoriesOrThreshold": [], "numCategories": -1} if leaf else
                     {"featureIndex": int(t.feature[i]), "leftCategoriesOrThreshold": [float(t.threshold[i])], "numCategories": -1},
        })
    return out


def stage_dir_name(idx: int, n_stages: int, uid: str) -> str:
    """``SharedReadWrite.getStagePath``: index padded to the digit count of the number of stages."""
    return f"{idx:0{len(str(n_stages))}d}_{uid}"
