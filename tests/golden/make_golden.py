"""Generate the committed golden fixtures from the reference's Spark-written artefacts.

Run ONCE in the build container, where /root/reference exists:  python tests/golden/make_golden.py
(nothing under tests/, smoke() or bench.py reads /root/reference at run time; they read the files
this script wrote).

Outputs (committed):
  tests/golden/shipped_model.tar.gz   the reference's saved PipelineModel directory
                                      (dialogue_classification_model/, Spark 3.5.5-written JSON metadata
                                      + snappy Parquet + Hadoop .crc sidecars), byte-for-byte, as data.
  tests/golden/golden_vectors.json    G1/G2/G3 known answers DERIVED FROM those Spark-written numbers
                                      (SURVEY.md section 8c), plus the one real dialogue the reference
                                      embeds at utils/agent_api.py:224 as a test input, plus
                                      "provisional" end-to-end outputs produced by oracle/spark_oracle.py
                                      (NOT Spark output; they pin the oracle against regressions only).
"""
import ast
import io
import json
import os
import sys
import tarfile

import numpy as np
import pyarrow.parquet as pq

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
MODEL = os.path.join(REF, "dialogue_classification_model")
sys.path.insert(0, ROOT)

from oracle import spark_oracle as so  # noqa: E402


def _tar_model():
    out = os.path.join(HERE, "shipped_model.tar.gz")
    buf = io.BytesIO()
    with tarfile.open(fileobj=buf, mode="w") as tf:
        for dirpath, dirnames, filenames in sorted(os.walk(MODEL)):
            dirnames.sort()
            for fn in sorted(filenames):
                full = os.path.join(dirpath, fn)
                arc = os.path.join("dialogue_classification_model", os.path.relpath(full, MODEL))
                ti = tf.gettarinfo(full, arcname=arc)
                ti.mtime, ti.uid, ti.gid, ti.uname, ti.gname = 0, 0, 0, "", ""
                with open(full, "rb") as f:
                    tf.addfile(ti, f)
    import gzip
    with open(out, "wb") as f:
        with gzip.GzipFile(fileobj=f, mode="wb", mtime=0) as gz:
            gz.write(buf.getvalue())
    return out


def _sample_dialogue():
    """The only real dataset row in the reference: the string literal in the commented-out usage
    example at utils/agent_api.py:224."""
    src = open(os.path.join(REF, "utils", "agent_api.py"), encoding="utf-8").read().splitlines()
    line = src[223].lstrip("# ").strip()
    return ast.literal_eval(line)


def main():
    idf_t = pq.read_table([os.path.join(MODEL, "stages/3_IDF_58bd96296a82/data", f) for f in
                           os.listdir(os.path.join(MODEL, "stages/3_IDF_58bd96296a82/data")) if f.endswith(".parquet")][0])
    idf = np.asarray(idf_t.column("idf")[0].as_py()["values"], dtype=np.float64)
    doc_freq = np.asarray(idf_t.column("docFreq")[0].as_py(), dtype=np.int64)
    num_docs = int(idf_t.column("numDocs")[0].as_py())
    lr_t = pq.read_table([os.path.join(MODEL, "stages/4_LogisticRegression_65ef69d6dfd9/data", f) for f in
                          os.listdir(os.path.join(MODEL, "stages/4_LogisticRegression_65ef69d6dfd9/data")) if f.endswith(".parquet")][0])
    cm = lr_t.column("coefficientMatrix")[0].as_py()
    intercept = lr_t.column("interceptVector")[0].as_py()["values"][0]
    coef = np.zeros(cm["numCols"], dtype=np.float64)
    coef[np.asarray(cm["rowIndices"])] = np.asarray(cm["values"])
    sw_meta = json.load(open(os.path.join(MODEL, "stages/1_StopWordsRemover_8c0b00b256b3/metadata/part-00000")))
    stop_words = sw_meta["defaultParamMap"]["stopWords"]

    g = {}
    # G1: the IDF formula against every Spark-stored double (checked bit-exactly in the test)
    g["G1"] = {"num_docs": num_docs, "n": int(idf.size),
               "examples": {str(int(df)): float(idf[np.nonzero(doc_freq == df)[0][0]]) for df in (0, 22, 1150)}}
    # G2: buckets present in every training document, and word -> (hash, bucket, docFreq)
    g["G2"] = {"full_df_buckets": [int(i) for i in np.nonzero(doc_freq == num_docs)[0]], "words": {}}
    for w in ["innocent", "", "suspect", "hello", "process", "social", "security", "number", "please", "scheduled",
              "insurance", "pm", "prize", "device", "verify"]:
        h = so.murmur3_x86_32(w.encode(), 42)
        b = so.non_negative_mod(h, idf.size)
        g["G2"]["words"][w] = {"hash": h, "bucket": b, "doc_freq": int(doc_freq[b])}
    g["G2"]["legacy_rejected"] = {w: {"bucket": so.non_negative_mod(so.murmur3_x86_32_legacy(w.encode(), 42), idf.size),
                                      "doc_freq": int(doc_freq[so.non_negative_mod(so.murmur3_x86_32_legacy(w.encode(), 42), idf.size)])}
                                  for w in ["suspect", "hello"]}
    # G3: LR support
    g["G3"] = {"coef_nnz": int(len(cm["values"])), "df_pos": int((doc_freq > 0).sum()),
               "df_full": int((doc_freq == num_docs).sum()), "intercept": float(intercept),
               "is_transposed": bool(cm["isTransposed"]), "col_ptrs": list(cm["colPtrs"])}

    sample = _sample_dialogue()
    texts = ["", "   ", "Hello.",
             "Hi, I'm calling to confirm your dentist appointment scheduled for 3 PM tomorrow.",
             "Your Social Security number has been suspended. Please verify your identity immediately to avoid legal action.",
             sample, "İK test K", "a  b", " a", "a ", "  a b  ", "tab\tsep\nline", "I'M  NOT  A  Robot!!"]
    pipe = so.OraclePipeline(stop_words, False, int(idf.size), False, idf,
                             {"kind": "lr", "coef": coef, "intercept": float(intercept), "threshold": 0.5})
    prov = []
    for t in texts:
        raw = t.encode("utf-8")
        idx, val, ntok, nkept = pipe.featurize(raw, True)
        pred, prob, rawp = pipe.score(idx, val)
        prov.append({"text": t, "clean": so.clean_text(t), "n_tokens": ntok, "n_kept": nkept, "idx": idx, "val": val,
                     "margin": rawp[1], "probability": prob, "prediction": pred})
    g["provisional_oracle_outputs"] = prov
    g["sample_dialogue_source"] = "reference utils/agent_api.py:224 (string literal in a commented-out usage example)"
    with open(os.path.join(HERE, "golden_vectors.json"), "w") as f:
        json.dump(g, f, indent=1)
    print("wrote", _tar_model(), "and golden_vectors.json")


if __name__ == "__main__":
    main()
