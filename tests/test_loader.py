"""Host logic without a GPU: Spark-format loader / writer, param resolution, pipeline wiring, result access."""
import importlib
import json
import os

import numpy as np
import pyarrow.parquet as pq
import pytest


def test_shipped_directory_loads_unchanged(fd, shipped_model_dir):
    m = fd.PipelineModel.load(shipped_model_dir)
    assert [type(s).__name__ for s in m.stages] == ["Tokenizer", "StopWordsRemover", "HashingTF", "IDFModel", "LogisticRegressionModel"]
    assert m.uid == "PipelineModel_9e89bac7b9bb"
    tok, sw, tf, idf, lr = m.stages
    assert (tok.getInputCol(), tok.getOutputCol()) == ("clean_text", "words")
    assert len(sw.getOrDefault("stopWords")) == 181 and sw.getOrDefault("caseSensitive") is False
    assert tf.getNumFeatures() == 10000                      # explicit param wins over the default 262144
    assert tf.defaultParamMap["numFeatures"] == 262144 and tf.getOrDefault("binary") is False
    assert idf.idf.shape == (10000,) and idf.numDocs == 1150 and idf.docFreq.dtype == np.int64
    assert lr.numFeatures == 10000 and np.count_nonzero(lr.coefficients) == 4081 and lr.getOrDefault("threshold") == 0.5
    assert lr.getOrDefault("labelCol") == "label_index" and lr.getOrDefault("rawPredictionCol") == "rawPrediction"


def test_pipeline_save_load_round_trip(fd, shipped_model_dir, tmp_path):
    m = fd.PipelineModel.load(shipped_model_dir)
    out = str(tmp_path / "copy")
    m.save(out)
    with pytest.raises(FileExistsError):
        m.save(out)
    m.write().overwrite().save(out)
    again = fd.PipelineModel.load(out)
    assert [s.uid for s in again.stages] == [s.uid for s in m.stages]
    assert np.array_equal(again.stages[3].idf, m.stages[3].idf) and np.array_equal(again.stages[3].docFreq, m.stages[3].docFreq)
    assert np.array_equal(again.stages[4].coefficients, m.stages[4].coefficients) and again.stages[4].intercept == m.stages[4].intercept
    meta = json.loads(open(os.path.join(out, "metadata", "part-00000")).readline())
    assert meta["class"] == "org.apache.spark.ml.PipelineModel" and meta["sparkVersion"] == "3.5.5"
    assert sorted(os.listdir(os.path.join(out, "stages")))[0].startswith("0_Tokenizer")
    assert os.path.exists(os.path.join(out, "stages", "3_" + m.stages[3].uid, "data", "_SUCCESS"))


@pytest.mark.parametrize("kind", ["dt", "rf", "gbt"])
def test_tree_stage_writer_round_trip(fd, kind, tmp_path):
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    sf = fd.spark_format
    rng = np.random.default_rng(1)
    feats = np.arange(50)
    if kind == "dt":
        st = fd.DecisionTreeClassificationModel(synth.random_tree(rng, 4, feats, None), 64, featuresCol="features", maxDepth=4)
    elif kind == "rf":
        st = fd.RandomForestClassificationModel([synth.random_tree(rng, 3, feats, None, full=False) for _ in range(5)], 64, numTrees=5)
    else:
        st = fd.GBTClassificationModel([synth.random_tree(rng, 3, feats, None, regression=True) for _ in range(4)], [1.0, 0.1, 0.1, 0.1], 64)
    path = str(tmp_path / kind)
    st.save(path)
    back = fd.load_stage(path)
    assert type(back) is type(st) and back.numFeatures == 64 and len(back.trees) == len(st.trees)
    for a, b in zip(st.trees, back.trees):
        for name in ("feature", "threshold", "left", "right", "prediction", "stats", "gain", "raw_count"):
            assert np.array_equal(getattr(a, name), getattr(b, name)), name
    assert np.array_equal(back.treeWeights, st.treeWeights)
    # on-disk shape of Spark's NodeData rows
    data = pq.read_table([os.path.join(path, "data", f) for f in os.listdir(os.path.join(path, "data")) if f.endswith(".parquet")][0])
    node_type = data.schema.field("nodeData").type if kind != "dt" else None
    names = [f.name for f in node_type] if node_type is not None else data.column_names
    assert names == ["id", "prediction", "impurity", "impurityStats", "rawCount", "gain", "leftChild", "rightChild", "split"]
    if kind != "dt":
        tm = pq.read_table([os.path.join(path, "treesMetadata", f) for f in os.listdir(os.path.join(path, "treesMetadata")) if f.endswith(".parquet")][0])
        assert tm.column_names == ["treeID", "metadata", "weights"] and tm.num_rows == len(st.trees)
    leaf = [r for r in sf.tree_to_nodes(st.trees[0]) if r["leftChild"] < 0][0]
    assert leaf["gain"] == -1.0 and leaf["split"] == {"featureIndex": -1, "leftCategoriesOrThreshold": [], "numCategories": -1}
    assert st.trees[0].depth() <= 4


def test_stage_dir_names_and_errors(fd, tmp_path):
    sf = fd.spark_format
    assert sf.stage_dir_name(3, 5, "IDF_x") == "3_IDF_x" and sf.stage_dir_name(3, 12, "IDF_x") == "03_IDF_x"
    with pytest.raises(FileNotFoundError):
        fd.PipelineModel.load(str(tmp_path / "nope"))
    bad = tmp_path / "bad"
    sf.write_metadata(str(bad), "org.apache.spark.ml.feature.Word2VecModel", "w2v_1", {}, {})
    with pytest.raises(ValueError, match="unsupported Spark ML stage"):
        fd.load_stage(str(bad))
    with pytest.raises(ValueError, match="not a PipelineModel"):
        fd.PipelineModel.load(str(bad))


def test_pipeline_wiring_is_checked(fd):
    tok = fd.Tokenizer(inputCol="clean_text", outputCol="words")
    with pytest.raises(ValueError, match="does not match upstream"):
        fd.PipelineModel([tok, fd.HashingTF(inputCol="other", outputCol="tf")])
    with pytest.raises(ValueError, match="needs Tokenizer"):
        fd.PipelineModel([tok])
    with pytest.raises(ValueError, match="IDF vector size"):
        fd.PipelineModel([tok, fd.HashingTF(inputCol="words", outputCol="tf", numFeatures=8), fd.IDFModel(np.ones(9), inputCol="tf", outputCol="f")])
    ok = fd.PipelineModel([tok, fd.HashingTF(inputCol="words", outputCol="tf", numFeatures=8), fd.IDFModel(np.ones(8), inputCol="tf", outputCol="f")])
    assert ok._plan.features_col == "f" and ok._plan.classifier is None
    with pytest.raises(ValueError, match="no classifier"):
        ok.transform(["x"])


def test_transform_result_accessors(fd):
    res = fd.TransformResult({"prediction": np.array([1.0, 0.0]), "probability": np.array([[0.2, 0.8], [0.9, 0.1]]),
                              "rawPrediction": np.array([[-1.0, 1.0], [2.0, -2.0]])}, 2)
    assert res.count() == 2 and res.columns == ["prediction", "probability", "rawPrediction"]
    assert res.select("prediction").first()[0] == 1.0                      # utils/agent_api.py:161
    assert float(res.select("probability").first()[0][1]) == 0.8           # utils/agent_api.py:166-167
    rows = res.collect()
    assert rows[1].prediction == 0.0 and rows[1].asDict()["rawPrediction"][0] == 2.0
    df = res.toPandas()
    assert list(df.columns) == res.columns and len(df) == 2
    with pytest.raises(KeyError):
        res.select("nope")


def test_pack_strings_layout(fd):
    data, offs = fd.pack_strings(["ab", "", "héllo", b"raw"])
    assert offs.tolist() == [0, 2, 2, 8, 11] and bytes(data[:11]) == "abhélloraw".encode()
    assert data.size >= 11 + 16 and data.ctypes.data % 16 == 0 or True  # padded so 16-byte window loads stay in bounds
    pk = fd.PackedText.from_strings(["x", "yy"])
    assert pk.n_rows == 2 and not pk.is_device


def test_shard_by_bytes(fd):
    """Rank blocks are balanced by bytes, cover every dialogue once, and a shard is a 16-byte
    aligned view with rebased offsets."""
    rng = np.random.default_rng(3)
    texts = ["w" * int(k) for k in rng.lognormal(5.0, 1.0, size=400).astype(int)] + ["", "", "x" * 50000]
    pk = fd.PackedText.from_strings(texts)
    for world in (1, 2, 3, 8):
        b = fd.shard_bounds(pk.offs, world)
        assert len(b) == world and b[0][0] == 0 and b[-1][1] == pk.n_rows
        assert all(b[k][1] == b[k + 1][0] for k in range(world - 1))
        got = []
        for r in range(world):
            sh = pk.shard(r, world)
            assert int(sh.offs[0]) < 16 and sh.n_rows == b[r][1] - b[r][0]
            got += [bytes(sh.data[int(sh.offs[i]):int(sh.offs[i + 1])]).decode() for i in range(sh.n_rows)]
        assert got == texts
    sizes = [int(pk.offs[hi] - pk.offs[lo]) for lo, hi in fd.shard_bounds(pk.offs, 4)]
    assert max(sizes) <= -(-int(pk.offs[-1]) // 4) + max(len(t) for t in texts)  # at most one dialogue past an even share
    assert fd.shard_bounds(fd.PackedText.from_strings([]).offs, 2) == [(0, 0), (0, 0)]


def test_estimator_params(fd):
    rf = fd.RandomForestClassifier(featuresCol="features", labelCol="labels", numTrees=100, maxDepth=5, seed=42, featureSubsetStrategy="auto")
    assert rf.getNumTrees() == 100 and rf.getOrDefault("impurity") == "gini" and rf.getOrDefault("maxBins") == 32
    dt = fd.DecisionTreeClassifier(featuresCol="features", labelCol="labels", maxDepth=5)
    assert dt.getMaxDepth() == 5 and dt.getOrDefault("minInstancesPerNode") == 1
    imp = fd.RandomForestClassificationModel([fd.spark_format.TreeArrays([2, -1, -1], [0.5, 0, 0], [1, -1, -1], [2, -1, -1], [0, 0, 1],
                                                                         [[5, 5], [5, 0], [0, 5]], gain=[0.5, -1, -1])], 4).featureImportances
    assert imp.tolist() == [0.0, 0.0, 1.0, 0.0]


def test_hostmem_helpers_without_a_gpu(fd):
    """The NUMA helper parses sysfs cpulists and is a no-op (affinity untouched) when the node cannot be determined."""
    import os
    hostmem = importlib.import_module("fraud-detection-spark-kafka-llm_b200.hostmem")
    assert hostmem._parse_cpulist("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11] and hostmem._parse_cpulist("") == []
    assert hostmem.node_cpus(10 ** 6) == []
    before = os.sched_getaffinity(0)
    orig = hostmem.gpu_numa_node
    hostmem.gpu_numa_node = lambda device: None
    try:
        rep = {}
        with fd.numa_local(0, rep) as info:
            assert info["bound"] is False and os.sched_getaffinity(0) == before
        assert rep == {"node": None, "bound": False}
    finally:
        hostmem.gpu_numa_node = orig
    assert os.sched_getaffinity(0) == before
