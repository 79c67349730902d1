#!/usr/bin/env python
"""The reference's offline trainer/evaluator (``fraud_detection_spark.py:326-395``) on the B200 path.

Same flow -- load dialogues with labels, drop rows whose cleaned text is empty (:45), 70/10/20 split (:338-339), fit
Tokenizer -> StopWordsRemover -> CountVectorizer(vocabSize=20000) -> IDF -> {DecisionTree(maxDepth=5),
RandomForest(100 trees, maxDepth=5, seed=42)} (:47-91), evaluate AUC / accuracy / weighted precision / recall / F1 and
the confusion matrix on validation and test (:93-123), print the most important vocabulary terms (:224-246), save the
decision-tree pipeline (:388-393) -- without Spark, plots, or the XGBoost benchmark model.  The reference's CSV is not
distributable, so by default a synthetic corpus with planted scam vocabulary stands in; pass --csv to use a file with
`dialogue` and `labels` columns.
"""
import argparse
import importlib
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

import fdb200 as fd  # noqa: E402


def load_dialogues(csv, n_rows, seed):
    if csv:
        import pandas as pd
        df = pd.read_csv(csv)
        df = df[df["labels"].astype(str).str.strip().isin(["0", "1"])]
        return df["dialogue"].astype(str).tolist(), df["labels"].astype(int).to_numpy()
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    corpus = synth.SynthCorpus(n_words=20000, seed=seed)
    planted = {"ids": list(range(corpus.n_stop + 100, corpus.n_stop + 120)), "rate": 0.06}
    pk, lab = corpus.sample(n_rows, tokens=375, device="cpu", seed=seed + 1, planted=planted)
    d, o = pk.data.numpy(), pk.offs.numpy()
    return [bytes(d[o[i]:o[i + 1]]).decode() for i in range(n_rows)], lab.numpy().astype(int)


def build_feature_pipeline():
    return [fd.Tokenizer(inputCol="clean_text", outputCol="words"),
            fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
            fd.CountVectorizer(inputCol="filtered_words", outputCol="raw_features", vocabSize=20000),
            fd.IDF(inputCol="raw_features", outputCol="features")]


def evaluate_model(model, datasets):
    results = {}
    for name, (texts, labels) in datasets.items():
        preds = model.transform(texts)
        cols = {"rawPrediction": preds["rawPrediction"], "prediction": preds["prediction"], "labels": labels.astype(np.float64)}
        multi = fd.MulticlassClassificationEvaluator(labelCol="labels")
        metrics = {"Accuracy": multi.setMetricName("accuracy").evaluate(cols),
                   "Precision": multi.setMetricName("weightedPrecision").evaluate(cols),
                   "Recall": multi.setMetricName("weightedRecall").evaluate(cols),
                   "F1": multi.setMetricName("f1").evaluate(cols),
                   "AUC": fd.BinaryClassificationEvaluator(labelCol="labels", rawPredictionCol="rawPrediction").evaluate(cols)}
        results[name] = {"metrics": metrics, "confusion_matrix": fd.confusion_matrix(cols["labels"], preds["prediction"])["matrix"]}
    return results


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--csv", default=None)
    ap.add_argument("--rows", type=int, default=1600)
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--trees", type=int, default=100)
    ap.add_argument("--save", default="fraud_detection_model")
    a = ap.parse_args(argv)
    texts, labels = load_dialogues(a.csv, a.rows, a.seed)
    keep = [i for i, t in enumerate(texts) if re.sub("[^a-zA-Z ]", "", t.lower()) != ""]   # .filter(col("clean_text") != "")
    texts, labels = [texts[i] for i in keep], labels[keep]
    rng = np.random.default_rng(a.seed)
    u = rng.random(len(texts))
    parts = {"Training": u < 0.7, "Validation": (u >= 0.7) & (u < 0.8), "Test": u >= 0.8}
    data = {k: ([t for t, m in zip(texts, mask) if m], labels[mask]) for k, mask in parts.items()}
    classifiers = {"DecisionTree": fd.DecisionTreeClassifier(featuresCol="features", labelCol="labels", maxDepth=5),
                   "RandomForest": fd.RandomForestClassifier(featuresCol="features", labelCol="labels", numTrees=a.trees, maxDepth=5, seed=42,
                                                             featureSubsetStrategy="auto")}
    models = {name: fd.Pipeline(stages=build_feature_pipeline() + [c]).fit(data["Training"][0], labels=data["Training"][1])
              for name, c in classifiers.items()}
    report = {}
    for name, model in models.items():
        report[name] = evaluate_model(model, {k: data[k] for k in ("Validation", "Test")})
        for split, r in report[name].items():
            print(f"=== {name} / {split} ===")
            for k, v in r["metrics"].items():
                print(f"{k}: {v:.4f}")
            print("Confusion matrix (rows = labels, columns = prediction):\n", r["confusion_matrix"])
    dt = models["DecisionTree"]
    vocab, imp = dt.stages[2].vocabulary, dt.stages[-1].featureImportances
    top = np.argsort(-imp)[:10]
    print("Top features:", [(vocab[i], round(float(imp[i]), 4)) for i in top if imp[i] > 0])
    dt.write().overwrite().save(a.save)
    print(f"Model saved to {a.save}")
    return report, models


if __name__ == "__main__":
    main()
