"""Debug helper: featurize a few texts on the GPU and print them next to the oracle's answer."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import fdb200 as fd
from helpers import oracle_for, csr_rows

nf = 1 << 18
stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered"),
          fd.HashingTF(inputCol="filtered", outputCol="raw_features", numFeatures=nf)]
model = fd.PipelineModel(stages)
texts = ["hello", "hello world", "the cat sat on the mat", "a  b", "abcdefghijkl mnopqrstuvwxyz ab", "x " * 40, "social security number please verify"]
ora = oracle_for(model)
rowptr, idx, val = model.featurize(texts, cleaned=False)
for t, (ix, vl) in zip(texts, csr_rows(rowptr, idx, val)):
    eix, evl, _, _ = ora.featurize(t.encode(), do_clean=True)
    print(repr(t[:40]), "OK" if (ix == eix and vl == evl) else "MISMATCH", ix, vl, "| oracle", eix, evl)
