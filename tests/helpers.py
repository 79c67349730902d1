"""Shared helpers for the parity tests: oracle construction from a loaded model, messy test text."""
import numpy as np

from oracle import spark_oracle as so


def oracle_for(model):
    """Build the CPU oracle pipeline from a loaded/constructed fdb200 PipelineModel."""
    p = model._plan
    sw = p.stop.getOrDefault("stopWords") if p.stop is not None else []
    cs = bool(p.stop.getOrDefault("caseSensitive")) if p.stop is not None else False
    c = p.classifier
    if c is None:
        cl = None
    elif c.__class__.__name__ == "LogisticRegressionModel":
        cl = {"kind": "lr", "coef": c.coefficients, "intercept": c.intercept, "threshold": float(c.getOrDefault("threshold"))}
    else:
        trees = [so.Tree(t.feature, t.threshold, t.left, t.right, t.prediction if c.kind == "gbt" else t.stats[:, :2]) for t in c.trees]
        cl = {"kind": c.kind, "trees": trees, "weights": list(c.treeWeights)}
    vocab = getattr(p.hashing, "vocabulary", None)
    return so.OraclePipeline(sw, cs, p.num_features, bool(p.hashing.getOrDefault("binary")),
                             p.idf.idf if p.idf is not None else None, cl, vocabulary=vocab)


_WORDS = ["hello", "Hello", "HELLO", "social", "Security", "number", "the", "The", "a", "I", "i'm", "don't", "you're", "is",
          "verify", "account", "3pm", "PM", "x", "zz", "İstanbul", "K", "naïve", "über", "résumé", "call-back", "SSA-123456789",
          "won't", "cannot", "prize", "WINNER", "bank", "of", "to", "and", "antidisestablishmentarianism", "q"]
_SEPS = [" ", " ", " ", " ", "  ", ". ", ", ", "! ", "? ", ": ", "\t", "\n", " - ", "   ", "; "]


def messy_texts(n, seed=0, max_tokens=60):
    """Dialogues with punctuation, mixed case, digits, non-ASCII (incl. U+0130 / U+212A), tabs/newlines,
    runs of spaces, leading/trailing blanks and the occasional empty or separator-only row."""
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        r = rng.random()
        if r < 0.02:
            out.append("")
            continue
        if r < 0.04:
            out.append(" " * int(rng.integers(1, 5)))
            continue
        if r < 0.05:
            out.append("123 456 !!!")
            continue
        k = int(rng.integers(1, max_tokens))
        parts = []
        if rng.random() < 0.2:
            parts.append(" " * int(rng.integers(1, 3)))
        for j in range(k):
            parts.append(_WORDS[int(rng.integers(0, len(_WORDS)))])
            if j + 1 < k:
                parts.append(_SEPS[int(rng.integers(0, len(_SEPS)))])
        if rng.random() < 0.3:
            parts.append(_SEPS[int(rng.integers(0, len(_SEPS)))])
        out.append("".join(parts))
    return out


def csr_rows(rowptr, idx, val):
    rp = rowptr.cpu().numpy()
    ix = idx.cpu().numpy()
    vl = val.cpu().numpy()
    return [(ix[rp[i]:rp[i + 1]].tolist(), vl[rp[i]:rp[i + 1]].tolist()) for i in range(len(rp) - 1)]
