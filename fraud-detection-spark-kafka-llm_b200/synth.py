"""Synthetic dialogues and random tree models of the shapes BASELINE.json names.

The reference's dataset (agent_conversation_all.csv) is not distributable here, so benchmarks and parity
tests run on seeded synthetic text: a 50,000-word lower-case vocabulary plus the 181 English stop words,
word ids drawn from a Zipf law with the stop words at the head (about 49 % of draws, as in the one real
dialogue the reference embeds), one space between tokens and an occasional double space (empty token).
Generation is written in torch so the 1M x 200-token set can be produced on the GPU in well under a
second; the same code runs on CPU for small test sets.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np

from . import _lib
from . import spark_format as sf
from .packing import PackedText
from .stages import load_default_stop_words


def make_vocab(n_words: int = 50000, seed: int = 20261106) -> List[bytes]:
    """n_words distinct lower-case ASCII words, lengths 2..12 (mean about 4.8)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    lens_p = np.array([0.10, 0.17, 0.20, 0.17, 0.12, 0.08, 0.06, 0.04, 0.03, 0.02, 0.01])  # lengths 2..12
    stop = {w.encode() for w in load_default_stop_words()}
    words, seen = [], set(stop)
    while len(words) < n_words:
        L = rng.choice(np.arange(2, 13), p=lens_p / lens_p.sum())
        w = bytes(rng.integers(97, 123, size=L, dtype=np.uint8))
        if w not in seen:
            seen.add(w)
            words.append(w)
    return words


class SynthCorpus:
    """Vocabulary (stop words first) + sampling law; ``sample`` produces packed dialogues."""

    def __init__(self, n_words: int = 50000, seed: int = 20261106, stop_mass: float = 0.49, zipf_s: float = 1.07,
                 stop_words: Optional[Sequence[str]] = None):
        sw = [w.encode() for w in (load_default_stop_words() if stop_words is None else stop_words)]
        sw = [w for w in sw if w.isalpha()]  # cleaned text cannot contain the apostrophe forms
        self.words: List[bytes] = sw + make_vocab(n_words, seed)
        self.n_stop = len(sw)
        self.seed = seed
        maxlen = max(len(w) for w in self.words)
        mat = np.full((len(self.words), maxlen), 0x20, dtype=np.uint8)
        for i, w in enumerate(self.words):
            mat[i, :len(w)] = np.frombuffer(w, dtype=np.uint8)
        self.word_mat = mat
        self.word_len = np.array([len(w) for w in self.words], dtype=np.int64)
        r_s = 1.0 / np.arange(1, self.n_stop + 1) ** 1.0
        r_v = 1.0 / np.arange(1, n_words + 1) ** zipf_s
        p = np.concatenate([stop_mass * r_s / r_s.sum(), (1.0 - stop_mass) * r_v / r_v.sum()])
        self.cdf = np.cumsum(p)
        self.cdf[-1] = 1.0

    def sample(self, n_rows: int, tokens: int = 200, device="cpu", seed: Optional[int] = None, skew: bool = False,
               double_space_p: float = 0.01, chunk_rows: int = 1 << 16, planted: Optional[dict] = None):
        """-> (PackedText on `device`, labels or None).  ``skew``: token counts LogNormal(median `tokens`,
        sigma 0.6) clipped to [1, 4096] instead of exactly `tokens`.  ``planted`` = {"ids": word ids,
        "rate": p}: label-1 rows replace a fraction p of their tokens by those words (training signal)."""
        import torch

        dev = torch.device(device)
        g = torch.Generator(device=dev)
        g.manual_seed(self.seed if seed is None else seed)
        cdf = torch.from_numpy(self.cdf).to(dev)
        wlen = torch.from_numpy(self.word_len).to(dev)
        wmat = torch.from_numpy(self.word_mat).to(dev)
        maxlen = wmat.shape[1]
        bufs, row_lens, labels = [], [], []
        for r0 in range(0, n_rows, chunk_rows):
            nr = min(chunk_rows, n_rows - r0)
            if skew:
                z = torch.randn(nr, generator=g, device=dev, dtype=torch.float64)
                cnt = torch.exp(np.log(tokens) + 0.6 * z).round().clamp_(1, 4096).to(torch.int64)
            else:
                cnt = torch.full((nr,), tokens, dtype=torch.int64, device=dev)
            nt = int(cnt.sum())
            row_of = torch.repeat_interleave(torch.arange(nr, device=dev), cnt)
            ids = torch.searchsorted(cdf, torch.rand(nt, generator=g, device=dev, dtype=torch.float64)).clamp_(max=len(self.words) - 1)
            if planted is not None:
                lab = (torch.rand(nr, generator=g, device=dev) < 0.5)
                pid = torch.as_tensor(planted["ids"], device=dev, dtype=torch.int64)
                swap = lab[row_of] & (torch.rand(nt, generator=g, device=dev) < planted["rate"])
                ids = torch.where(swap, pid[torch.randint(0, pid.numel(), (nt,), generator=g, device=dev)], ids)
                labels.append(lab.to(torch.uint8))
            last = torch.zeros(nt, dtype=torch.bool, device=dev)
            last[torch.cumsum(cnt, 0) - 1] = True
            sep = 1 + (torch.rand(nt, generator=g, device=dev) < double_space_p).to(torch.int64)
            sep = torch.where(last, torch.zeros_like(sep), sep)
            tl = wlen[ids]
            span = tl + sep
            end = torch.cumsum(span, 0)
            start = end - span
            total = int(end[-1]) if nt else 0
            buf = torch.full((total,), 0x20, dtype=torch.uint8, device=dev)
            for j in range(maxlen):
                m = tl > j
                if not bool(m.any()):
                    break
                buf[start[m] + j] = wmat[ids[m], j]
            rl = torch.zeros(nr, dtype=torch.int64, device=dev)
            rl.index_add_(0, row_of, span)
            bufs.append(buf)
            row_lens.append(rl)
        rl = torch.cat(row_lens) if row_lens else torch.zeros(0, dtype=torch.int64, device=dev)
        offs = torch.zeros(n_rows + 1, dtype=torch.int64, device=dev)
        torch.cumsum(rl, 0, out=offs[1:])
        data = torch.cat(bufs + [torch.zeros(32, dtype=torch.uint8, device=dev)])
        lab = torch.cat(labels) if labels else None
        return PackedText(data, offs), lab


def hash_terms(terms: Sequence[bytes], num_features: int) -> np.ndarray:
    """``HashingTF.indexOf`` for many terms through the C-ABI's host-side ``fd_hash_terms``."""
    lib = _lib.load_library()
    offs = np.zeros(len(terms) + 1, dtype=np.int32)
    np.cumsum([len(t) for t in terms], out=offs[1:])
    blob = np.frombuffer(b"".join(terms) + b"\0", dtype=np.uint8).copy()
    out = np.empty(len(terms), dtype=np.int32)
    rc = lib.fd_hash_terms(_lib.ptr(blob), _lib.ptr(offs), len(terms), int(num_features), _lib.ptr(out))
    if rc != 0:
        raise _lib.FdError(rc, "fd_hash_terms failed")
    return out


def random_tree(rng: np.random.Generator, depth: int, features: np.ndarray, idf: Optional[np.ndarray], regression: bool = False,
                full: bool = True, weights: Optional[np.ndarray] = None) -> sf.TreeArrays:
    """A random binary tree of the given depth in Spark's pre-order numbering.  Split features come from
    `features` (buckets that actually occur; drawn uniformly, or in proportion to `weights` -- e.g. their document
    frequencies, as a trained tree prefers features that occur), thresholds sit between the tf-idf values of 0..3
    occurrences so that both branches are taken.  Classification leaves carry random class counts."""
    pw = None if weights is None else np.asarray(weights, dtype=np.float64) / float(np.sum(weights))
    feature, threshold, left, right, pred, stats = [], [], [], [], [], []

    def grow(d):
        me = len(feature)
        leaf = d == depth or (not full and d > 0 and rng.random() < 0.15)
        feature.append(-1)
        threshold.append(0.0)
        left.append(-1)
        right.append(-1)
        if regression:
            v = float(rng.normal(0.0, 0.2))
            cnt = float(rng.integers(1, 1000))
            pred.append(v)
            stats.append([cnt, v * cnt, v * v * cnt])
        else:
            c = rng.integers(0, 500, size=2).astype(np.float64)
            if c.sum() == 0:
                c[0] = 1.0
            pred.append(float(np.argmax(c)))
            stats.append(list(c))
        if not leaf:
            f = int(features[rng.integers(0, len(features)) if pw is None else rng.choice(len(features), p=pw)])
            w = 1.0 if idf is None else float(idf[f])
            feature[me] = f
            threshold[me] = w * (float(rng.integers(0, 3)) + 0.5)
            left[me] = grow(d + 1)
            right[me] = grow(d + 1)
        return me

    grow(0)
    return sf.TreeArrays(feature, threshold, left, right, pred, np.asarray(stats, dtype=np.float64))
