"""Random-forest / decision-tree training on binned sparse TF-IDF rows, row-sharded over ranks.

Restates ``ml.tree.impl.RandomForest.run`` (Spark 3.5.5), reached in the reference through
``Pipeline.fit`` at ``fraud_detection_spark.py:85-91`` with the estimators built at ``:59-74``.  The host code
here only orchestrates: quantile thresholds from a sample (``findSplits``), then level-wise growth where each
level is  histogram kernel -> all-reduce(sum) over ranks -> best-split kernel -> row-advance kernel.  All
per-entry work runs in csrc/fd_train.cu through the C-ABI; histograms are integer counts, so the cross-rank
sum is order-independent and every rank selects identical splits.

The compute backend is injected (``CudaTreeBackend`` in production); tests drive the same orchestration under
``gloo`` with a stand-in backend to cover the multi-rank logic without GPUs.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from . import spark_format as sf

EPSILON = 2.220446049250313e-16  # ml.impl.Utils.EPSILON
MAX_BINS_SUPPORTED = 32
SEG_LEN = 8192                   # column entries per histogram work unit
MAX_ITEMS_PER_GROUP = 1 << 21    # (node, feature) items per histogram pass: 2M x 256 B = 512 MiB


# ------------------------------------------------------------------------------------------ findSplits
def find_splits(feat: np.ndarray, val: np.ndarray, n_features: int, weighted_num_samples: float, num_splits: int):
    """``RandomForest.findSplitsBySorting`` for continuous features with unit instance weights.

    ``feat`` / ``val``: the NON-ZERO entries of the sampled rows.  Returns ``(thr_ptr int32[F+1], thr f64[],
    n_splits int32[F], bin_of_zero u8[F])``; a feature never seen non-zero in the sample has no split."""
    feat = np.asarray(feat, dtype=np.int64)
    val = np.asarray(val, dtype=np.float64)
    keep = val != 0.0
    feat, val = feat[keep], val[keep]
    per_feature: Dict[int, np.ndarray] = {}
    if feat.size:
        order = np.lexsort((val, feat))
        f, v = feat[order], val[order]
        new = np.ones(f.size, dtype=bool)
        new[1:] = (f[1:] != f[:-1]) | (v[1:] != v[:-1])
        starts = np.nonzero(new)[0]
        uf, uv = f[starts], v[starts]
        cnt = np.diff(np.append(starts, f.size)).astype(np.float64)
        fstart = np.nonzero(np.append(True, uf[1:] != uf[:-1]))[0]
        fend = np.append(fstart[1:], uf.size)
        for a, b in zip(fstart, fend):
            fid = int(uf[a])
            values, counts = uv[a:b], cnt[a:b]
            part = float(counts.sum())
            tol = EPSILON * part * 100
            if weighted_num_samples - part > tol:  # add the implied zero-valued rows
                pos = int(np.searchsorted(values, 0.0))
                values = np.insert(values, pos, 0.0)
                counts = np.insert(counts, pos, weighted_num_samples - part)
            possible = values.size - 1
            if possible == 0:
                continue
            if possible <= num_splits:
                per_feature[fid] = (values[:-1] + values[1:]) / 2.0
                continue
            stride = weighted_num_samples / (num_splits + 1)
            out = []
            current = counts[0]
            target = stride
            for i in range(1, values.size):
                previous = current
                current = current + counts[i]
                if abs(previous - target) < abs(current - target):
                    out.append((values[i - 1] + values[i]) / 2.0)
                    target += stride
            if out:
                per_feature[fid] = np.asarray(out, dtype=np.float64)
    n_splits = np.zeros(n_features, dtype=np.int32)
    for fid, t in per_feature.items():
        n_splits[fid] = t.size
    thr_ptr = np.zeros(n_features + 1, dtype=np.int32)
    np.cumsum(n_splits, out=thr_ptr[1:])
    thr = np.zeros(int(thr_ptr[-1]), dtype=np.float64)
    bin_of_zero = np.zeros(n_features, dtype=np.uint8)
    for fid, t in per_feature.items():
        thr[thr_ptr[fid]:thr_ptr[fid + 1]] = t
        bin_of_zero[fid] = int(np.searchsorted(t, 0.0, side="left"))
    return thr_ptr, thr, n_splits, bin_of_zero


def gini(counts: Sequence[float]) -> float:
    total = 0.0
    for c in counts:
        total += c
    if total == 0.0:
        return 0.0
    imp = 1.0
    for c in counts:
        f = c / total
        imp -= f * f
    return imp


def features_per_node(strategy: str, n_trees: int, n_features: int) -> int:
    """``DecisionTreeMetadata.numFeaturesPerNode`` for classification."""
    s = strategy.lower()
    if s == "auto":
        s = "all" if n_trees == 1 else "sqrt"
    if s == "all":
        return n_features
    if s == "sqrt":
        return int(math.ceil(math.sqrt(n_features)))
    if s == "log2":
        return max(1, int(math.ceil(math.log2(n_features))))
    if s == "onethird":
        return int(math.ceil(n_features / 3.0))
    try:
        x = float(s)
    except ValueError:
        raise ValueError(f"unsupported featureSubsetStrategy '{strategy}'")
    return max(1, int(x)) if x >= 1.0 and x == int(x) else int(math.ceil(x * n_features))


def default_subset_fn(seed: int, n_features: int, k: int) -> Callable[[int, int], np.ndarray]:
    """Per-(tree, node) feature subset.  Spark draws it with a reservoir sample seeded per node group
    (XORShiftRandom); that stream is not reproduced -- subsets are an explicit input of the parity contract."""
    if k >= n_features:
        full = np.arange(n_features, dtype=np.int32)
        return lambda t, n: full

    def fn(t: int, n: int) -> np.ndarray:
        rng = np.random.default_rng([seed & 0x7FFFFFFF, t, n])
        out = np.empty(0, dtype=np.int64)
        while out.size < k:
            draw = rng.integers(0, n_features, size=int((k - out.size) * 1.2) + 8)
            cat = np.concatenate([out, draw])
            _, first = np.unique(cat, return_index=True)
            out = cat[np.sort(first)]
        return out[:k].astype(np.int32)

    return fn


# ------------------------------------------------------------------------------------------ backend
class CudaTreeBackend:
    """Thin adapter: torch CUDA tensors in, C-ABI calls (fd_bin_csr, fd_rf_*) on the current stream."""

    def __init__(self, handle: _lib.Handle):
        import torch
        self.torch = torch
        self.h = handle
        self.device = torch.device("cuda", handle.device)

    def _st(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def bin_csr(self, idx, val, thr_ptr, thr):
        out = self.torch.empty(idx.numel(), dtype=self.torch.uint8, device=self.device)
        if thr.numel() == 0:
            thr = self.torch.zeros(1, dtype=self.torch.float64, device=self.device)
        self.h.call("fd_bin_csr", _lib.ptr(idx), _lib.ptr(val), idx.numel(), _lib.ptr(thr_ptr), _lib.ptr(thr), _lib.ptr(out), self._st())
        return out

    def histogram(self, csc, label, bagw, node_id, items, units):
        colptr, rowidx, bins = csc
        it_t, it_n, it_f = items
        u_item, u_seg = units
        n_items = it_t.numel()
        hist = self.torch.empty((n_items, 32, 2), dtype=self.torch.int32, device=self.device)
        self.h.call("fd_rf_histogram", _lib.ptr(colptr), _lib.ptr(rowidx), _lib.ptr(bins), _lib.ptr(label), _lib.ptr(bagw),
                    _lib.ptr(node_id), node_id.shape[1], n_items, _lib.ptr(it_t), _lib.ptr(it_n), _lib.ptr(it_f), u_item.numel(),
                    _lib.ptr(u_item), _lib.ptr(u_seg), SEG_LEN, _lib.ptr(hist), self._st())
        return hist

    def best_split(self, hist, node_item_ptr, item_feat, node_total, n_splits, bin_of_zero, min_instances, min_info_gain):
        t = self.torch
        n = node_item_ptr.numel() - 1
        bi = t.empty(n, dtype=t.int32, device=self.device)
        bs = t.empty(n, dtype=t.int32, device=self.device)
        bg = t.empty(n, dtype=t.float64, device=self.device)
        left = t.empty((n, 2), dtype=t.float64, device=self.device)
        right = t.empty((n, 2), dtype=t.float64, device=self.device)
        self.h.call("fd_rf_best_split", _lib.ptr(hist), n, _lib.ptr(node_item_ptr), _lib.ptr(item_feat), _lib.ptr(node_total),
                    _lib.ptr(n_splits), _lib.ptr(bin_of_zero), int(min_instances), C.c_double(min_info_gain), _lib.ptr(bi), _lib.ptr(bs),
                    _lib.ptr(bg), _lib.ptr(left), _lib.ptr(right), self._st())
        return bi, bs, bg, left, right

    def advance(self, csc, node_id, level_base, splits, units, action):
        colptr, rowidx, bins = csc
        s_t, s_n, s_f, s_b = splits
        u_item, u_seg = units
        self.h.call("fd_rf_advance", _lib.ptr(colptr), _lib.ptr(rowidx), _lib.ptr(bins), _lib.ptr(node_id), node_id.shape[1],
                    node_id.shape[0], int(level_base), s_t.numel(), _lib.ptr(s_t), _lib.ptr(s_n), _lib.ptr(s_f), _lib.ptr(s_b),
                    u_item.numel(), _lib.ptr(u_item), _lib.ptr(u_seg), SEG_LEN, _lib.ptr(action), self._st())


# ------------------------------------------------------------------------------------------ trainer
def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


def _units(col_len: np.ndarray, feats: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """One work unit per SEG_LEN entries of each item's column -> (unit_item int32[], unit_seg int32[])."""
    nseg = (col_len[feats] + SEG_LEN - 1) // SEG_LEN
    unit_item = np.repeat(np.arange(feats.size, dtype=np.int32), nseg)
    first = np.cumsum(nseg) - nseg
    unit_seg = (np.arange(unit_item.size, dtype=np.int64) - np.repeat(first, nseg)).astype(np.int32)
    return unit_item, unit_seg


class ForestTrainer:
    """Level-wise growth of ``n_trees`` classification trees (Gini, binary labels) on this rank's row shard."""

    def __init__(self, backend, n_features: int, n_trees: int = 1, max_depth: int = 5, max_bins: int = 32,
                 min_instances_per_node: int = 1, min_info_gain: float = 0.0, feature_subset_strategy: str = "auto",
                 subsampling_rate: float = 1.0, seed: int = 0, prune: bool = True):
        if max_bins > MAX_BINS_SUPPORTED:
            raise ValueError(f"maxBins <= {MAX_BINS_SUPPORTED} on this path")
        if max_depth > 30:
            raise ValueError("maxDepth <= 30 (Spark's own limit)")
        self.b = backend
        self.F, self.T, self.max_depth, self.max_bins = int(n_features), int(n_trees), int(max_depth), int(max_bins)
        self.min_instances, self.min_info_gain = int(min_instances_per_node), float(min_info_gain)
        self.strategy, self.rate, self.seed, self.prune = feature_subset_strategy, float(subsampling_rate), int(seed), bool(prune)
        self.stats: Dict[str, float] = {}

    # -- thresholds from a row sample gathered over ranks
    def _thresholds(self, rowptr, idx, val, n_total: int, rank: int):
        import torch
        dist = _dist()
        n_local = rowptr.numel() - 1
        required = max(self.max_bins * self.max_bins, 10000)
        fraction = min(1.0, required / float(n_total)) if n_total > 0 else 1.0
        if fraction >= 1.0:
            sel_f, sel_v = idx, val
        else:
            g = torch.Generator(device="cpu")
            g.manual_seed(self.seed + 7919 * (rank + 1))
            pick = (torch.rand(n_local, generator=g) < fraction).to(rowptr.device)
            counts = rowptr[1:] - rowptr[:-1]
            mask = torch.repeat_interleave(pick, counts)
            sel_f, sel_v = idx[mask], val[mask]
        f_np, v_np = sel_f.cpu().numpy(), sel_v.cpu().numpy()
        if dist is not None and dist.get_world_size() > 1:
            parts: List = [None] * dist.get_world_size()
            dist.all_gather_object(parts, (f_np, v_np))
            f_np = np.concatenate([p[0] for p in parts])
            v_np = np.concatenate([p[1] for p in parts])
        num_splits = min(self.max_bins, n_total) - 1
        return find_splits(f_np, v_np, self.F, fraction * n_total, num_splits)

    def fit(self, rowptr, idx, val, labels, bag_weights=None, subset_fn=None, thresholds=None) -> List[sf.TreeArrays]:
        """``rowptr/idx/val``: this rank's CSR rows (torch tensors on the backend device); ``labels`` uint8.
        ``bag_weights`` uint8 [n_trees, n_rows], ``subset_fn(tree, node)`` and ``thresholds`` (the
        ``find_splits`` tuple) may be given explicitly -- they are the parity contract's inputs."""
        import torch
        dist = _dist()
        dev = self.b.device
        rank = dist.get_rank() if dist is not None else 0
        n_local = int(rowptr.numel()) - 1
        nt = torch.tensor([n_local], dtype=torch.int64, device=dev)
        if dist is not None:
            dist.all_reduce(nt)
        n_total = int(nt.item())
        # ---- thresholds, bins, CSC
        thr_ptr, thr, n_splits, bin_of_zero = thresholds if thresholds is not None else self._thresholds(rowptr, idx, val, n_total, rank)
        d_thr_ptr = torch.from_numpy(thr_ptr).to(dev)
        d_thr = torch.from_numpy(thr).to(dev)
        d_nsplits = torch.from_numpy(n_splits).to(dev)
        d_bz = torch.from_numpy(bin_of_zero).to(dev)
        bins = self.b.bin_csr(idx, val, d_thr_ptr, d_thr)
        counts = rowptr[1:] - rowptr[:-1]
        row_of = torch.repeat_interleave(torch.arange(n_local, dtype=torch.int32, device=dev), counts)
        order = torch.argsort(idx, stable=True)
        csc = (torch.zeros(self.F + 1, dtype=torch.int64, device=dev), row_of[order].contiguous(), bins[order].contiguous())
        torch.cumsum(torch.bincount(idx.long(), minlength=self.F), 0, out=csc[0][1:])
        col_len = np.diff(csc[0].cpu().numpy())
        del order, row_of, bins
        labels = labels.to(dev).to(torch.uint8).contiguous()
        # ---- bagging
        if bag_weights is None:
            if self.T == 1:
                bag_weights = torch.ones((1, n_local), dtype=torch.uint8, device=dev)
            else:
                g = torch.Generator(device=dev)
                g.manual_seed(self.seed + rank + 1)
                lam = torch.full((self.T, n_local), self.rate, dtype=torch.float32, device=dev)
                bag_weights = torch.poisson(lam, generator=g).clamp_(max=255).to(torch.uint8)
        bagw = bag_weights.to(dev).to(torch.uint8).contiguous()
        node_id = torch.ones((self.T, n_local), dtype=torch.int32, device=dev)
        onehot = torch.stack([(labels == 0), (labels == 1)], 1).to(torch.float64)
        root = bagw.to(torch.float64) @ onehot  # [T, 2]
        if dist is not None:
            dist.all_reduce(root)
        root = root.cpu().numpy()
        k = features_per_node(self.strategy, self.T, self.F)
        subset_fn = subset_fn or default_subset_fn(self.seed, self.F, k)
        # ---- level-wise growth; nodes[t][heap id] = record
        nodes: List[Dict[int, dict]] = [{1: {"stats": [float(root[t, 0]), float(root[t, 1])], "leaf": True}} for t in range(self.T)]
        frontier = [(t, 1) for t in range(self.T)]
        hist_bytes = allreduce_calls = 0
        for depth in range(self.max_depth):
            base = 1 << depth
            active = [(t, n) for (t, n) in frontier if depth == 0 or abs(gini(nodes[t][n]["stats"])) >= EPSILON]
            if not active:
                break
            splits = []  # (t, n, f, bin)
            next_frontier = []
            g0 = 0
            while g0 < len(active):
                # ---- one node group: items (node x feature subset, features without thresholds dropped)
                it_t, it_n, it_f, ptr = [], [], [], [0]
                g1 = g0
                while g1 < len(active) and (g1 == g0 or ptr[-1] < MAX_ITEMS_PER_GROUP):
                    t, n = active[g1]
                    feats = np.asarray(subset_fn(t, n), dtype=np.int32)
                    feats = feats[n_splits[feats] > 0]
                    it_t.append(np.full(feats.size, t, np.int32))
                    it_n.append(np.full(feats.size, n, np.int32))
                    it_f.append(feats)
                    ptr.append(ptr[-1] + feats.size)
                    g1 += 1
                group = active[g0:g1]
                g0 = g1
                f_all = np.concatenate(it_f) if it_f else np.zeros(0, np.int32)
                if f_all.size == 0:
                    continue
                items = tuple(torch.from_numpy(np.concatenate(a)).to(dev) for a in (it_t, it_n, it_f))
                units = tuple(torch.from_numpy(a).to(dev) for a in _units(col_len, f_all))
                hist = self.b.histogram(csc, labels, bagw, node_id, items, units)
                if dist is not None and dist.get_world_size() > 1:
                    dist.all_reduce(hist)  # integer counts: exact, order-independent
                    allreduce_calls += 1
                hist_bytes += hist.numel() * 4
                totals = torch.tensor([nodes[t][n]["stats"] for (t, n) in group], dtype=torch.float64, device=dev)
                bi, bs, bg, left, right = self.b.best_split(hist, torch.tensor(ptr, dtype=torch.int32, device=dev), items[2], totals,
                                                            d_nsplits, d_bz, self.min_instances, self.min_info_gain)
                bi, bs, bg, left, right = (x.cpu().numpy() for x in (bi, bs, bg, left, right))
                del hist
                for j, (t, n) in enumerate(group):
                    if bi[j] < 0 or not (bg[j] > 0.0):
                        continue  # stays a leaf
                    f = int(f_all[bi[j]])
                    sb = int(bs[j])
                    rec = nodes[t][n]
                    rec.update(leaf=False, feature=f, split_bin=sb, threshold=float(thr[thr_ptr[f] + sb]), gain=float(bg[j]))
                    nodes[t][2 * n] = {"stats": [float(left[j, 0]), float(left[j, 1])], "leaf": True}
                    nodes[t][2 * n + 1] = {"stats": [float(right[j, 0]), float(right[j, 1])], "leaf": True}
                    splits.append((t, n, f, sb))
                    next_frontier += [(t, 2 * n), (t, 2 * n + 1)]
            if not splits:
                break
            sp = np.asarray(splits, dtype=np.int32)
            action = np.full((self.T, base), -1, dtype=np.int32)
            action[sp[:, 0], sp[:, 1] - base] = np.where(bin_of_zero[sp[:, 2]] <= sp[:, 3], 2 * sp[:, 1], 2 * sp[:, 1] + 1)
            d_sp = tuple(torch.from_numpy(np.ascontiguousarray(sp[:, c])).to(dev) for c in range(4))
            units = tuple(torch.from_numpy(a).to(dev) for a in _units(col_len, sp[:, 2]))
            self.b.advance(csc, node_id, base, d_sp, units, torch.from_numpy(action).to(dev))
            frontier = next_frontier
        self.stats = {"hist_bytes": float(hist_bytes), "allreduce_calls": float(allreduce_calls), "n_total": float(n_total)}
        return [self._to_tree(nd) for nd in nodes]

    # -- LearningNode.toNode(prune): heap-indexed records -> pre-order arrays
    def _to_tree(self, nd: Dict[int, dict]) -> sf.TreeArrays:
        def build(i):
            r = nd[i]
            pred = 1 if r["stats"][1] > r["stats"][0] else 0
            if r["leaf"]:
                return dict(leaf=True, rec=r, pred=pred)
            l, rr = build(2 * i), build(2 * i + 1)
            if self.prune and l["leaf"] and rr["leaf"] and l["pred"] == rr["pred"]:
                return dict(leaf=True, rec=r, pred=pred)
            return dict(leaf=False, rec=r, pred=pred, l=l, r=rr)

        feature, threshold, left, right, pred, stats, imp, gain = [], [], [], [], [], [], [], []

        def emit(x):
            me = len(feature)
            r = x["rec"]
            feature.append(-1 if x["leaf"] else r["feature"])
            threshold.append(0.0 if x["leaf"] else r["threshold"])
            left.append(-1)
            right.append(-1)
            pred.append(float(x["pred"]))
            stats.append(list(r["stats"]))
            imp.append(gini(r["stats"]))
            gain.append(-1.0 if x["leaf"] else r["gain"])
            if not x["leaf"]:
                left[me] = emit(x["l"])
                right[me] = emit(x["r"])
            return me

        emit(build(1))
        return sf.TreeArrays(feature, threshold, left, right, pred, np.asarray(stats, dtype=np.float64), imp, gain)
