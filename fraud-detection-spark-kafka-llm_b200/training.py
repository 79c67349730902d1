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
PROFILE_DEFAULT = False          # new trainers time their phases (with device synchronisation) when set


# ------------------------------------------------------------------------------------------ findSplits
def find_splits(feat: np.ndarray, val: np.ndarray, n_features: int, weighted_num_samples: float, num_splits: int):
    """``RandomForest.findSplitsBySorting`` for continuous features with unit instance weights.

    ``feat`` / ``val``: the NON-ZERO entries of the sampled rows.  Returns ``(thr_ptr int32[F+1], thr f64[],
    n_splits int32[F], bin_of_zero u8[F])``; a feature never seen non-zero in the sample has no split.
    Vectorised for the usual case (at most ``num_splits`` + 1 distinct values: every midpoint is a threshold);
    features with more distinct values take Spark's stride walk one by one."""
    feat = np.asarray(feat, dtype=np.int64)
    val = np.asarray(val, dtype=np.float64)
    keep = val != 0.0
    feat, val = feat[keep], val[keep]
    n_splits = np.zeros(n_features, dtype=np.int32)
    bin_of_zero = np.zeros(n_features, dtype=np.uint8)
    thr_ptr = np.zeros(n_features + 1, dtype=np.int32)
    if feat.size == 0:
        return thr_ptr, np.zeros(0, dtype=np.float64), n_splits, bin_of_zero
    # distinct (feature, value) pairs with their counts
    order = np.lexsort((val, feat))
    f, v = feat[order], val[order]
    new = np.ones(f.size, dtype=bool)
    new[1:] = (f[1:] != f[:-1]) | (v[1:] != v[:-1])
    starts = np.nonzero(new)[0]
    uf, uv = f[starts], v[starts]
    cnt = np.diff(np.append(starts, f.size)).astype(np.float64)
    fstart = np.nonzero(np.append(True, uf[1:] != uf[:-1]))[0]
    fids = uf[fstart]
    part = np.add.reduceat(cnt, fstart)
    # the implied zero-valued rows of each feature
    need_zero = (weighted_num_samples - part) > (EPSILON * part * 100)
    af = np.concatenate([uf, fids[need_zero]])
    av = np.concatenate([uv, np.zeros(int(need_zero.sum()))])
    ac = np.concatenate([cnt, weighted_num_samples - part[need_zero]])
    order = np.lexsort((av, af))
    af, av, ac = af[order], av[order], ac[order]
    gstart = np.nonzero(np.append(True, af[1:] != af[:-1]))[0]
    gsize = np.diff(np.append(gstart, af.size))
    gf = af[gstart]
    possible = gsize - 1
    same = np.zeros(af.size, dtype=bool)
    same[1:] = af[1:] == af[:-1]
    mids = np.zeros(af.size)
    mids[1:] = (av[:-1] + av[1:]) / 2.0
    simple = np.repeat(possible <= num_splits, gsize) & same      # midpoint rows of the simple features
    per_feature: Dict[int, np.ndarray] = {}
    for gi in np.nonzero(possible > num_splits)[0]:                # Spark's stride walk for the others
        a0, b0 = gstart[gi], gstart[gi] + gsize[gi]
        values, counts = av[a0:b0], ac[a0:b0]
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
        per_feature[int(gf[gi])] = np.asarray(out, dtype=np.float64)
    np.add.at(n_splits, af[simple], 1)
    for fid, t in per_feature.items():
        n_splits[fid] = t.size
    np.cumsum(n_splits, out=thr_ptr[1:])
    thr = np.zeros(int(thr_ptr[-1]), dtype=np.float64)
    sf_, sm_ = af[simple], mids[simple]                            # already ordered by (feature, value)
    first = np.nonzero(np.append(True, sf_[1:] != sf_[:-1]))[0] if sf_.size else np.zeros(0, dtype=np.int64)
    within = np.arange(sf_.size) - np.repeat(first, np.diff(np.append(first, sf_.size))) if sf_.size else np.zeros(0, dtype=np.int64)
    thr[thr_ptr[sf_] + within] = sm_
    np.add.at(bin_of_zero, sf_[sm_ < 0.0], 1)                      # lower_bound(thresholds, 0.0)
    for fid, t in per_feature.items():
        thr[thr_ptr[fid]:thr_ptr[fid + 1]] = t
        bin_of_zero[fid] = int(np.searchsorted(t, 0.0, side="left"))
    return thr_ptr, thr, n_splits, bin_of_zero


def find_splits_torch(feat, val, n_features: int, weighted_num_samples: float, num_splits: int):
    """:func:`find_splits` with the sorting and grouping done by torch on the tensors' device (the sampled entries
    never leave the GPU); only features with more than ``num_splits`` + 1 distinct values are walked on the host.
    Returns the same four numpy arrays."""
    import torch
    dev = feat.device
    keep = val != 0.0
    f, v = feat[keep].long(), val[keep].to(torch.float64)
    n_splits = torch.zeros(n_features, dtype=torch.int64, device=dev)
    boz = torch.zeros(n_features, dtype=torch.int64, device=dev)
    if f.numel() == 0:
        return (np.zeros(n_features + 1, np.int32), np.zeros(0, np.float64), np.zeros(n_features, np.int32), np.zeros(n_features, np.uint8))

    def sort_fv(ff, vv, *rest):
        o = torch.argsort(vv, stable=True)
        ff, vv, rest = ff[o], vv[o], [r[o] for r in rest]
        o = torch.argsort(ff, stable=True)
        return (ff[o], vv[o], *[r[o] for r in rest])

    f, v = sort_fv(f, v)
    new = torch.ones(f.numel(), dtype=torch.bool, device=dev)
    new[1:] = (f[1:] != f[:-1]) | (v[1:] != v[:-1])
    starts = torch.nonzero(new).flatten()
    uf, uv = f[starts], v[starts]
    cnt = torch.diff(torch.cat([starts, torch.tensor([f.numel()], device=dev)])).to(torch.float64)
    fnew = torch.ones(uf.numel(), dtype=torch.bool, device=dev)
    fnew[1:] = uf[1:] != uf[:-1]
    gid = torch.cumsum(fnew, 0) - 1
    fids = uf[fnew]
    part = torch.zeros(fids.numel(), dtype=torch.float64, device=dev).index_add_(0, gid, cnt)
    need_zero = (weighted_num_samples - part) > (EPSILON * part * 100)
    af = torch.cat([uf, fids[need_zero]])
    av = torch.cat([uv, torch.zeros(int(need_zero.sum()), dtype=torch.float64, device=dev)])
    ac = torch.cat([cnt, weighted_num_samples - part[need_zero]])
    af, av, ac = sort_fv(af, av, ac)
    gnew = torch.ones(af.numel(), dtype=torch.bool, device=dev)
    gnew[1:] = af[1:] != af[:-1]
    g2 = torch.cumsum(gnew, 0) - 1
    gstart = torch.nonzero(gnew).flatten()
    gsize = torch.bincount(g2)
    possible = gsize - 1
    mids = torch.zeros(af.numel(), dtype=torch.float64, device=dev)
    mids[1:] = (av[:-1] + av[1:]) / 2.0
    simple = (possible <= num_splits)[g2] & ~gnew                      # midpoint rows of the simple features
    n_splits.index_add_(0, af[simple], torch.ones(int(simple.sum()), dtype=torch.int64, device=dev))
    hard = {}
    for gi in torch.nonzero(possible > num_splits).flatten().tolist():    # Spark's stride walk, on the host
        a0, b0 = int(gstart[gi]), int(gstart[gi] + gsize[gi])
        values, counts = av[a0:b0].cpu().numpy(), ac[a0:b0].cpu().numpy()
        stride = weighted_num_samples / (num_splits + 1)
        out, current, target = [], counts[0], stride
        for i in range(1, values.size):
            previous = current
            current = current + counts[i]
            if abs(previous - target) < abs(current - target):
                out.append((values[i - 1] + values[i]) / 2.0)
                target += stride
        hard[int(af[a0])] = np.asarray(out, dtype=np.float64)
        n_splits[int(af[a0])] = len(out)
    thr_ptr = torch.zeros(n_features + 1, dtype=torch.int64, device=dev)
    thr_ptr[1:] = torch.cumsum(n_splits, 0)
    thr = torch.zeros(int(thr_ptr[-1]), dtype=torch.float64, device=dev)
    sf_, sm_ = af[simple], mids[simple]
    if sf_.numel():
        first = torch.ones(sf_.numel(), dtype=torch.bool, device=dev)
        first[1:] = sf_[1:] != sf_[:-1]
        fpos = torch.nonzero(first).flatten()
        within = torch.arange(sf_.numel(), device=dev) - fpos[torch.cumsum(first, 0) - 1]
        thr[thr_ptr[sf_] + within] = sm_
        boz.index_add_(0, sf_[sm_ < 0.0], torch.ones(int((sm_ < 0.0).sum()), dtype=torch.int64, device=dev))
    thr_ptr_np = thr_ptr.cpu().numpy().astype(np.int32)
    thr_np, ns_np, boz_np = thr.cpu().numpy(), n_splits.cpu().numpy().astype(np.int32), boz.cpu().numpy().astype(np.uint8)
    for fid, t in hard.items():
        thr_np[thr_ptr_np[fid]:thr_ptr_np[fid + 1]] = t
        boz_np[fid] = int(np.searchsorted(t, 0.0, side="left"))
    return thr_ptr_np, thr_np, ns_np, boz_np


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


def _gini_rows(stats: np.ndarray) -> np.ndarray:
    """:func:`gini` over the rows of ``stats`` [m, 2], same operation order (1 - f0*f0 - f1*f1, 0 when the row is empty)."""
    total = stats[:, 0] + stats[:, 1]
    safe = np.where(total == 0.0, 1.0, total)
    f0, f1 = stats[:, 0] / safe, stats[:, 1] / safe
    return np.where(total == 0.0, 0.0, (1.0 - f0 * f0) - f1 * f1)


def trees_from_levels(n_trees: int, tree: np.ndarray, heap: np.ndarray, stats: np.ndarray, split: np.ndarray, feature: np.ndarray,
                      threshold: np.ndarray, gain: np.ndarray, prune: bool) -> List[sf.TreeArrays]:
    """``LearningNode.toNode(prune)`` for a whole forest at once.  One row per grown node: its tree, heap id (root 1,
    children 2n / 2n+1), class counts, and -- where ``split`` -- the split it received.  With ``prune`` a split whose
    two (pruned) children are leaves predicting the same class collapses into a leaf, bottom-up.  Returns each tree
    in Spark's pre-order numbering."""
    tree = np.asarray(tree, dtype=np.int64)
    heap = np.asarray(heap, dtype=np.int64)
    n = tree.size
    depth = np.zeros(n, dtype=np.int64)
    tmp = heap.copy()
    while True:
        tmp >>= 1
        more = tmp > 0
        if not more.any():
            break
        depth += more
    maxd = int(depth.max()) if n else 0
    key = (tree << 32) | heap
    by_key = np.argsort(key, kind="stable")
    skey = key[by_key]

    def index_of(t, h):  # row index of node (t, h); every queried node exists
        return by_key[np.searchsorted(skey, (t << 32) | h)]

    pred = (stats[:, 1] > stats[:, 0]).astype(np.int64)
    leaf = ~np.asarray(split, dtype=bool)
    if prune:
        for d in range(maxd - 1, -1, -1):
            cand = np.nonzero((depth == d) & ~leaf)[0]
            if cand.size == 0:
                continue
            li, ri = index_of(tree[cand], 2 * heap[cand]), index_of(tree[cand], 2 * heap[cand] + 1)
            leaf[cand[leaf[li] & leaf[ri] & (pred[li] == pred[ri])]] = True
    kept = depth == 0
    for d in range(1, maxd + 1):
        cand = np.nonzero(depth == d)[0]
        if cand.size == 0:
            continue
        par = index_of(tree[cand], heap[cand] >> 1)
        kept[cand[kept[par] & ~leaf[par]]] = True
    rows = np.nonzero(kept)[0]
    # pre-order = paths in lexicographic order with an ancestor before its descendants: heap ids left-aligned, then depth
    aligned = heap[rows] << (maxd - depth[rows])
    rows = rows[np.lexsort((depth[rows], aligned, tree[rows]))]
    t_sorted = tree[rows]
    starts = np.searchsorted(t_sorted, np.arange(n_trees + 1))
    pos = np.empty(n, dtype=np.int64)
    pos[rows] = np.arange(rows.size) - starts[t_sorted]
    inner = rows[~leaf[rows]]
    left = np.full(n, -1, dtype=np.int64)
    right = np.full(n, -1, dtype=np.int64)
    left[inner] = pos[index_of(tree[inner], 2 * heap[inner])]
    right[inner] = pos[index_of(tree[inner], 2 * heap[inner] + 1)]
    imp = _gini_rows(stats)
    out = []
    for t in range(n_trees):
        r = rows[starts[t]:starts[t + 1]]
        lf = leaf[r]
        out.append(sf.TreeArrays(np.where(lf, -1, feature[r]), np.where(lf, 0.0, threshold[r]), left[r], right[r], pred[r].astype(np.float64),
                                 stats[r], imp[r], np.where(lf, -1.0, gain[r])))
    return out


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
    # DecisionTreeMetadata's order: an integer N > 0 means min(N, numFeatures) features; otherwise a fraction in (0, 1]
    # means ceil(fraction * numFeatures); anything else is rejected ("1" is one feature, "1.0" is all of them)
    try:
        n = int(s)
    except ValueError:
        n = None
    if n is not None:
        if n <= 0:
            raise ValueError(f"unsupported featureSubsetStrategy '{strategy}': an integer must be > 0")
        return min(n, n_features)
    try:
        x = float(s)
    except ValueError:
        raise ValueError(f"unsupported featureSubsetStrategy '{strategy}'")
    if not (0.0 < x <= 1.0):
        raise ValueError(f"unsupported featureSubsetStrategy '{strategy}': a fraction must be in (0, 1]")
    return int(math.ceil(x * n_features))


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


def default_subset_batch(seed: int, n_features: int, k: int, device):
    """Default subsets for a whole node group at once, drawn ON THE DEVICE: ``fn(depth, first, m) -> int64 [m, k]``
    (k distinct features per node, in order of first appearance).  Every rank calls it with the same arguments
    and the same seeded generator, so the subsets agree across ranks without any exchange."""
    import torch

    def fn(depth: int, first: int, m: int):
        if k >= n_features:
            return torch.arange(n_features, device=device).expand(m, n_features)
        g = torch.Generator(device=device)
        g.manual_seed(((seed & 0x7FFFFFFF) * 1000003 + depth * 7919 + first) & 0x7FFFFFFFFFFFFFFF)
        width = int(k * 1.15) + 16
        while True:
            draws = torch.randint(0, n_features, (m, width), generator=g, device=device)
            srt, order = torch.sort(draws, dim=1, stable=True)
            dup_sorted = torch.zeros_like(draws, dtype=torch.bool)
            dup_sorted[:, 1:] = srt[:, 1:] == srt[:, :-1]
            dup = torch.zeros_like(dup_sorted).scatter_(1, order, dup_sorted)  # later occurrences are the duplicates
            keep = ~dup
            keep &= torch.cumsum(keep, dim=1) <= k
            if bool((keep.sum(dim=1) == k).all()):
                return draws[keep].view(m, k)
            width *= 2  # vanishingly rare: some node drew fewer than k distinct features

    return fn


def _seg_len(nodes_per_tree: float, row_nnz: float, subset_fraction: float) -> int:
    """Column entries per work unit of a level whose trees have ``nodes_per_tree`` open nodes.  The column walks gather
    state[tree, row] at scattered rows; the ~9,500 units in flight should belong to one or two trees, whose state rows then
    stay in the L2.  A tree's walk visits about nodes x nnz x (features per node / features) entries, so near the root the
    units have to be short for one tree to fill the machine; deep levels keep SEG_LEN."""
    want = nodes_per_tree * row_nnz * subset_fraction * 850.0
    seg = 256
    while seg * 2 <= want and seg < SEG_LEN:
        seg *= 2
    return seg


def _units_device(col_len, feats, seg_len: int = SEG_LEN):
    """Device version of :func:`_units` -> (unit_item, unit_seg, seg_len)."""
    import torch
    SEG_LEN = int(seg_len)
    nseg = (col_len[feats] + SEG_LEN - 1) // SEG_LEN
    unit_item = torch.repeat_interleave(torch.arange(feats.numel(), dtype=torch.int32, device=feats.device), nseg)
    first = torch.cumsum(nseg, 0) - nseg
    unit_seg = (torch.arange(unit_item.numel(), dtype=torch.int64, device=feats.device) - torch.repeat_interleave(first, nseg)).to(torch.int32)
    return unit_item, unit_seg, SEG_LEN


# ------------------------------------------------------------------------------------------ backend
class CudaTreeBackend:
    """Thin adapter: torch CUDA tensors in, C-ABI calls (fd_bin_csr, fd_rf_*) on the current stream."""
    lazy_default = True   # rows without a split feature are not rewritten; items / splits carry chain lengths instead

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

    def _csc_scratch(self, nnz, n_features) -> int:
        need = int(self.h.lib.fd_csc_scratch_bytes(self.h.h, nnz, n_features))
        if need < 0:
            self.h.check(need)
        return need

    def csc_build(self, rowptr, idx, bins, labels, n_features: int):
        """-> (rowidx i32 [nnz] = row | label << 31, bin u8 [nnz]) ordered by (feature, row): fd_csc_build."""
        t = self.torch
        nnz = idx.numel()
        rowidx = t.empty(nnz, dtype=t.int32, device=self.device)
        out_bin = t.empty(nnz, dtype=t.uint8, device=self.device)
        scratch = t.empty(self._csc_scratch(nnz, int(n_features)), dtype=t.uint8, device=self.device)  # (from torch's caching allocator)
        self.h.call("fd_csc_build", _lib.ptr(rowptr), _lib.ptr(idx), _lib.ptr(bins), _lib.ptr(labels), rowptr.numel() - 1, nnz, int(n_features),
                    _lib.ptr(rowidx), _lib.ptr(out_bin), _lib.ptr(scratch), scratch.numel(), self._st())
        return rowidx, out_bin

    def bag_poisson(self, seed: int, rate: float, n_trees: int, label):
        """-> (bagw u8 [n_trees, n_rows], root i64 [n_trees, 2]): Poisson(rate) bootstrap weights and the weighted class
        counts of this rank's rows, one kernel (fd_rf_bag_poisson)."""
        t = self.torch
        n = label.numel()
        bagw = t.empty((n_trees, n), dtype=t.uint8, device=self.device)
        root = t.empty((n_trees, 2), dtype=t.int64, device=self.device)
        self.h.call("fd_rf_bag_poisson", C.c_uint64(seed & 0xFFFFFFFFFFFFFFFF), C.c_double(rate), int(n_trees), n, _lib.ptr(label),
                    _lib.ptr(bagw), _lib.ptr(root), self._st())
        return bagw, root

    def histogram(self, csc, state, items, units):
        """csc = (colptr i64, rowidx i32 with the row's label in bit 31, bin u8); state i32 [n_trees, n_rows] =
        node_id << 8 | bag weight."""
        colptr, rowidx, bins = csc
        it_t, it_n, it_f = items[0], items[1], items[2]
        it_j = items[3] if len(items) > 3 else None   # chain lengths (None: every row's state names its node)
        u_item, u_seg = units[0], units[1]
        seg_len = int(units[2]) if len(units) > 2 else SEG_LEN
        n_items = it_t.numel()
        hist = self.torch.empty((n_items, 32, 2), dtype=self.torch.int32, device=self.device)
        self.h.call("fd_rf_histogram", _lib.ptr(colptr), _lib.ptr(rowidx), _lib.ptr(bins), _lib.ptr(state), state.shape[1], n_items,
                    _lib.ptr(it_t), _lib.ptr(it_n), _lib.ptr(it_f), _lib.ptr(it_j) if it_j is not None else None, u_item.numel(), _lib.ptr(u_item),
                    _lib.ptr(u_seg), seg_len,
                    _lib.ptr(hist), self._st())
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

    def advance(self, csc, state, level_base, splits, units, action):
        colptr, rowidx, bins = csc
        s_t, s_n, s_f, s_b = splits[0], splits[1], splits[2], splits[3]
        s_j = splits[4] if len(splits) > 4 else None  # chain lengths: with them the default moves need no pass (action unused)
        u_item, u_seg = units[0], units[1]
        seg_len = int(units[2]) if len(units) > 2 else SEG_LEN
        self.h.call("fd_rf_advance", _lib.ptr(colptr), _lib.ptr(rowidx), _lib.ptr(bins), _lib.ptr(state), state.shape[1],
                    state.shape[0], int(level_base), s_t.numel(), _lib.ptr(s_t), _lib.ptr(s_n), _lib.ptr(s_f), _lib.ptr(s_b),
                    _lib.ptr(s_j) if s_j is not None else None, u_item.numel(), _lib.ptr(u_item), _lib.ptr(u_seg), seg_len,
                    _lib.ptr(action) if s_j is None else None, self._st())


# ------------------------------------------------------------------------------------------ trainer
def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


def _units(col_len: np.ndarray, feats: np.ndarray, seg_len: int = SEG_LEN) -> Tuple[np.ndarray, np.ndarray]:
    """One work unit per ``seg_len`` entries of each item's column -> (unit_item int32[], unit_seg int32[])."""
    SEG_LEN = int(seg_len)
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
        self.profile = PROFILE_DEFAULT  # when set, phases are timed with a device synchronise on both sides (slower)
        self._t: Dict[str, float] = {}

    def _tick(self, name: str, t0: float) -> float:
        import time
        if self.profile:
            self.b.torch.cuda.synchronize() if hasattr(self.b, "torch") else None
            now = time.perf_counter()
            self._t[name] = self._t.get(name, 0.0) + now - t0
            return now
        return t0

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
            # Bernoulli(fraction) row sample drawn where the rows live; only the picked rows' entries are gathered
            g = torch.Generator(device=rowptr.device)
            g.manual_seed(self.seed + 7919 * (rank + 1))
            rows = torch.nonzero(torch.rand(n_local, generator=g, device=rowptr.device) < fraction).flatten()
            cnt = rowptr[rows + 1] - rowptr[rows]
            start = torch.repeat_interleave(rowptr[rows], cnt)
            first = torch.repeat_interleave(torch.cumsum(cnt, 0) - cnt, cnt)
            pos = start + (torch.arange(int(cnt.sum()), device=rowptr.device) - first)
            sel_f, sel_v = idx[pos], val[pos]
        if dist is not None and dist.get_world_size() > 1:
            # every rank needs the SAME sample.  The exchange uses all-reduce only (the collective the histograms use, so
            # no second set of channels has to be set up): sizes first, then every rank adds its entries into its own slice
            # of a zero buffer -- sums with zeros are exact, feature ids are exact in fp64
            ws = dist.get_world_size()
            sizes = torch.zeros(ws, dtype=torch.int64, device=sel_f.device)
            sizes[rank] = sel_f.numel()
            dist.all_reduce(sizes)
            sz = sizes.cpu().tolist()
            off = int(sum(sz[:rank]))
            buf = torch.zeros((int(sum(sz)), 2), dtype=torch.float64, device=sel_f.device)
            buf[off:off + sel_f.numel(), 0] = sel_f.to(torch.float64)
            buf[off:off + sel_f.numel(), 1] = sel_v
            dist.all_reduce(buf)
            sel_f, sel_v = buf[:, 0].to(torch.int32), buf[:, 1].contiguous()
        num_splits = min(self.max_bins, n_total) - 1
        return find_splits_torch(sel_f, sel_v, self.F, fraction * n_total, num_splits)

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
        import time
        self._t = {}
        tt = self._tick("start", time.perf_counter())
        thr_ptr, thr, n_splits, bin_of_zero = thresholds if thresholds is not None else self._thresholds(rowptr, idx, val, n_total, rank)
        tt = self._tick("thresholds", tt)
        d_thr_ptr = torch.from_numpy(thr_ptr).to(dev)
        d_thr = torch.from_numpy(thr).to(dev)
        d_nsplits = torch.from_numpy(n_splits).to(dev)
        d_bz = torch.from_numpy(bin_of_zero).to(dev)
        bins = self.b.bin_csr(idx, val, d_thr_ptr, d_thr)
        labels = labels.to(dev).to(torch.uint8).contiguous()
        if self.max_depth > 22:
            raise ValueError("maxDepth > 22 is not supported (node ids share a 32-bit word with the bag weight)")
        # column entries carry (row | label << 31): the histogram kernel then needs a single gather per entry
        colptr = torch.zeros(self.F + 1, dtype=torch.int64, device=dev)
        torch.cumsum(torch.bincount(idx, minlength=self.F), 0, out=colptr[1:])
        if hasattr(self.b, "csc_build"):
            rowidx, cbins = self.b.csc_build(rowptr, idx, bins, labels, self.F)
        else:  # stand-in backends of the CPU orchestration tests
            counts = rowptr[1:] - rowptr[:-1]
            row_of = torch.repeat_interleave(torch.arange(n_local, dtype=torch.int32, device=dev), counts)
            order = torch.argsort(idx, stable=True)
            row_lab = row_of | torch.repeat_interleave(labels.to(torch.int32) << 31, counts)
            rowidx, cbins = row_lab[order].contiguous(), bins[order].contiguous()
            del order, row_of, row_lab
        csc = (colptr, rowidx, cbins)
        col_len = np.diff(colptr.cpu().numpy())
        del bins
        tt = self._tick("bin+csc", tt)
        # ---- bagging
        root = None
        if bag_weights is None:
            if self.T == 1:
                bag_weights = torch.ones((1, n_local), dtype=torch.uint8, device=dev)
            elif hasattr(self.b, "bag_poisson"):
                # BaggedPoint.convertToBaggedRDD: Poisson(subsamplingRate) per (tree, row), drawn on the device
                bag_weights, root = self.b.bag_poisson((self.seed + 1) * 1000003 + rank, self.rate, self.T, labels)
                root = root.to(torch.float64)
            else:  # stand-in backends of the CPU orchestration tests
                g = torch.Generator(device=dev)
                g.manual_seed(self.seed + rank + 1)
                lam = torch.full((self.T, n_local), self.rate, dtype=torch.float32, device=dev)
                bag_weights = torch.poisson(lam, generator=g).clamp_(max=255).to(torch.uint8)
        bagw = bag_weights.to(dev).to(torch.uint8).contiguous()
        state = bagw.to(torch.int32) | (1 << 8)  # per (tree, row): node_id << 8 | bag weight; every row starts in the root
        if root is None:
            lab1 = labels.to(torch.bool)
            w1 = (bagw * lab1).sum(dim=1, dtype=torch.int64)
            root = torch.stack([bagw.sum(dim=1, dtype=torch.int64) - w1, w1], 1).to(torch.float64)  # [T, 2]
        if dist is not None:
            dist.all_reduce(root)
        root = root.cpu().numpy()
        # histogram cells are 32-bit and summed over ranks as int32: no cell can exceed a tree's total bag weight
        if root.size and float(root.sum(axis=1).max()) >= 2.0 ** 31:
            raise ValueError("the total bag weight of a tree reaches 2^31: histogram cells (int32, all-reduced) could overflow; "
                             "train on fewer rows per forest or lower subsamplingRate")
        k = features_per_node(self.strategy, self.T, self.F)
        batch_fn = default_subset_batch(self.seed, self.F, k, dev) if subset_fn is None else None
        d_col_len = csc[0][1:] - csc[0][:-1]
        tt = self._tick("bagging+root", tt)
        # ---- level-wise growth.  Host bookkeeping is array-valued: one record per grown node, kept per level
        #      (tree, heap id, class counts, and -- once it splits -- feature / threshold / gain)
        lt = np.arange(self.T, dtype=np.int32)                  # this level's nodes: tree ...
        ln = np.ones(self.T, dtype=np.int32)                    # ... heap id (root 1) ...
        ls = np.ascontiguousarray(root, dtype=np.float64)       # ... class counts [m, 2]
        # ... and, for backends that do not rewrite the rows that follow a default child (lazy_default), how many of the
        # node's nearest ancestors lead to it by default links only
        lazy = bool(getattr(self.b, "lazy_default", False))
        lj = np.zeros(self.T, dtype=np.int32)
        levels = []
        hist_bytes = allreduce_calls = 0
        for depth in range(self.max_depth + 1):
            m_lvl = lt.size
            l_split = np.zeros(m_lvl, dtype=bool)
            l_feat = np.full(m_lvl, -1, dtype=np.int32)
            l_thr = np.zeros(m_lvl, dtype=np.float64)
            l_gain = np.full(m_lvl, -1.0, dtype=np.float64)
            levels.append((lt, ln, ls, l_split, l_feat, l_thr, l_gain))
            if depth == self.max_depth or m_lvl == 0:
                break
            base = 1 << depth
            aidx = np.arange(m_lvl) if depth == 0 else np.nonzero(np.abs(_gini_rows(ls)) >= EPSILON)[0]
            if aidx.size == 0:
                break
            sp_parts, child_parts = [], []
            # work-unit length of this level's column walks (see _seg_len)
            open_trees = max(1, int(np.unique(lt[aidx]).size))
            seg_len = _seg_len(aidx.size / open_trees, float(idx.numel()) / max(1, n_local), min(1.0, k / float(self.F)))
            g0 = 0
            while g0 < aidx.size:
                # ---- one node group: items = node x feature subset (features without thresholds dropped), on the device
                g1 = g0
                if batch_fn is not None:
                    per_node = max(1, min(k, self.F))
                    g1 = min(aidx.size, g0 + max(1, MAX_ITEMS_PER_GROUP // per_node))
                    gi = aidx[g0:g1]
                    grp_t = torch.from_numpy(lt[gi]).to(dev)
                    grp_n = torch.from_numpy(ln[gi]).to(dev)
                    sub = batch_fn(depth, g0, g1 - g0)                           # [m, k]
                    ok = d_nsplits[sub] > 0
                    cnt = ok.sum(dim=1)
                    items = (torch.repeat_interleave(grp_t, cnt), torch.repeat_interleave(grp_n, cnt), sub[ok].to(torch.int32))
                    if lazy:
                        items = items + (torch.repeat_interleave(torch.from_numpy(lj[gi]).to(dev), cnt),)
                    d_ptr = torch.zeros(g1 - g0 + 1, dtype=torch.int32, device=dev)
                    d_ptr[1:] = torch.cumsum(cnt, 0)
                else:
                    it_t, it_n, it_f, it_j, ptr = [], [], [], [], [0]
                    while g1 < aidx.size and (g1 == g0 or ptr[-1] < MAX_ITEMS_PER_GROUP):
                        t, n = int(lt[aidx[g1]]), int(ln[aidx[g1]])
                        feats = np.asarray(subset_fn(t, n), dtype=np.int32)
                        feats = feats[n_splits[feats] > 0]
                        it_t.append(np.full(feats.size, t, np.int32))
                        it_n.append(np.full(feats.size, n, np.int32))
                        it_f.append(feats)
                        it_j.append(np.full(feats.size, lj[aidx[g1]], np.int32))
                        ptr.append(ptr[-1] + feats.size)
                        g1 += 1
                    gi = aidx[g0:g1]
                    items = tuple(torch.from_numpy(np.concatenate(x)).to(dev) for x in ((it_t, it_n, it_f, it_j) if lazy else (it_t, it_n, it_f)))
                    d_ptr = torch.tensor(ptr, dtype=torch.int32, device=dev)
                g0 = g1
                tt = self._tick("subsets+items", tt)
                if items[2].numel() == 0:
                    continue
                units = _units_device(d_col_len, items[2].long(), seg_len)
                tt = self._tick("units", tt)
                hist = self.b.histogram(csc, state, items, units)
                tt = self._tick("kernel:histogram", tt)
                if dist is not None and dist.get_world_size() > 1:
                    dist.all_reduce(hist)  # integer counts: exact, order-independent
                    allreduce_calls += 1
                hist_bytes += hist.numel() * 4
                tt = self._tick("allreduce", tt)
                totals = torch.from_numpy(np.ascontiguousarray(ls[gi])).to(dev)
                bi, bs, bg, left, right = self.b.best_split(hist, d_ptr, items[2], totals, d_nsplits, d_bz, self.min_instances,
                                                            self.min_info_gain)
                bf = items[2][bi.clamp(min=0).long()]  # feature of the winning item
                bi, bs, bg, left, right, bf = (x.cpu().numpy() for x in (bi, bs, bg, left, right, bf))
                del hist
                tt = self._tick("kernel:best_split+d2h", tt)
                won = (bi >= 0) & (bg > 0.0)             # the others stay leaves
                wi = gi[won]
                f, sb = bf[won].astype(np.int32), bs[won].astype(np.int32)
                l_split[wi] = True
                l_feat[wi] = f
                l_thr[wi] = thr[thr_ptr[f] + sb]
                l_gain[wi] = bg[won]
                sp_parts.append(np.stack([lt[wi], ln[wi], f, sb], 1).astype(np.int32))
                child_parts.append((lt[wi], ln[wi], left[won], right[won], lj[wi]))
            tt = self._tick("host:grow", tt)
            sp = np.concatenate(sp_parts) if sp_parts else np.zeros((0, 4), np.int32)
            if sp.shape[0] == 0:
                break
            action = np.full((self.T, base), -1, dtype=np.int32)
            action[sp[:, 0], sp[:, 1] - base] = np.where(bin_of_zero[sp[:, 2]] <= sp[:, 3], 2 * sp[:, 1], 2 * sp[:, 1] + 1)
            d_sp = tuple(torch.from_numpy(np.ascontiguousarray(sp[:, c])).to(dev) for c in range(4))
            pj = np.concatenate([c[4] for c in child_parts]).astype(np.int32)   # the split nodes' own chain lengths
            if lazy:
                d_sp = d_sp + (torch.from_numpy(pj).to(dev),)
            adv_seg = _seg_len(sp.shape[0] / max(1, int(np.unique(sp[:, 0]).size)), 1.0, 1.0 / 16.0)  # (a split feature's column: ~1/16 of the rows)
            units = _units_device(d_col_len, d_sp[2].long(), adv_seg)
            self.b.advance(csc, state, base, d_sp, units, None if lazy else torch.from_numpy(action).to(dev))
            # next level: the two children of every split node, left then right, in split order
            ct = np.concatenate([c[0] for c in child_parts])
            cn = np.concatenate([c[1] for c in child_parts])
            cl = np.concatenate([c[2] for c in child_parts])
            cr = np.concatenate([c[3] for c in child_parts])
            lt = np.repeat(ct, 2).astype(np.int32)
            ln = np.stack([2 * cn, 2 * cn + 1], 1).reshape(-1).astype(np.int32)
            ls = np.stack([cl, cr], 1).reshape(-1, 2).astype(np.float64)
            dflt = action[sp[:, 0], sp[:, 1] - base]                             # each split node's default child
            lj = np.stack([np.where(dflt == 2 * cn, pj + 1, 0), np.where(dflt == 2 * cn + 1, pj + 1, 0)], 1).reshape(-1).astype(np.int32)
            tt = self._tick("kernel:advance", tt)
        cat = [np.concatenate([lv[c] for lv in levels]) for c in range(7)]
        trees = trees_from_levels(self.T, cat[0], cat[1], cat[2], cat[3], cat[4], cat[5], cat[6], self.prune)
        tt = self._tick("host:assemble trees", tt)
        self.stats = {"phases_s": dict(self._t), "hist_bytes": float(hist_bytes), "allreduce_calls": float(allreduce_calls), "n_total": float(n_total)}
        return trees
