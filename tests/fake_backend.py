"""A numpy stand-in for the CUDA tree backend -- TEST INFRASTRUCTURE ONLY.  It lets the gloo tests drive the
product's multi-rank training orchestration (training.ForestTrainer) on CPU; the arithmetic it performs is
the oracle's (oracle/spark_oracle.py)."""
import numpy as np
import torch

from oracle import spark_oracle as so


class NumpyTreeBackend:
    device = torch.device("cpu")

    def bin_csr(self, idx, val, thr_ptr, thr):
        tp, th = thr_ptr.numpy(), thr.numpy()
        out = np.zeros(idx.numel(), dtype=np.uint8)
        for k, (f, v) in enumerate(zip(idx.numpy(), val.numpy())):
            out[k] = so.bin_value(th[tp[f]:tp[f + 1]], float(v))
        return torch.from_numpy(out)

    def histogram(self, csc, state, items, units):
        colptr, rowidx, bins = (x.numpy() for x in csc)
        st = state.numpy()
        it_t, it_n, it_f = (x.numpy() for x in items)
        hist = np.zeros((it_t.size, 32, 2), dtype=np.int32)
        for i, (t, n, f) in enumerate(zip(it_t, it_n, it_f)):
            for e in range(colptr[f], colptr[f + 1]):
                r, lab = int(rowidx[e]) & 0x7FFFFFFF, (int(rowidx[e]) >> 31) & 1
                if st[t, r] >> 8 == n:
                    hist[i, bins[e], lab] += int(st[t, r] & 255)
        return torch.from_numpy(hist)

    def best_split(self, hist, node_item_ptr, item_feat, node_total, n_splits, bin_of_zero, min_instances, min_info_gain):
        h, ptr, feats, tot = hist.numpy().astype(np.float64), node_item_ptr.numpy(), item_feat.numpy(), node_total.numpy()
        ns, bz = n_splits.numpy(), bin_of_zero.numpy()
        n = ptr.size - 1
        bi = np.full(n, -1, np.int32)
        bs = np.full(n, -1, np.int32)
        bg = np.full(n, -1.7976931348623157e308)
        left = np.zeros((n, 2))
        right = np.zeros((n, 2))
        for j in range(n):
            a, b = ptr[j], ptr[j + 1]
            if a == b:
                continue
            hh = h[a:b].copy()
            for i in range(b - a):  # implied zero-valued rows
                hh[i, bz[feats[a + i]], :] += tot[j] - hh[i].sum(axis=0)
            gain, jj, s, lc, rc, _ = so.best_split_for_node(hh, [ns[f] for f in feats[a:b]], list(feats[a:b]), min_instances, min_info_gain)
            if jj >= 0:
                bi[j], bs[j], bg[j], left[j], right[j] = a + jj, s, gain, lc, rc
        return tuple(torch.from_numpy(x) for x in (bi, bs, bg, left, right))

    def advance(self, csc, state, level_base, splits, units, action):
        colptr, rowidx, bins = (x.numpy() for x in csc)
        st = state.numpy()
        act = action.numpy()
        for t, n, f, sb in zip(*(x.numpy() for x in splits)):
            for e in range(colptr[f], colptr[f + 1]):
                r = int(rowidx[e]) & 0x7FFFFFFF
                if st[t, r] >> 8 == n:
                    st[t, r] = ((2 * n if bins[e] <= sb else 2 * n + 1) << 8) | (st[t, r] & 255)
        T, N = st.shape
        for t in range(T):
            for r in range(N):
                n = st[t, r] >> 8
                if level_base <= n < 2 * level_base and act[t, n - level_base] >= 0:
                    st[t, r] = (act[t, n - level_base] << 8) | (st[t, r] & 255)
