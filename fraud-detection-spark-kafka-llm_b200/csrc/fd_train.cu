// fd_train.cu -- random-forest training kernels: TreePoint binning, per-node class-count histograms over
// CSC columns (shared-memory atomics), Gini best-split selection, and the row -> child update.
//
// Restates ml.tree.impl.RandomForest (Spark 3.5.5) reached through Pipeline.fit at
// fraud_detection_spark.py:85-91 with the classifiers built at :59-74.  Spark densifies every row to
// numFeatures bins; here rows stay sparse: a histogram pass visits, for every (node, feature-in-subset)
// item, only that feature's COLUMN (CSC over this rank's row shard) and keeps the entries whose row
// currently sits in the node.  Rows without the feature are accounted for after the cross-rank sum:
// hist[bin_of_zero] += node_total - sum(explicit), which is exact because all counts are integers.
#include <cub/cub.cuh>

#include "fd_common.cuh"

namespace fd {

constexpr int HB = 32;  // bins per feature (maxBins)

// ---- TreePoint.findBin: bin = lower_bound(thresholds of the feature, value) --------------------------
__global__ void k_bin_csr(const int32_t* __restrict__ idx, const double* __restrict__ val, long long nnz,
                          const int32_t* __restrict__ thr_ptr, const double* __restrict__ thr, uint8_t* __restrict__ out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride) {
        const int f = idx[k];
        const double v = val[k];
        int lo = thr_ptr[f], hi = thr_ptr[f + 1];
        const int base = lo;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (__ldg(thr + mid) < v) lo = mid + 1;
            else hi = mid;
        }
        out[k] = (uint8_t)(lo - base);
    }
}

// ---- CSR -> CSC for the trainer: column entries (row | label << 31, bin), columns in feature order ------------
// One warp per row packs its entries as (row | label << 31) << 8 | bin; a key-limited radix sort by feature id (CUB, only
// the bits a feature id can have) brings every column together with its rows ascending; the sorted values are unpacked.
__global__ void k_csc_pack(const int64_t* __restrict__ rowptr, const uint8_t* __restrict__ label, const uint8_t* __restrict__ bin,
                           long long n_rows, unsigned long long* __restrict__ packed) {
    const int lane = threadIdx.x & 31;
    const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long r = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < n_rows; r += nwarps) {
        const long long a = rowptr[r], b = rowptr[r + 1];
        const unsigned long long head = ((unsigned long long)(uint32_t)r | ((unsigned long long)(label[r] ? 1u : 0u) << 31)) << 8;
        for (long long k = a + lane; k < b; k += 32) packed[k] = head | (unsigned long long)bin[k];
    }
}
__global__ void k_csc_unpack(const unsigned long long* __restrict__ packed, long long nnz, uint32_t* __restrict__ rowidx, uint8_t* __restrict__ bin) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride) {
        const unsigned long long v = packed[k];
        rowidx[k] = (uint32_t)(v >> 8);
        bin[k] = (uint8_t)v;
    }
}

// ---- IDF.fit, counting part: docFreq[j] = number of dialogues with a value > 0 at feature j ----------------
// (a dialogue holds a feature at most once in CSR, so every stored entry with a positive value is one document)
__global__ void k_doc_freq(const int32_t* __restrict__ idx, const double* __restrict__ val, long long nnz, unsigned long long* __restrict__ df) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride)
        if (val[k] > 0.0) atomicAdd(df + idx[k], 1ULL);
}
// The same with `copies` private 32-bit count vectors (CTA b adds into copy b % copies) and a sum over the copies: a
// frequent term is in every second dialogue, and millions of atomic adds to ONE address run one after the other
// (measured: 20.6 ms for 753M entries with a single vector, against the 1.4 ms its 9 GB of reads take).
__global__ void k_doc_freq_private(const int32_t* __restrict__ idx, const double* __restrict__ val, long long nnz, unsigned int* __restrict__ part,
                                   int copies, int num_features) {
    unsigned int* mine = part + (size_t)(blockIdx.x % copies) * num_features;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride)
        if (__ldcs(val + k) > 0.0) atomicAdd(mine + __ldcs(idx + k), 1u);
}
__global__ void k_doc_freq_sum(const unsigned int* __restrict__ part, int copies, int num_features, unsigned long long* __restrict__ df) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= num_features) return;
    unsigned long long s = 0;
    for (int c = 0; c < copies; ++c) s += part[(size_t)c * num_features + f];
    df[f] = s;
}

// ---- IDFModel.transform on a CSR matrix already on the device: val[k] *= idf[idx[k]] (one exact fp64 multiply) ----------
__global__ void k_idf_scale(const int32_t* __restrict__ idx, double* __restrict__ val, long long nnz, const double* __restrict__ idf) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride) val[k] = __dmul_rn(val[k], __ldg(idf + idx[k]));
}

// ---- bagging: Poisson(rate) bootstrap weights + weighted class totals of the root ------------------------
// One thread draws 16 consecutive rows of one tree (one 16-byte store).  u = top 32 bits of splitmix64(key) with
// key = seed, tree, row; the draw is the number of CDF thresholds <= u (thresholds precomputed on the host).
constexpr int BAG_CDF = 16;
struct BagCdf {
    uint32_t thr[BAG_CDF];
};
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}
__global__ void __launch_bounds__(256) k_rf_bag_poisson(uint64_t seed, BagCdf cdf, int n_trees, long long n_rows,
                                                        const uint8_t* __restrict__ label, uint8_t* __restrict__ bagw,
                                                        unsigned long long* __restrict__ root) {
    const long long groups = (n_rows + 15) / 16;  // per tree
    const long long total = groups * n_trees;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total + (31 - (total + 31) % 32); g += (long long)gridDim.x * blockDim.x) {
        // (the loop bound is rounded up to whole warps so that the shuffles below always see 32 lanes)
        const bool live = g < total;
        const int t = live ? (int)(g / groups) : 0;
        const long long r0 = live ? (g % groups) * 16 : 0;
        uint32_t w[4] = {0u, 0u, 0u, 0u};
        unsigned int c0 = 0, c1 = 0;
        if (live) {
            const uint64_t tkey = splitmix64(seed ^ ((uint64_t)t * 0xD6E8FEB86659FD93ULL));
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const long long r = r0 + j;
                if (r < n_rows) {
                    const uint32_t u = (uint32_t)(splitmix64(tkey + (uint64_t)r) >> 32);
                    unsigned int k = 0;
                    while (k < BAG_CDF && u >= cdf.thr[k]) ++k;
                    w[j >> 2] |= k << (8 * (j & 3));
                    if (label[r]) c1 += k; else c0 += k;
                }
            }
            uint8_t* dst = bagw + (long long)t * n_rows + r0;
            if (r0 + 16 <= n_rows && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
                *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
            } else {
                for (int j = 0; j < 16 && r0 + j < n_rows; ++j) dst[j] = (uint8_t)(w[j >> 2] >> (8 * (j & 3)));
            }
        }
        // lanes of a warp usually share the tree: reduce when they all do, else add individually
        const int t0 = __shfl_sync(0xFFFFFFFFu, t, 0);
        if (__all_sync(0xFFFFFFFFu, t == t0 || !live)) {
            for (int d = 16; d > 0; d >>= 1) {
                c0 += __shfl_down_sync(0xFFFFFFFFu, c0, d);
                c1 += __shfl_down_sync(0xFFFFFFFFu, c1, d);
            }
            if ((threadIdx.x & 31) == 0) {
                if (c0) atomicAdd(root + 2 * t0, (unsigned long long)c0);
                if (c1) atomicAdd(root + 2 * t0 + 1, (unsigned long long)c1);
            }
        } else if (live) {
            if (c0) atomicAdd(root + 2 * t, (unsigned long long)c0);
            if (c1) atomicAdd(root + 2 * t + 1, (unsigned long long)c1);
        }
    }
}

// ---- cache policy of the column walks ---------------------------------------------------------------------
// A histogram / advance pass streams column entries (read once) and gathers state[tree, row] at scattered rows.  The
// warps in flight work on one or two trees at a time, whose state rows (4 B x rows) fit the L2 -- unless the stream
// pushes them out, and then every gather is a 32-byte DRAM sector (measured: 38 B of DRAM per visited entry).  So the
// stream is loaded evict-first and the state words with an evict-last L2 policy.
__device__ __forceinline__ uint64_t l2_keep_policy() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint32_t ld_keep_u32(const uint32_t* p, uint64_t policy) {
    uint32_t v;
    asm volatile("ld.global.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(policy) : "memory");
    return v;
}
__device__ __forceinline__ void st_keep_u32(uint32_t* p, uint32_t v, uint64_t policy) {
    asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(p), "r"(v), "l"(policy) : "memory");
}

// ---- binSeqOp: one warp per (item, column segment) ---------------------------------------------------
// item = (tree, node, feature).  The warp walks its segment of the feature's column, keeps entries whose
// row is in the node, and adds the row's bag weight to a warp-private shared-memory histogram
// [32 bins][2 classes]; the non-zero cells are then added to the item's global histogram.
// Per entry ONE gather: the row's label rides in bit 31 of the column's row index, and the row's node id and
// bag weight share one word, state[tree, row] = node_id << 8 | bag_weight.
// Rows that do NOT hold a split feature follow the node's default child.  They are not rewritten: their state keeps naming
// the ancestor they last moved to by an explicit entry, and a node of the current level is told how many of its nearest
// ancestors lead to it by default links only (`chain`, kept by the trainer: 0 for the non-default child).  A row whose
// state is one of those ancestors sits in the node -- it was never moved explicitly on the way, or its state would name a
// later node.  Heap numbering: the ancestor d levels up is node >> d.  (This replaces a pass over all of state per level.)
__device__ __forceinline__ bool in_node(uint32_t state_node, uint32_t node, int node_depth, uint32_t chain) {
    const int d = node_depth - (31 - __clz(state_node));
    return (unsigned)d <= chain && (node >> d) == state_node;
}

constexpr int WALK_WARPS = 2;
constexpr int WALK_ILP = 4;    // entries per lane and trip of a column walk  // warps per CTA of the column walks (see the scheduling note below)

// Scheduling: ONE unit per warp and a grid as long as the unit list, not a persistent grid-stride loop.  Units are in
// (tree, node, feature) order and differ in length by three orders of magnitude; under a grid-stride loop the warps
// with short units run far ahead, the units in flight spread over dozens of trees and every state gather becomes a DRAM
// sector (measured: 332 GB read by one launch).  The hardware hands CTAs out in index order, so with small CTAs the
// units in flight are always the next ~9,500 of the list -- one or two trees, whose state rows stay in the L2.
__global__ void __launch_bounds__(WALK_WARPS * 32) k_rf_histogram(const int64_t* __restrict__ colptr, const uint32_t* __restrict__ rowidx,
                                                      const uint8_t* __restrict__ bin, const uint32_t* __restrict__ state,
                                                      long long n_rows, int n_units, const int32_t* __restrict__ unit_item,
                                                      const int32_t* __restrict__ unit_seg, int seg_len,
                                                      const int32_t* __restrict__ item_tree, const int32_t* __restrict__ item_node,
                                                      const int32_t* __restrict__ item_feat, const int32_t* __restrict__ item_chain,
                                                      unsigned int* __restrict__ hist) {
    __shared__ unsigned int sh[WALK_WARPS][HB * 2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned int* h = sh[warp];
    const uint64_t keep = l2_keep_policy();
    const int u = blockIdx.x * WALK_WARPS + warp;
    if (u < n_units) {
        h[lane] = 0u;
        h[32 + lane] = 0u;
        __syncwarp();
        const int item = unit_item[u];
        const int t = item_tree[item], f = item_feat[item];
        const uint32_t node = (uint32_t)item_node[item];
        // a row is in `node` when its state names the node or one of the node's `chain` nearest ancestors (see in_node)
        const uint32_t chain = item_chain ? (uint32_t)item_chain[item] : 0u;
        const int dn = 31 - __clz(node);
        const long long c0 = colptr[f] + (long long)unit_seg[u] * seg_len;
        const long long c1 = min((long long)colptr[f + 1], c0 + (long long)seg_len);
        const uint32_t* st = state + (long long)t * n_rows;
        // WALK_ILP entries per lane and trip: the walk is bound by the latency of the dependent state gather, so the loads of
        // a trip are issued together (entries past the segment's end read its last entry and are masked out)
        for (long long e0 = c0 + lane; e0 < c1; e0 += 32 * WALK_ILP) {
            uint32_t rl[WALK_ILP], s[WALK_ILP];
#pragma unroll
            for (int j = 0; j < WALK_ILP; ++j) rl[j] = __ldcs(rowidx + min(e0 + 32 * j, c1 - 1));
#pragma unroll
            for (int j = 0; j < WALK_ILP; ++j) s[j] = ld_keep_u32(st + (rl[j] & 0x7FFFFFFFu), keep);
#pragma unroll
            for (int j = 0; j < WALK_ILP; ++j) {
                const uint32_t w = s[j] & 255u;
                if (in_node(s[j] >> 8, node, dn, chain) && w && e0 + 32 * j < c1) atomicAdd(&h[(int)__ldcs(bin + e0 + 32 * j) * 2 + (int)(rl[j] >> 31)], w);
            }
        }
        __syncwarp();
        unsigned int* g = hist + (size_t)item * (HB * 2);
        const unsigned int a = h[lane], b = h[32 + lane];
        if (a) atomicAdd(g + lane, a);
        if (b) atomicAdd(g + 32 + lane, b);
    }
}

// ---- binsToBestSplit: one warp per node --------------------------------------------------------------
__device__ __forceinline__ double gini2(double c0, double c1, double total) {
    if (total == 0.0) return 0.0;
    double imp = 1.0;
    const double f0 = c0 / total;
    imp = __dsub_rn(imp, __dmul_rn(f0, f0));
    const double f1 = c1 / total;
    imp = __dsub_rn(imp, __dmul_rn(f1, f1));
    return imp;
}

constexpr double MIN_GAIN = -1.7976931348623157e308;  // Double.MinValue marks an invalid split

__global__ void __launch_bounds__(128) k_rf_best_split(const unsigned int* __restrict__ hist, int n_nodes,
                                                       const int32_t* __restrict__ node_item_ptr, const int32_t* __restrict__ item_feat,
                                                       const double* __restrict__ node_total, const int32_t* __restrict__ n_splits,
                                                       const uint8_t* __restrict__ bin_of_zero, int min_instances, double min_info_gain,
                                                       int32_t* __restrict__ best_item, int32_t* __restrict__ best_split,
                                                       double* __restrict__ best_gain, double* __restrict__ out_left,
                                                       double* __restrict__ out_right) {
    const int lane = threadIdx.x & 31;
    const int node = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (node >= n_nodes) return;
    const double t0 = node_total[2 * node], t1 = node_total[2 * node + 1];
    const double ptotal = __dadd_rn(t0, t1);
    const double pimp = gini2(t0, t1, ptotal);
    double bg = MIN_GAIN, bl0 = 0, bl1 = 0;
    int bi = -1, bs = -1;
    for (int it = node_item_ptr[node] + lane; it < node_item_ptr[node + 1]; it += 32) {
        const int f = item_feat[it];
        const int ns = n_splits[f];
        if (ns == 0) continue;
        const unsigned int* h = hist + (size_t)it * (HB * 2);
        // explicit totals, then the implied zero-valued rows go to bin_of_zero
        unsigned int e0 = 0, e1 = 0;
        for (int b = 0; b <= ns; ++b) {
            e0 += h[2 * b];
            e1 += h[2 * b + 1];
        }
        const double z0 = t0 - (double)e0, z1 = t1 - (double)e1;
        const int bz = bin_of_zero[f];
        double l0 = 0.0, l1 = 0.0, fg = MIN_GAIN, fl0 = 0, fl1 = 0;
        int fs = -1;
        for (int s = 0; s < ns; ++s) {  // cumulative left stats = bins 0..s
            l0 += (double)h[2 * s] + (s == bz ? z0 : 0.0);
            l1 += (double)h[2 * s + 1] + (s == bz ? z1 : 0.0);
            const double r0 = t0 - l0, r1 = t1 - l1;
            const double lc = __dadd_rn(l0, l1), rc = __dadd_rn(r0, r1);
            double g;
            if (lc < (double)min_instances || rc < (double)min_instances) {
                g = MIN_GAIN;
            } else {
                const double tc = __dadd_rn(lc, rc);
                const double lw = lc / tc, rw = rc / tc;
                g = __dsub_rn(__dsub_rn(pimp, __dmul_rn(lw, gini2(l0, l1, lc))), __dmul_rn(rw, gini2(r0, r1, rc)));
                if (g < min_info_gain) g = MIN_GAIN;
            }
            if (g > fg) {  // first maximum over splits
                fg = g;
                fs = s;
                fl0 = l0;
                fl1 = l1;
            }
        }
        if (fs >= 0 && fg > bg) {  // first maximum over this lane's items (ascending item order)
            bg = fg;
            bi = it;
            bs = fs;
            bl0 = fl0;
            bl1 = fl1;
        }
    }
    // warp arg-max: larger gain wins, ties go to the smaller item index (= first in subset order)
    for (int d = 16; d > 0; d >>= 1) {
        const double og = __shfl_down_sync(0xFFFFFFFFu, bg, d);
        const int oi = __shfl_down_sync(0xFFFFFFFFu, bi, d), os = __shfl_down_sync(0xFFFFFFFFu, bs, d);
        const double o0 = __shfl_down_sync(0xFFFFFFFFu, bl0, d), o1 = __shfl_down_sync(0xFFFFFFFFu, bl1, d);
        const bool take = oi >= 0 && (bi < 0 || og > bg || (og == bg && oi < bi));
        if (take) {
            bg = og;
            bi = oi;
            bs = os;
            bl0 = o0;
            bl1 = o1;
        }
    }
    if (lane == 0) {
        best_item[node] = bi;
        best_split[node] = bs;
        best_gain[node] = bi >= 0 ? bg : MIN_GAIN;
        out_left[2 * node] = bl0;
        out_left[2 * node + 1] = bl1;
        out_right[2 * node] = t0 - bl0;
        out_right[2 * node + 1] = t1 - bl1;
    }
}

// ---- move rows to their children ------------------------------------------------------------------------
// explicit entries of the split feature: one warp per (split node, column segment)
__global__ void __launch_bounds__(WALK_WARPS * 32) k_rf_advance_explicit(const int64_t* __restrict__ colptr, const uint32_t* __restrict__ rowidx,
                                                             const uint8_t* __restrict__ bin, uint32_t* __restrict__ state,
                                                             long long n_rows, int n_units, const int32_t* __restrict__ unit_item,
                                                             const int32_t* __restrict__ unit_seg, int seg_len,
                                                             const int32_t* __restrict__ split_tree, const int32_t* __restrict__ split_node,
                                                             const int32_t* __restrict__ split_feat, const int32_t* __restrict__ split_bin,
                                                             const int32_t* __restrict__ split_chain) {
    const int lane = threadIdx.x & 31;
    const uint64_t keep = l2_keep_policy();
    const int u = blockIdx.x * WALK_WARPS + (threadIdx.x >> 5);
    if (u < n_units) {
        const int s = unit_item[u];
        const uint32_t node = (uint32_t)split_node[s];
        const int f = split_feat[s], sb = split_bin[s];
        const uint32_t chain = split_chain ? (uint32_t)split_chain[s] : 0u;
        const int dn = 31 - __clz(node);
        uint32_t* st = state + (long long)split_tree[s] * n_rows;
        const long long c0 = colptr[f] + (long long)unit_seg[u] * seg_len;
        const long long c1 = min((long long)colptr[f + 1], c0 + (long long)seg_len);
        for (long long e0 = c0 + lane; e0 < c1; e0 += 32 * WALK_ILP) {  // (several entries per lane and trip, as in the histogram walk)
            uint32_t r[WALK_ILP], v[WALK_ILP];
#pragma unroll
            for (int j = 0; j < WALK_ILP; ++j) r[j] = __ldcs(rowidx + min(e0 + 32 * j, c1 - 1)) & 0x7FFFFFFFu;
#pragma unroll
            for (int j = 0; j < WALK_ILP; ++j) v[j] = ld_keep_u32(st + r[j], keep);
#pragma unroll
            for (int j = 0; j < WALK_ILP; ++j)
                if (in_node(v[j] >> 8, node, dn, chain) && e0 + 32 * j < c1)
                    st_keep_u32(st + r[j], ((((int)__ldcs(bin + e0 + 32 * j) <= sb) ? 2u * node : 2u * node + 1u) << 8) | (v[j] & 255u), keep);
        }
    }
}

__global__ void k_rf_advance_default(uint32_t* __restrict__ state, long long n_rows, int level_base,
                                     const int32_t* __restrict__ action) {
    // blockIdx.y = tree: no 64-bit division per element, and the tree's action row is read through one pointer
    uint32_t* st = state + (long long)blockIdx.y * n_rows;
    const int32_t* act = action + (long long)blockIdx.y * level_base;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n_rows; k += stride) {
        const uint32_t v = st[k];
        const int n = (int)(v >> 8);
        if (n >= level_base && n < 2 * level_base) {
            const int a = act[n - level_base];
            if (a >= 0) st[k] = ((uint32_t)a << 8) | (v & 255u);
        }
    }
}

}  // namespace fd

using namespace fd;

extern "C" {

int fd_bin_csr(fd_handle* h, const int32_t* idx_dev, const double* val_dev, int64_t nnz, const int32_t* thr_ptr_dev,
               const double* thr_dev, uint8_t* out_bin_dev, void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (nnz < 0 || (nnz > 0 && (!idx_dev || !val_dev || !thr_ptr_dev || !out_bin_dev))) return set_err(h, FD_ERR_INVALID, "bad fd_bin_csr arguments");
    if (nnz == 0) return FD_OK;
    FD_CUDA(h, cudaSetDevice(h->device));
    long long g = (nnz + 255) / 256;
    if (g > (long long)h->sm_count * 32) g = (long long)h->sm_count * 32;
    k_bin_csr<<<(unsigned)g, 256, 0, (cudaStream_t)stream>>>(idx_dev, val_dev, nnz, thr_ptr_dev, thr_dev, out_bin_dev);
    h->launches += 1;
    FD_CUDA(h, cudaGetLastError());
    return FD_OK;
}

// bits of a feature id; bytes of scratch fd_csc_build needs: sorted keys + packed values in and out + the sort's workspace
static int csc_end_bit(int32_t num_features) {
    int end_bit = 1;
    while (end_bit < 31 && (1LL << end_bit) < (long long)num_features) ++end_bit;
    return end_bit;
}
static size_t csc_scratch_need(int64_t nnz, int end_bit, size_t* tmp_bytes) {
    size_t tb = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tb, (const int32_t*)nullptr, (int32_t*)nullptr, (const unsigned long long*)nullptr,
                                    (unsigned long long*)nullptr, (int)nnz, 0, end_bit, (cudaStream_t)0);
    if (tmp_bytes) *tmp_bytes = tb;
    const size_t N = (size_t)nnz;
    return ((N * 4 + 255) & ~(size_t)255) + 2 * N * 8 + 256 + tb;
}

int64_t fd_csc_scratch_bytes(fd_handle* h, int64_t nnz, int32_t num_features) {
    if (!h) return FD_ERR_INVALID;
    if (nnz < 0 || nnz >= (1LL << 31) || num_features <= 0) return set_err(h, FD_ERR_INVALID, "bad fd_csc_scratch_bytes arguments");
    return (int64_t)csc_scratch_need(nnz, csc_end_bit(num_features), nullptr);
}

int fd_csc_build(fd_handle* h, const int64_t* rowptr_dev, const int32_t* idx_dev, const uint8_t* bin_dev, const uint8_t* label_dev,
                 int64_t n_rows, int64_t nnz, int32_t num_features, uint32_t* out_rowidx_dev, uint8_t* out_bin_dev, void* scratch_dev,
                 int64_t scratch_bytes, void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (n_rows < 0 || nnz < 0 || num_features <= 0 || (nnz > 0 && (!rowptr_dev || !idx_dev || !bin_dev || !label_dev || !out_rowidx_dev || !out_bin_dev)))
        return set_err(h, FD_ERR_INVALID, "bad fd_csc_build arguments");
    if (n_rows >= (1LL << 31) || nnz >= (1LL << 31)) return set_err(h, FD_ERR_LIMIT, "fd_csc_build: at most 2^31 - 1 rows and entries per rank");
    if (nnz == 0) return FD_OK;
    FD_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int end_bit = csc_end_bit(num_features);
    const size_t N = (size_t)nnz;
    size_t tmp_bytes = 0;
    const size_t need = csc_scratch_need(nnz, end_bit, &tmp_bytes);
    if (!scratch_dev || (size_t)scratch_bytes < need)
        return set_err(h, FD_ERR_CAPACITY, "fd_csc_build needs %zu bytes of scratch (fd_csc_scratch_bytes), got %lld", need,
                       (long long)scratch_bytes);
    char* buf = static_cast<char*>(scratch_dev);
    int32_t* keys_out = reinterpret_cast<int32_t*>(buf);
    unsigned long long* vin = reinterpret_cast<unsigned long long*>(buf + ((N * 4 + 255) & ~(size_t)255));
    unsigned long long* vout = vin + N;
    void* tmp = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(vout + N) + 255) & ~(uintptr_t)255);
    long long g = (n_rows + 7) / 8;
    if (g > (long long)h->sm_count * 16) g = (long long)h->sm_count * 16;
    if (g < 1) g = 1;
    k_csc_pack<<<(unsigned)g, 256, 0, st>>>(rowptr_dev, label_dev, bin_dev, n_rows, vin);
    FD_CUDA(h, cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, idx_dev, keys_out, vin, vout, (int)nnz, 0, end_bit, st));
    long long g2 = (nnz + 255) / 256;
    if (g2 > (long long)h->sm_count * 32) g2 = (long long)h->sm_count * 32;
    k_csc_unpack<<<(unsigned)g2, 256, 0, st>>>(vout, nnz, out_rowidx_dev, out_bin_dev);
    h->launches += 2;
    FD_CUDA(h, cudaGetLastError());
    return FD_OK;
}

int fd_doc_freq(fd_handle* h, const int32_t* idx_dev, const double* val_dev, int64_t nnz, int32_t num_features, int64_t* out_doc_freq_dev,
                void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (nnz < 0 || num_features <= 0 || !out_doc_freq_dev || (nnz > 0 && (!idx_dev || !val_dev))) return set_err(h, FD_ERR_INVALID, "bad fd_doc_freq arguments");
    FD_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    FD_CUDA(h, cudaMemsetAsync(out_doc_freq_dev, 0, (size_t)num_features * sizeof(int64_t), st));
    if (nnz == 0) return FD_OK;
    long long g = (nnz + 255) / 256;
    if (g > (long long)h->sm_count * 32) g = (long long)h->sm_count * 32;
    // private count vectors in the handle's workspace when the matrix is large enough to pay for them (at most 256 MiB)
    const int copies = (int)std::min<long long>(std::min<long long>(g, h->sm_count), ((long long)256 << 20) / ((long long)num_features * 4));
    if (copies >= 2 && nnz >= (1LL << 22) && nnz < (1LL << 32)) {
        const size_t bytes = (size_t)copies * num_features * 4;
        const int rc = ensure_ws(h, bytes + 4096);
        if (rc) return rc;
        h->feat_rows = -1;  // (the workspace is also the staging area of an unfinished fd_featurize_begin)
        unsigned int* part = reinterpret_cast<unsigned int*>(h->ws + 4096);  // (past the control block)
        FD_CUDA(h, cudaMemsetAsync(part, 0, bytes, st));
        k_doc_freq_private<<<(unsigned)g, 256, 0, st>>>(idx_dev, val_dev, nnz, part, copies, num_features);
        k_doc_freq_sum<<<(num_features + 255) / 256, 256, 0, st>>>(part, copies, num_features, reinterpret_cast<unsigned long long*>(out_doc_freq_dev));
        h->launches += 2;
        FD_CUDA(h, cudaGetLastError());
        return FD_OK;
    }
    k_doc_freq<<<(unsigned)g, 256, 0, st>>>(idx_dev, val_dev, nnz, reinterpret_cast<unsigned long long*>(out_doc_freq_dev));
    h->launches += 1;
    FD_CUDA(h, cudaGetLastError());
    return FD_OK;
}

int fd_idf_scale(fd_handle* h, const int32_t* idx_dev, double* val_dev, int64_t nnz, const double* idf_dev, int32_t num_features, void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (nnz < 0 || num_features <= 0 || (nnz > 0 && (!idx_dev || !val_dev || !idf_dev))) return set_err(h, FD_ERR_INVALID, "bad fd_idf_scale arguments");
    if (nnz == 0) return FD_OK;
    FD_CUDA(h, cudaSetDevice(h->device));
    long long g = (nnz + 255) / 256;
    if (g > (long long)h->sm_count * 32) g = (long long)h->sm_count * 32;
    k_idf_scale<<<(unsigned)g, 256, 0, (cudaStream_t)stream>>>(idx_dev, val_dev, nnz, idf_dev);
    h->launches += 1;
    FD_CUDA(h, cudaGetLastError());
    return FD_OK;
}

int fd_rf_histogram(fd_handle* h, const int64_t* colptr_dev, const uint32_t* rowidx_dev, const uint8_t* bin_dev,
                    const uint32_t* state_dev, int64_t n_rows, int32_t n_items, const int32_t* item_tree_dev, const int32_t* item_node_dev, const int32_t* item_feat_dev,
                    const int32_t* item_chain_dev, int32_t n_units, const int32_t* unit_item_dev, const int32_t* unit_seg_dev, int32_t seg_len,
                    uint32_t* out_hist_dev, void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (n_items < 0 || n_units < 0 || seg_len <= 0 || !out_hist_dev) return set_err(h, FD_ERR_INVALID, "bad fd_rf_histogram arguments");
    FD_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    FD_CUDA(h, cudaMemsetAsync(out_hist_dev, 0, (size_t)n_items * HB * 2 * sizeof(uint32_t), st));
    if (n_units == 0) return FD_OK;
    const int g = (n_units + WALK_WARPS - 1) / WALK_WARPS;
    k_rf_histogram<<<g, WALK_WARPS * 32, 0, st>>>(colptr_dev, rowidx_dev, bin_dev, state_dev, n_rows, n_units, unit_item_dev,
                                      unit_seg_dev, seg_len, item_tree_dev, item_node_dev, item_feat_dev, item_chain_dev, out_hist_dev);
    h->launches += 1;
    FD_CUDA(h, cudaGetLastError());
    return FD_OK;
}

int fd_rf_bag_poisson(fd_handle* h, uint64_t seed, double rate, int32_t n_trees, int64_t n_rows, const uint8_t* label_dev,
                      uint8_t* out_bagw_dev, uint64_t* out_root_dev, void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (n_trees <= 0 || n_rows < 0 || !(rate > 0.0) || rate > 32.0 || !out_bagw_dev || !out_root_dev || (n_rows > 0 && !label_dev))
        return set_err(h, FD_ERR_INVALID, "bad fd_rf_bag_poisson arguments (0 < rate <= 32)");
    FD_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    FD_CUDA(h, cudaMemsetAsync(out_root_dev, 0, (size_t)n_trees * 2 * sizeof(uint64_t), st));
    if (n_rows == 0) return FD_OK;
    // thr[k] = P(X <= k) scaled to 2^32 (saturating); the draw is the number of thresholds <= u, so it is capped at 16
    // for rates where P(X > 16) is not negligible the tail is folded into 16 -- rate <= 32 is accepted, Spark's default is 1
    BagCdf cdf;
    double p = std::exp(-rate), c = 0.0;
    for (int k = 0; k < BAG_CDF; ++k) {
        c += p;
        const double scaled = c * 4294967296.0;
        cdf.thr[k] = scaled >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)scaled;
        p *= rate / (double)(k + 1);
    }
    const long long total = ((n_rows + 15) / 16) * (long long)n_trees;
    long long g = (total + 255) / 256;
    if (g > (long long)h->sm_count * 16) g = (long long)h->sm_count * 16;
    k_rf_bag_poisson<<<(unsigned)g, 256, 0, st>>>(seed, cdf, n_trees, n_rows, label_dev, out_bagw_dev,
                                                  reinterpret_cast<unsigned long long*>(out_root_dev));
    h->launches += 1;
    FD_CUDA(h, cudaGetLastError());
    return FD_OK;
}

int fd_rf_best_split(fd_handle* h, const uint32_t* hist_dev, int32_t n_nodes, const int32_t* node_item_ptr_dev,
                     const int32_t* item_feat_dev, const double* node_total_dev, const int32_t* n_splits_dev,
                     const uint8_t* bin_of_zero_dev, int32_t min_instances, double min_info_gain, int32_t* out_best_item_dev,
                     int32_t* out_best_split_dev, double* out_best_gain_dev, double* out_left_dev, double* out_right_dev,
                     void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (n_nodes < 0) return set_err(h, FD_ERR_INVALID, "n_nodes < 0");
    if (n_nodes == 0) return FD_OK;
    FD_CUDA(h, cudaSetDevice(h->device));
    k_rf_best_split<<<(n_nodes + 3) / 4, 128, 0, (cudaStream_t)stream>>>(hist_dev, n_nodes, node_item_ptr_dev, item_feat_dev, node_total_dev,
                                                                        n_splits_dev, bin_of_zero_dev, min_instances, min_info_gain,
                                                                        out_best_item_dev, out_best_split_dev, out_best_gain_dev,
                                                                        out_left_dev, out_right_dev);
    h->launches += 1;
    FD_CUDA(h, cudaGetLastError());
    return FD_OK;
}

int fd_rf_advance(fd_handle* h, const int64_t* colptr_dev, const uint32_t* rowidx_dev, const uint8_t* bin_dev,
                  uint32_t* state_dev, int64_t n_rows, int32_t n_trees, int32_t level_base, int32_t n_split_nodes,
                  const int32_t* split_tree_dev, const int32_t* split_node_dev, const int32_t* split_feat_dev,
                  const int32_t* split_bin_dev, const int32_t* split_chain_dev, int32_t n_units, const int32_t* unit_item_dev,
                  const int32_t* unit_seg_dev, int32_t seg_len, const int32_t* action_dev, void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (n_split_nodes < 0 || n_units < 0 || seg_len <= 0 || level_base <= 0) return set_err(h, FD_ERR_INVALID, "bad fd_rf_advance arguments");
    if (n_split_nodes == 0) return FD_OK;
    FD_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    if (n_units > 0) {
        const int g = (n_units + WALK_WARPS - 1) / WALK_WARPS;
        k_rf_advance_explicit<<<g, WALK_WARPS * 32, 0, st>>>(colptr_dev, rowidx_dev, bin_dev, state_dev, n_rows, n_units, unit_item_dev, unit_seg_dev,
                                                 seg_len, split_tree_dev, split_node_dev, split_feat_dev, split_bin_dev, split_chain_dev);
        h->launches += 1;
    }
    if (!action_dev) {  // default moves are implied by the chain lengths (in_node): no pass over state
        FD_CUDA(h, cudaGetLastError());
        return FD_OK;
    }
    long long g2 = (n_rows + 255) / 256;
    const long long cap = std::max<long long>(1, (long long)h->sm_count * 16 / std::max(1, n_trees));
    if (g2 > cap) g2 = cap;
    if (g2 < 1) g2 = 1;
    if (n_trees > 65535) return set_err(h, FD_ERR_LIMIT, "at most 65535 trees per advance call");
    k_rf_advance_default<<<dim3((unsigned)g2, (unsigned)n_trees), 256, 0, st>>>(state_dev, n_rows, level_base, action_dev);
    h->launches += 1;
    FD_CUDA(h, cudaGetLastError());
    return FD_OK;
}

}  // extern "C"
