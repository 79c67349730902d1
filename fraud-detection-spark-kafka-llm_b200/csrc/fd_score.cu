// fd_score.cu -- classifier stage on CSR features (the path taken when the caller already holds the
// `features` column): LogisticRegression / DecisionTree / RandomForest / GBT -> prediction,
// probability, rawPrediction.  One warp per dialogue; sparse feature reads are binary searches over the
// dialogue's ascending bucket ids (SparseVector.apply), absent -> 0.0.
//
// Spark semantics restated: ProbabilisticClassificationModel.transform; reference call
// sites utils/agent_api.py:158-167 and fraud_detection_spark.py:98-118.
#include "fd_rows.cuh"

namespace fd {

__device__ __forceinline__ double csr_value(const int32_t* __restrict__ idx, const double* __restrict__ val, int n, int f) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(idx + mid) < f) lo = mid + 1;
        else hi = mid;
    }
    return (lo < n && __ldg(idx + lo) == f) ? __ldg(val + lo) : 0.0;
}

// Forests read ~1,000 features per dialogue, nearly all of them absent from it, and a binary search pays its full
// depth for every one.  A warp therefore first builds a loop-free front over its dialogue in shared memory:
// front[g(f)] = 1 + position of THE feature of the row that maps to g, 0 when none does (absent: answered by one
// 16-bit load), FRONT_MANY when several do (binary search as before).
constexpr int FRONT_LOG2 = 10;
constexpr uint32_t FRONT_MANY = 0xFFFFu;
__device__ __forceinline__ uint32_t front_slot(uint32_t f) { return (f * 0x85EBCA6Bu) >> (32 - FRONT_LOG2); }

__device__ __forceinline__ void build_front(uint16_t* front, const int32_t* __restrict__ idx, int n, int lane) {
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (int i = lane; i < (2 << FRONT_LOG2) / 16; i += 32) reinterpret_cast<uint4*>(front)[i] = z;
    __syncwarp();
    const int m = min(n, (int)FRONT_MANY - 1);  // positions beyond 65,533 cannot be encoded: those features ...
    for (int i = lane; i < m; i += 32) front[front_slot((uint32_t)__ldg(idx + i))] = (uint16_t)(i + 1);
    __syncwarp();
    for (int i = lane; i < n; i += 32) {  // ... and every feature that lost its slot mark it shared
        const uint32_t g = front_slot((uint32_t)__ldg(idx + i));
        if (i >= m || front[g] != (uint16_t)(i + 1)) front[g] = (uint16_t)FRONT_MANY;
    }
    __syncwarp();
}

__device__ __forceinline__ double csr_value(const uint16_t* front, const int32_t* __restrict__ idx, const double* __restrict__ val, int n, int f) {
    if (front == nullptr) return csr_value(idx, val, n, f);
    const uint32_t j = front[front_slot((uint32_t)f)];
    if (j == 0u) return 0.0;
    if (j == FRONT_MANY) return csr_value(idx, val, n, f);
    return __ldg(idx + (j - 1u)) == f ? __ldg(val + (j - 1u)) : 0.0;
}

__device__ __forceinline__ int walk_csr(const Model& m, const uint16_t* front, const int32_t* idx, const double* val, int n, int root,
                                        double& leaf_pred) {
    int nd = root;
    for (;;) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(m.nodes + nd));
        const int feat = (int)q.z;
        const double thr = __hiloint2double((int)q.y, (int)q.x);
        if (feat < 0) {
            leaf_pred = thr;
            return nd;
        }
        nd = (int)q.w + ((csr_value(front, idx, val, n, feat) <= thr) ? 0 : 1);
    }
}

__global__ void __launch_bounds__(256) k_score_csr(const int64_t* __restrict__ rowptr, const int32_t* __restrict__ idx_all,
                                                   const double* __restrict__ val_all, long long n_rows, const Model m,
                                                   double* __restrict__ pred, double* __restrict__ prob,
                                                   double* __restrict__ raw) {
    __shared__ double scratch_all[8][64];
    __shared__ __align__(16) uint16_t front_all[8][1 << FRONT_LOG2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double* scratch = scratch_all[warp];
    const bool use_front = m.classifier == 2 && m.n_trees >= 8;
    const long long nwarps = (long long)gridDim.x * 8;
    for (long long row = (long long)blockIdx.x * 8 + warp; row < n_rows; row += nwarps) {
        const long long o = rowptr[row];
        const int n = (int)(rowptr[row + 1] - o);
        const int32_t* idx = idx_all + o;
        const double* val = val_all + o;
        double out_pred, p0, p1, r0, r1;
        const uint16_t* front = nullptr;
        if (use_front) {
            build_front(front_all[warp], idx, n, lane);
            front = front_all[warp];
        }
        if (m.classifier == 1) {
            double sum = 0.0;
            for (int b = 0; b < n; b += 32) {
                const int i = b + lane;
                if (i < n) scratch[lane] = __dmul_rn(val[i], __ldg(m.coef + idx[i]));
                __syncwarp();
                if (lane == 0) {
                    const int cnt = min(32, n - b);
                    for (int l = 0; l < cnt; ++l) sum = __dadd_rn(sum, scratch[l]);
                }
                __syncwarp();
            }
            sum = __shfl_sync(0xFFFFFFFFu, sum, 0);
            const double mg = __dadd_rn(sum, m.intercept);
            p0 = 1.0 / (1.0 + exp(mg));
            p1 = 1.0 - p0;
            r0 = -mg;
            r1 = mg;
            out_pred = mg > m.raw_threshold ? 1.0 : 0.0;
        } else if (m.kind == FD_KIND_DT) {
            double lp;
            const int nd = walk_csr(m, nullptr, idx, val, n, __ldg(m.roots), lp);
            const double2 st = __ldg(m.leaf2 + nd), pr = __ldg(m.leaf_prob + nd);
            r0 = st.x;
            r1 = st.y;
            p0 = pr.x;
            p1 = pr.y;
            out_pred = st.y > st.x ? 1.0 : 0.0;
        } else if (m.kind == FD_KIND_RF) {
            double v = 0.0;
            for (int b = 0; b < m.n_trees; b += 32) {
                const int t = b + lane;
                if (t < m.n_trees) {
                    double lp;
                    const int nd = walk_csr(m, front, idx, val, n, __ldg(m.roots + t), lp);
                    const double2 st = __ldg(m.leaf2 + nd);
                    scratch[lane] = st.x;
                    scratch[32 + lane] = st.y;
                }
                __syncwarp();
                if (lane < 2) {
                    const int cnt = min(32, m.n_trees - b);
                    const double* src = scratch + 32 * lane;
                    for (int l = 0; l < cnt; ++l) v = __dadd_rn(v, src[l]);
                }
                __syncwarp();
            }
            r0 = __shfl_sync(0xFFFFFFFFu, v, 0);
            r1 = __shfl_sync(0xFFFFFFFFu, v, 1);
            const double tot = __dadd_rn(r0, r1);
            p0 = tot != 0.0 ? r0 / tot : r0;
            p1 = tot != 0.0 ? r1 / tot : r1;
            out_pred = r1 > r0 ? 1.0 : 0.0;
        } else {  // GBT
            double mg = 0.0;
            for (int b = 0; b < m.n_trees; b += 32) {
                const int t = b + lane;
                if (t < m.n_trees) {
                    double lp;
                    walk_csr(m, front, idx, val, n, __ldg(m.roots + t), lp);
                    scratch[lane] = __dmul_rn(lp, __ldg(m.tree_w + t));
                }
                __syncwarp();
                if (lane == 0) {
                    const int cnt = min(32, m.n_trees - b);
                    for (int l = 0; l < cnt; ++l) mg = __dadd_rn(mg, scratch[l]);
                }
                __syncwarp();
            }
            mg = __shfl_sync(0xFFFFFFFFu, mg, 0);
            r0 = -mg;
            r1 = mg;
            p0 = 1.0 / (1.0 + exp(-2.0 * r0));
            p1 = 1.0 - p0;
            out_pred = r1 > r0 ? 1.0 : 0.0;
        }
        if (lane == 0) pred[row] = out_pred;
        else if (lane == 1) prob[2 * row] = p0;
        else if (lane == 2) prob[2 * row + 1] = p1;
        else if (lane == 3) raw[2 * row] = r0;
        else if (lane == 4) raw[2 * row + 1] = r1;
    }
}

cudaError_t launch_score_csr(fd_handle* h, const int64_t* rowptr, const int32_t* idx, const double* val, long long n_rows,
                             double* pred, double* prob, double* raw, cudaStream_t st) {
    long long g = (n_rows + 7) / 8;
    if (g > (long long)h->sm_count * 8) g = (long long)h->sm_count * 8;
    if (g < 1) g = 1;
    k_score_csr<<<(unsigned)g, 256, 0, st>>>(rowptr, idx, val, n_rows, h->m, pred, prob, raw);
    h->launches += 1;
    return cudaGetLastError();
}

}  // namespace fd
