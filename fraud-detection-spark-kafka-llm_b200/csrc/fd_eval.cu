// fd_eval.cu -- the evaluators the reference applies to the transform output (fraud_detection_spark.py:101-119):
// BinaryClassificationEvaluator(areaUnderROC) on rawPrediction[1], and the confusion counts behind
// MulticlassClassificationEvaluator (accuracy / weightedPrecision / weightedRecall / f1) and crosstab(labels, prediction).
//
// Spark semantics restated: mllib BinaryClassificationMetrics on one partition -- counts per DISTINCT
// score in descending score order (java.lang.Double equality and order, i.e. by bit pattern: -0.0 and 0.0 differ),
// down-sampled in chunks of (distinct / numBins) consecutive scores when that quotient is >= 2, ROC points (0,0) ...
// (1,1), trapezoid rule.  The sort is a library radix sort (CUB, as cuBLAS would be for a GEMM); everything after it --
// group heads, cumulative positives, chunk points, the trapezoid sum -- is this file's kernels.
#include <cub/cub.cuh>

#include "fd_common.cuh"

namespace fd {

// score -> key whose unsigned order is the score's order; value = 1 for a positive label
__global__ void k_auc_keys(const double* __restrict__ score, long long stride_elems, const double* __restrict__ label, long long n,
                           unsigned long long* __restrict__ key, unsigned int* __restrict__ pos) {
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += step) {
        const unsigned long long b = (unsigned long long)__double_as_longlong(score[i * stride_elems]);
        key[i] = (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
        pos[i] = label[i] > 0.5 ? 1u : 0u;
    }
}

// head[i] = 1 where a new distinct score starts (input sorted descending)
__global__ void k_auc_heads(const unsigned long long* __restrict__ key, long long n, unsigned int* __restrict__ head) {
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += step) head[i] = (i == 0 || key[i] != key[i - 1]) ? 1u : 0u;
}

// gpos[d] = first position of the d-th distinct score (gidx = inclusive scan of head)
__global__ void k_auc_group_starts(const unsigned int* __restrict__ head, const unsigned int* __restrict__ gidx, long long n,
                                   unsigned int* __restrict__ gpos) {
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += step)
        if (head[i]) gpos[gidx[i] - 1u] = (unsigned int)i;
}

// one thread per chunk of `grouping` consecutive distinct scores: the trapezoid between the ROC point before the chunk and
// the point after it.  cpos = inclusive scan of the positive flags.  out[0] += area, out[1] = positives, out[2] = negatives
__global__ void k_auc_area(const unsigned int* __restrict__ gpos, const unsigned int* __restrict__ cpos, const unsigned int* __restrict__ gidx,
                           long long n, int num_bins, double* __restrict__ out) {
    const long long nd = gidx[n - 1];
    const long long grouping = (num_bins > 0 && nd / num_bins >= 2) ? nd / num_bins : 1;
    const long long chunks = (nd + grouping - 1) / grouping;
    const double tp = (double)cpos[n - 1], tn = (double)(n - (long long)cpos[n - 1]);
    double acc = 0.0;
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x; c < chunks; c += step) {
        const long long s = gpos[c * grouping];                                         // first position of the chunk
        const long long e = (c + 1) * grouping < nd ? (long long)gpos[(c + 1) * grouping] - 1 : n - 1;  // its last position
        const double cp0 = s > 0 ? (double)cpos[s - 1] : 0.0, cn0 = (double)s - cp0;
        const double cp1 = (double)cpos[e], cn1 = (double)(e + 1) - cp1;
        const double x0 = tn > 0.0 ? cn0 / tn : 0.0, x1 = tn > 0.0 ? cn1 / tn : 0.0;
        const double y0 = tp > 0.0 ? cp0 / tp : 0.0, y1 = tp > 0.0 ? cp1 / tp : 0.0;
        acc += (x1 - x0) * (y1 + y0) / 2.0;
    }
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_down_sync(0xFFFFFFFFu, acc, d);
    if ((threadIdx.x & 31) == 0 && acc != 0.0) atomicAdd(out, acc);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        out[1] = tp;
        out[2] = tn;
    }
}

// counts[2 y + p] over binary labels / predictions; counts[4] = entries with a value outside {0, 1}
__global__ void k_confusion2(const double* __restrict__ label, const double* __restrict__ pred, long long n, unsigned long long* __restrict__ counts) {
    unsigned int c[5] = {0u, 0u, 0u, 0u, 0u};
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += step) {
        const double y = label[i], p = pred[i];
        const bool ok = (y == 0.0 || y == 1.0) && (p == 0.0 || p == 1.0);
        if (ok) c[2 * (int)y + (int)p] += 1u;
        else c[4] += 1u;
    }
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        unsigned int v = c[k];
        for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, d);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(counts + k, (unsigned long long)v);
    }
}

}  // namespace fd

using namespace fd;

extern "C" {

int fd_eval_auc_roc(fd_handle* h, const double* score_dev, int64_t score_stride, const double* label_dev, int64_t n, int32_t num_bins,
                    double* out_dev, void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (n < 0 || score_stride <= 0 || !out_dev || (n > 0 && (!score_dev || !label_dev))) return set_err(h, FD_ERR_INVALID, "bad fd_eval_auc_roc arguments");
    if (n >= (1LL << 32)) return set_err(h, FD_ERR_LIMIT, "at most 2^32 - 1 rows per evaluation");
    FD_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    FD_CUDA(h, cudaMemsetAsync(out_dev, 0, 3 * sizeof(double), st));
    if (n == 0) return FD_OK;
    // scratch (an evaluation is far off the hot path: allocated and freed per call)
    const size_t N = (size_t)n;
    size_t tmp_sort = 0, tmp_scan = 0;
    cub::DeviceRadixSort::SortPairsDescending(nullptr, tmp_sort, (unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                              (unsigned int*)nullptr, (unsigned int*)nullptr, (int)n, 0, 64, st);
    cub::DeviceScan::InclusiveSum(nullptr, tmp_scan, (unsigned int*)nullptr, (unsigned int*)nullptr, (int)n, st);
    const size_t tmp_bytes = std::max(tmp_sort, tmp_scan);
    const size_t total = 2 * N * 8 + 5 * N * 4 + tmp_bytes + 1024;
    char* buf = nullptr;
    FD_CUDA(h, cudaMalloc((void**)&buf, total));
    unsigned long long* key_in = reinterpret_cast<unsigned long long*>(buf);
    unsigned long long* key = key_in + N;
    unsigned int* pos_in = reinterpret_cast<unsigned int*>(key + N);
    unsigned int* pos = pos_in + N;      // positive flags in sorted order, then their inclusive scan
    unsigned int* head = pos + N;        // group heads, then reused for the starts of the distinct groups
    unsigned int* gidx = head + N;       // inclusive scan of the heads
    unsigned int* gpos = gidx + N;
    void* tmp = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(gpos + N) + 255) & ~(uintptr_t)255);
    long long g = (n + 255) / 256;
    if (g > (long long)h->sm_count * 16) g = (long long)h->sm_count * 16;
    k_auc_keys<<<(unsigned)g, 256, 0, st>>>(score_dev, (long long)score_stride, label_dev, n, key_in, pos_in);
    size_t tb = tmp_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortPairsDescending(tmp, tb, key_in, key, pos_in, pos, (int)n, 0, 64, st);
    if (e == cudaSuccess) {
        k_auc_heads<<<(unsigned)g, 256, 0, st>>>(key, n, head);
        tb = tmp_bytes;
        e = cub::DeviceScan::InclusiveSum(tmp, tb, head, gidx, (int)n, st);
    }
    if (e == cudaSuccess) {
        k_auc_group_starts<<<(unsigned)g, 256, 0, st>>>(head, gidx, n, gpos);
        tb = tmp_bytes;
        e = cub::DeviceScan::InclusiveSum(tmp, tb, pos, pos_in, (int)n, st);  // cumulative positives into pos_in
    }
    if (e == cudaSuccess) {
        k_auc_area<<<(unsigned)g, 256, 0, st>>>(gpos, pos_in, gidx, n, num_bins, out_dev);
        h->launches += 4;
        e = cudaGetLastError();
    }
    cudaError_t e2 = cudaStreamSynchronize(st);  // (the scratch is freed below)
    cudaFree(buf);
    if (e != cudaSuccess) return cuda_fail(h, e, "fd_eval_auc_roc");
    if (e2 != cudaSuccess) return cuda_fail(h, e2, "fd_eval_auc_roc (synchronize)");
    return FD_OK;
}

int fd_eval_confusion2(fd_handle* h, const double* label_dev, const double* pred_dev, int64_t n, int64_t* out_counts_dev, void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (n < 0 || !out_counts_dev || (n > 0 && (!label_dev || !pred_dev))) return set_err(h, FD_ERR_INVALID, "bad fd_eval_confusion2 arguments");
    FD_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    FD_CUDA(h, cudaMemsetAsync(out_counts_dev, 0, 5 * sizeof(int64_t), st));
    if (n == 0) return FD_OK;
    long long g = (n + 255) / 256;
    if (g > (long long)h->sm_count * 16) g = (long long)h->sm_count * 16;
    k_confusion2<<<(unsigned)g, 256, 0, st>>>(label_dev, pred_dev, n, reinterpret_cast<unsigned long long*>(out_counts_dev));
    h->launches += 1;
    FD_CUDA(h, cudaGetLastError());
    return FD_OK;
}

}  // extern "C"
