// fd_noalloc.cu -- fd_transform_noalloc: the fused transform for callers that must not allocate.
//
// Who needs it: a process whose device memory is already spoken for -- an inference server or a Spark/RAPIDS
// executor that handed the whole GPU to its own pool allocator at start-up -- and latency-bound micro-batch
// consumers (the Kafka loop of the reference's app) that can afford neither the driver's allocation lock nor
// an out-of-memory failure in the middle of a stream.  fd_transform grows a handle-owned workspace with
// cudaMalloc on first use; this variant instead keeps ALL of its scratch in one fixed array that is part of
// the library image, so the call itself performs no cudaMalloc / cudaFree and calls no allocating routine.
//
// The price is a hard limit and shared state:
//   * at most FD_NOALLOC_MAX_ROWS dialogues per call (the deferred-dialogue list has one slot per dialogue,
//     so it cannot overflow); larger batches are rejected with FD_ERR_LIMIT before anything is launched;
//   * every GPU that loads the library carries the array (about 4.5 MiB) until the process exits;
//   * all calls on a device share the array and it keeps whatever the previous call left in it.  The call
//     clears the control block (counters and status must start at zero) on the stream before launching; the
//     wide tables are re-initialised by the kernel that uses them.  Two calls whose GPU work overlaps (different
//     streams, or different handles on one device) race on it -- serialise them.
//   * the first call on a device looks the array's device address up (cudaGetSymbolAddress, once per device,
//     with that device current).  Make that first call before starting any stream capture.
#include <cstring>
#include <mutex>

#include "fd_rows.cuh"

namespace {

constexpr long long NOALLOC_MAX_ROWS = 1 << 16;  // FD_NOALLOC_MAX_ROWS in fdb200.h
constexpr int NOALLOC_WORKERS = 8;               // large wide-table workers (512 KiB table each)
constexpr int NOALLOC_MED_CTAS = 16;             // medium wide-table CTAs (8 warps x 32 KiB each); shares the same 4 MiB
constexpr size_t NA_CTRL = 256;
constexpr size_t NA_TABS = fd::wide_pool_slots(NOALLOC_MED_CTAS, NOALLOC_WORKERS) * 8;
constexpr size_t NA_LIST = (size_t)NOALLOC_MAX_ROWS * 8;
constexpr size_t NA_BYTES = NA_CTRL + NA_TABS + NA_LIST;

// The scratch itself.  In HOST code this name is only a stand-in for the symbol, not a device address: it is
// never passed to a kernel; the address below comes from cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
void* g_scratch_addr[MAX_DEVICES] = {nullptr};

// once per device, with that device current
int scratch_address(fd_handle* h, char** out) {
    if (h->device < 0 || h->device >= MAX_DEVICES) return fd::set_err(h, FD_ERR_INVALID, "device index %d out of range", h->device);
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[h->device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return fd::cuda_fail(h, e, "cudaGetSymbolAddress(g_noalloc_scratch)");
        if (!p) return fd::set_err(h, FD_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        g_scratch_addr[h->device] = p;
    }
    *out = static_cast<char*>(g_scratch_addr[h->device]);
    return FD_OK;
}

}  // namespace

using namespace fd;

extern "C" int fd_transform_noalloc(fd_handle* h, const uint8_t* text_dev, const int64_t* offs_dev, int64_t n_rows,
                                    uint32_t flags, double* out_pred_dev, double* out_prob_dev, double* out_raw_dev,
                                    int32_t* out_status_dev, void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (n_rows < 0) return set_err(h, FD_ERR_INVALID, "n_rows < 0");
    if (n_rows > NOALLOC_MAX_ROWS)
        return set_err(h, FD_ERR_LIMIT, "fd_transform_noalloc takes at most %lld dialogues per call (got %lld): its scratch is a fixed "
                       "array; split the batch or use fd_transform", NOALLOC_MAX_ROWS, (long long)n_rows);
    if (n_rows > 0 && (!text_dev || !offs_dev || !out_pred_dev || !out_prob_dev || !out_raw_dev))
        return set_err(h, FD_ERR_INVALID, "null input or output pointer");
    if (((uintptr_t)text_dev & 15) != 0) return set_err(h, FD_ERR_INVALID, "text pointer must be 16-byte aligned");
    if (h->m.num_features <= 0) return set_err(h, FD_ERR_STATE, "HashingTF stage not set (fd_set_hashing)");
    int mode;
    if (h->m.classifier == 1) mode = MODE_LINEAR;
    else if (h->m.classifier == 2) mode = h->m.kind == FD_KIND_DT ? MODE_DT : (h->m.kind == FD_KIND_RF ? MODE_RF : MODE_GBT);
    else return set_err(h, FD_ERR_STATE, "no classifier stage set (fd_set_linear / fd_set_forest)");
    FD_CUDA(h, cudaSetDevice(h->device));
    if (h->fused_dirty)  // cannot happen: fd_set_idf / fd_set_forest prepare the model (and allocate) at upload time
        return set_err(h, FD_ERR_STATE, "model not prepared");
    char* base = nullptr;
    int rc = scratch_address(h, &base);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    RowArgs a;
    memset(&a, 0, sizeof(a));
    a.host_bytes = -1;
    a.text = text_dev;
    a.offs = offs_dev;
    a.n_rows = n_rows;
    a.flags = flags;
    a.m = h->m;  // (after prepare_fused)
    a.ctrl = reinterpret_cast<Ctrl*>(base);
    a.big_tabs = reinterpret_cast<unsigned long long*>(base + NA_CTRL);
    a.big2_tabs = a.big_tabs;  // one pool: the two wide-table kernels run one after the other here
    a.spill_tabs = nullptr;  // (a fixed scratch array cannot hold a table per warp: overflowing dialogues are deferred)
    a.big_rows = reinterpret_cast<long long*>(base + NA_CTRL + NA_TABS);
    a.big_cap = NOALLOC_MAX_ROWS;
    a.pred = out_pred_dev;
    a.prob = out_prob_dev;
    a.raw = out_raw_dev;
    // the array still holds the previous call's counters and status: they must start at zero
    FD_CUDA(h, cudaMemsetAsync(a.ctrl, 0, sizeof(Ctrl), st));
    if (n_rows > 0) FD_CUDA(h, launch_rows_workers(h, mode, a, NOALLOC_MED_CTAS, NOALLOC_WORKERS, false, st));
    if (out_status_dev)
        FD_CUDA(h, cudaMemcpyAsync(out_status_dev, &a.ctrl->status, sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    return FD_OK;
}
