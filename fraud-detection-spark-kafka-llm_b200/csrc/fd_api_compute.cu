// fd_api_compute.cu -- extern "C" compute entry points: featurize (two passes), score, fused transform
// on device buffers, and the host-buffer transform that pipelines H2D / kernels / D2H.
#include <algorithm>
#include <cstring>

#include "fd_rows.cuh"

namespace fd {

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

constexpr size_t SLOT_CTRL = 256;
constexpr size_t SLOT_MEDTABS = med_pool_slots(MED_CTAS) * 8;
constexpr size_t SLOT_BIGTABS = SLOT_MEDTABS + big_pool_slots(BIG_WORKERS) * 8;  // two disjoint pools (the variants run side by side)
// scratch slot = control block, wide tables, per-warp spill tables, then the deferred-dialogue list with ONE ENTRY PER DIALOGUE of the
// call (every dialogue may be deferred, so the list can never overflow)
constexpr size_t SLOT_SPILL = SPILL_SLOTS * 8;
static size_t slot_bytes(long long n_rows) { return align_up(SLOT_CTRL + SLOT_BIGTABS + SLOT_SPILL + (size_t)n_rows * 8, 256); }

static void fill_slot(RowArgs& a, char* base, long long n_rows) {
    a.ctrl = reinterpret_cast<Ctrl*>(base);
    a.big_tabs = reinterpret_cast<unsigned long long*>(base + SLOT_CTRL);
    a.big2_tabs = reinterpret_cast<unsigned long long*>(base + SLOT_CTRL + SLOT_MEDTABS);
    a.spill_tabs = reinterpret_cast<unsigned long long*>(base + SLOT_CTRL + SLOT_BIGTABS);
    a.big_rows = reinterpret_cast<long long*>(base + SLOT_CTRL + SLOT_BIGTABS + SLOT_SPILL);
    a.big_cap = n_rows;
}

static int classifier_mode(fd_handle* h, int* mode) {
    if (h->m.classifier == 1) *mode = MODE_LINEAR;
    else if (h->m.classifier == 2)
        *mode = h->m.kind == FD_KIND_DT ? MODE_DT : (h->m.kind == FD_KIND_RF ? MODE_RF : MODE_GBT);
    else return set_err(h, FD_ERR_STATE, "no classifier stage set (fd_set_linear / fd_set_forest)");
    return FD_OK;
}

static int check_text_args(fd_handle* h, const void* text, const void* offs, int64_t n_rows) {
    if (!h) return FD_ERR_INVALID;
    if (n_rows < 0) return set_err(h, FD_ERR_INVALID, "n_rows < 0");
    if (n_rows > 0 && (!text || !offs)) return set_err(h, FD_ERR_INVALID, "null text/offs pointer");
    if (((uintptr_t)text & 15) != 0) return set_err(h, FD_ERR_INVALID, "text pointer must be 16-byte aligned");
    if (h->m.num_features <= 0) return set_err(h, FD_ERR_STATE, "HashingTF stage not set (fd_set_hashing)");
    return FD_OK;
}

// read the control block back (synchronises the stream)
static int read_ctrl(fd_handle* h, const Ctrl* dev, Ctrl* out, cudaStream_t st) {
    if (!h->pinned_ctrl) FD_CUDA(h, cudaMallocHost((void**)&h->pinned_ctrl, 4 * sizeof(Ctrl)));
    FD_CUDA(h, cudaMemcpyAsync(h->pinned_ctrl, dev, sizeof(Ctrl), cudaMemcpyDeviceToHost, st));
    FD_CUDA(h, cudaStreamSynchronize(st));
    *out = h->pinned_ctrl[0];
    return FD_OK;
}

// featurize workspace: [slot fixed][row_nnz i32][row_tmpoff i64][bsum i64][tmp u64]
struct FeatLayout {
    size_t off_nnz, off_tmpoff, off_bsum, off_tmp, total;
};
static FeatLayout feat_layout(int64_t n_rows, int64_t tmp_cap) {
    FeatLayout L;
    const size_t nb = (size_t)((n_rows + 1023) / 1024);
    L.off_nnz = slot_bytes(n_rows);
    L.off_tmpoff = align_up(L.off_nnz + (size_t)n_rows * 4, 256);
    L.off_bsum = align_up(L.off_tmpoff + (size_t)n_rows * 8, 256);
    L.off_tmp = align_up(L.off_bsum + (nb + 1) * 8, 256);
    L.total = L.off_tmp + (size_t)tmp_cap * 8;
    return L;
}
// staging entries fd_featurize_begin starts with (it re-runs with the exact size when a batch needs more)
static int64_t feat_guess(int64_t n_rows) { return std::max<int64_t>(1 << 20, n_rows * 96); }

static int run_featurize_pass(fd_handle* h, const uint8_t* text, const int64_t* offs, int64_t n_rows, uint32_t flags,
                              int64_t tmp_cap, cudaStream_t st, Ctrl* result) {
    const FeatLayout L = feat_layout(n_rows, tmp_cap);
    const size_t off_nnz = L.off_nnz, off_tmpoff = L.off_tmpoff, off_bsum = L.off_bsum, off_tmp = L.off_tmp, total = L.total;
    int rc = ensure_ws(h, total);
    if (rc) return rc;
    h->off_nnz = off_nnz;
    h->off_tmpoff = off_tmpoff;
    h->off_bsum = off_bsum;
    h->off_tmp = off_tmp;
    h->tmp_cap = tmp_cap;
    RowArgs a;
    memset(&a, 0, sizeof(a));
    a.host_bytes = -1;
    a.text = text;
    a.offs = offs;
    a.n_rows = n_rows;
    a.flags = flags;
    a.m = h->m;  // (after prepare_fused)
    fill_slot(a, h->ws, n_rows);
    a.row_nnz = reinterpret_cast<int32_t*>(h->ws + off_nnz);
    a.row_tmpoff = reinterpret_cast<long long*>(h->ws + off_tmpoff);
    a.tmp = reinterpret_cast<unsigned long long*>(h->ws + off_tmp);
    a.tmp_cap = tmp_cap;
    FD_CUDA(h, cudaMemsetAsync(a.ctrl, 0, CTRL_CLEAR_BYTES, st));  // cursor, big_count, overflow (status is sticky)
    FD_CUDA(h, launch_rows(h, MODE_CSR, a, st));
    return read_ctrl(h, a.ctrl, result, st);
}

}  // namespace fd

using namespace fd;

extern "C" {

int fd_featurize_begin(fd_handle* h, const uint8_t* text_dev, const int64_t* offs_dev, int64_t n_rows, uint32_t flags,
                       int64_t* nnz_host, void* stream) {
    int rc = check_text_args(h, text_dev, offs_dev, n_rows);
    if (rc) return rc;
    if (!nnz_host) return set_err(h, FD_ERR_INVALID, "nnz_host is null");
    FD_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    h->feat_rows = -1;
    *nnz_host = 0;
    if (n_rows == 0) {
        h->feat_rows = 0;
        h->feat_nnz = 0;
        return FD_OK;
    }
    // staging capacity: total bytes are not known on the host, start from a per-dialogue guess and
    // re-run with the exact size if the kernel reports an overflow
    int64_t cap = std::max<int64_t>(h->tmp_cap, feat_guess(n_rows));
    Ctrl c;
    rc = run_featurize_pass(h, text_dev, offs_dev, n_rows, flags, cap, st, &c);
    if (rc) return rc;
    if (c.tmp_overflow) {
        cap = (int64_t)c.tmp_cursor + 1024;
        rc = run_featurize_pass(h, text_dev, offs_dev, n_rows, flags, cap, st, &c);
        if (rc) return rc;
        if (c.tmp_overflow) return set_err(h, FD_ERR_CUDA, "CSR staging overflow after resize");
    }
    if (c.status != FD_OK) {
        cudaMemsetAsync(&reinterpret_cast<Ctrl*>(h->ws)->status, 0, sizeof(int), st);
        return set_err(h, c.status, "a dialogue exceeds the supported number of distinct hashed terms (%d)",
                       (1 << BIG_LOG2) * 3 / 4);
    }
    h->feat_rows = n_rows;
    h->feat_unsorted = (int64_t)c.big_count + (int64_t)c.big2_count;  // dialogues the wide-table kernel staged (not ordered)
    h->feat_nnz = (int64_t)c.tmp_cursor;
    *nnz_host = h->feat_nnz;
    // exclusive scan of per-dialogue nnz (block sums now; row pointers in finish)
    FD_CUDA(h, launch_csr_assembly(h, reinterpret_cast<int32_t*>(h->ws + h->off_nnz), nullptr, nullptr, n_rows,
                                   reinterpret_cast<long long*>(h->ws + h->off_bsum), nullptr, nullptr, nullptr, true, 0, st));
    return FD_OK;
}

int fd_featurize_finish(fd_handle* h, int64_t* out_rowptr_dev, int32_t* out_idx_dev, double* out_val_dev, int64_t capacity,
                        void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (h->feat_rows < 0) return set_err(h, FD_ERR_STATE, "fd_featurize_finish without a successful fd_featurize_begin");
    if (!out_rowptr_dev) return set_err(h, FD_ERR_INVALID, "null rowptr");
    if (capacity < h->feat_nnz) return set_err(h, FD_ERR_CAPACITY, "capacity %lld < nnz %lld", (long long)capacity, (long long)h->feat_nnz);
    if (h->feat_nnz > 0 && (!out_idx_dev || !out_val_dev)) return set_err(h, FD_ERR_INVALID, "null idx/val");
    FD_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    if (h->feat_rows == 0) {
        FD_CUDA(h, cudaMemsetAsync(out_rowptr_dev, 0, 8, st));
        return FD_OK;
    }
    FD_CUDA(h, launch_csr_assembly(h, reinterpret_cast<int32_t*>(h->ws + h->off_nnz),
                                   reinterpret_cast<long long*>(h->ws + h->off_tmpoff),
                                   reinterpret_cast<unsigned long long*>(h->ws + h->off_tmp), h->feat_rows,
                                   reinterpret_cast<long long*>(h->ws + h->off_bsum), (long long*)out_rowptr_dev, out_idx_dev,
                                   out_val_dev, false, h->feat_unsorted, st));
    return FD_OK;
}

int fd_featurize(fd_handle* h, const uint8_t* text_dev, const int64_t* offs_dev, int64_t n_rows, uint32_t flags,
                 int64_t* out_rowptr_dev, int32_t* out_idx_dev, double* out_val_dev, int64_t capacity, int64_t* nnz_host,
                 void* stream) {
    int64_t nnz = 0;
    int rc = fd_featurize_begin(h, text_dev, offs_dev, n_rows, flags, &nnz, stream);
    if (nnz_host) *nnz_host = nnz;
    if (rc) return rc;
    return fd_featurize_finish(h, out_rowptr_dev, out_idx_dev, out_val_dev, capacity, stream);
}

int fd_score(fd_handle* h, const int64_t* rowptr_dev, const int32_t* idx_dev, const double* val_dev, int64_t n_rows,
             double* out_pred_dev, double* out_prob_dev, double* out_raw_dev, void* stream) {
    if (!h) return FD_ERR_INVALID;
    int mode;
    int rc = classifier_mode(h, &mode);
    if (rc) return rc;
    if (n_rows < 0 || !rowptr_dev || !out_pred_dev || !out_prob_dev || !out_raw_dev)
        return set_err(h, FD_ERR_INVALID, "bad fd_score arguments");
    if (n_rows == 0) return FD_OK;
    FD_CUDA(h, cudaSetDevice(h->device));
    FD_CUDA(h, launch_score_csr(h, rowptr_dev, idx_dev, val_dev, n_rows, out_pred_dev, out_prob_dev, out_raw_dev,
                                (cudaStream_t)stream));
    return FD_OK;
}

int fd_transform(fd_handle* h, const uint8_t* text_dev, const int64_t* offs_dev, int64_t n_rows, uint32_t flags,
                 double* out_pred_dev, double* out_prob_dev, double* out_raw_dev, void* stream) {
    int rc = check_text_args(h, text_dev, offs_dev, n_rows);
    if (rc) return rc;
    int mode;
    rc = classifier_mode(h, &mode);
    if (rc) return rc;
    if (n_rows == 0) return FD_OK;
    if (!out_pred_dev || !out_prob_dev || !out_raw_dev) return set_err(h, FD_ERR_INVALID, "null output column");
    FD_CUDA(h, cudaSetDevice(h->device));
    rc = prepare_fused(h);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    rc = ensure_ws(h, slot_bytes(n_rows));
    if (rc) return rc;
    h->feat_rows = -1;  // (the scratch slot overlaps the staging arrays of an unfinished fd_featurize_begin)
    RowArgs a;
    memset(&a, 0, sizeof(a));
    a.host_bytes = -1;
    a.text = text_dev;
    a.offs = offs_dev;
    a.n_rows = n_rows;
    a.flags = flags;
    a.m = h->m;  // (after prepare_fused)
    fill_slot(a, h->ws, n_rows);
    a.pred = out_pred_dev;
    a.prob = out_prob_dev;
    a.raw = out_raw_dev;
    FD_CUDA(h, cudaMemsetAsync(a.ctrl, 0, CTRL_CLEAR_BYTES, st));
    FD_CUDA(h, launch_rows(h, mode, a, st));
    return FD_OK;
}

int fd_vocab_count(fd_handle* h, const uint8_t* text_dev, const int64_t* offs_dev, int64_t n_rows, uint32_t flags,
                   uint64_t* keys_dev, uint32_t* info_dev, int32_t log2_slots, uint8_t* pool_dev, int64_t pool_bytes,
                   uint64_t* counters_dev, void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (n_rows < 0 || (n_rows > 0 && (!text_dev || !offs_dev))) return set_err(h, FD_ERR_INVALID, "bad text arguments");
    if (((uintptr_t)text_dev & 15) != 0) return set_err(h, FD_ERR_INVALID, "text pointer must be 16-byte aligned");
    if (!keys_dev || !info_dev || !counters_dev || log2_slots < 10 || log2_slots > 30 || pool_bytes < 0 || (pool_bytes > 0 && !pool_dev))
        return set_err(h, FD_ERR_INVALID, "bad term-table arguments");
    if (n_rows == 0) return FD_OK;
    FD_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    int rc = ensure_ws(h, slot_bytes(n_rows));
    if (rc) return rc;
    h->feat_rows = -1;  // (as in fd_transform)
    RowArgs a;
    memset(&a, 0, sizeof(a));
    a.host_bytes = -1;
    a.text = text_dev;
    a.offs = offs_dev;
    a.n_rows = n_rows;
    a.flags = flags;
    a.m = h->m;
    a.m.vocab_cap = 0;  // counting terms: neither the hashing trick nor an earlier vocabulary applies
    a.m.binary = 0;
    a.m.drop_empty = a.m.empty_is_stop;
    fill_slot(a, h->ws, n_rows);
    a.fit_keys = reinterpret_cast<unsigned long long*>(keys_dev);
    a.fit_info = reinterpret_cast<uint4*>(info_dev);
    a.fit_log2 = log2_slots;
    a.fit_pool = pool_dev;
    a.fit_pool_cap = pool_bytes;
    a.fit_counters = reinterpret_cast<unsigned long long*>(counters_dev);
    FD_CUDA(h, cudaMemsetAsync(a.ctrl, 0, CTRL_CLEAR_BYTES, st));
    FD_CUDA(h, launch_rows(h, MODE_FIT, a, st));
    return FD_OK;
}

int64_t fd_workspace_bytes(fd_handle* h, int64_t n_rows) {
    if (!h || n_rows < 0) return FD_ERR_INVALID;
    return (int64_t)std::max(slot_bytes(n_rows), feat_layout(n_rows, std::max<int64_t>(h->tmp_cap, feat_guess(n_rows))).total);
}

int fd_reserve(fd_handle* h, int64_t n_rows) {
    if (!h || n_rows < 0) return set_err(h, FD_ERR_INVALID, "n_rows < 0");
    FD_CUDA(h, cudaSetDevice(h->device));
    return ensure_ws(h, (size_t)fd_workspace_bytes(h, n_rows));
}

int fd_status(fd_handle* h, void* stream) {
    if (!h) return FD_ERR_INVALID;
    if (!h->ws) return FD_OK;
    FD_CUDA(h, cudaSetDevice(h->device));
    Ctrl c;
    int rc = read_ctrl(h, reinterpret_cast<Ctrl*>(h->ws), &c, (cudaStream_t)stream);
    if (rc) return rc;
    if (c.status != FD_OK) {
        cudaMemsetAsync(&reinterpret_cast<Ctrl*>(h->ws)->status, 0, sizeof(int), (cudaStream_t)stream);
        return set_err(h, c.status, "a dialogue exceeds the supported number of distinct hashed terms (%d)",
                       (1 << BIG_LOG2) * 3 / 4);
    }
    return FD_OK;
}

int fd_transform_host(fd_handle* h, const uint8_t* text_host, const int64_t* offs_host, int64_t n_rows, uint32_t flags,
                      double* out_pred_host, double* out_prob_host, double* out_raw_host) {
    if (!h) return FD_ERR_INVALID;
    if (n_rows < 0 || (n_rows > 0 && (!text_host || !offs_host || !out_pred_host || !out_prob_host || !out_raw_host)))
        return set_err(h, FD_ERR_INVALID, "bad fd_transform_host arguments");
    if (h->m.num_features <= 0) return set_err(h, FD_ERR_STATE, "HashingTF stage not set (fd_set_hashing)");
    int mode;
    int rc = classifier_mode(h, &mode);
    if (rc) return rc;
    if (n_rows == 0) return FD_OK;
    FD_CUDA(h, cudaSetDevice(h->device));
    rc = prepare_fused(h);
    if (rc) return rc;
    for (int i = 0; i < 3; ++i)
        if (!h->hs[i]) FD_CUDA(h, cudaStreamCreateWithFlags(&h->hs[i], cudaStreamNonBlocking));
    if (!h->pinned_ctrl) FD_CUDA(h, cudaMallocHost((void**)&h->pinned_ctrl, 4 * sizeof(Ctrl)));

    // ---- micro-batches whose buffers are pinned (mapped) host memory: no copies at all.  The kernels read the dialogues
    // and offsets from host memory through the same TMA windows and write the three columns straight back, so the transfer
    // overlaps the work window by window and the call is two launches, the status read and one synchronise -- instead
    // of two H2D and three D2H copies around them.  (A 16-byte block never straddles a page, so the kernels' reads up to the
    // next multiple of 16 stay inside the pinned pages.)  Larger batches keep the chunked three-stream copy pipeline below.
    constexpr int64_t DIRECT_MAX_BYTES = 8 << 20;
    if (n_rows <= (1 << 16) && offs_host[n_rows] - offs_host[0] <= DIRECT_MAX_BYTES && h->direct_host) {
        const void* dp[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
        const void* hp[5] = {text_host, offs_host, out_pred_host, out_prob_host, out_raw_host};
        bool ok = ((uintptr_t)text_host & 15) == 0;
        for (int i = 0; i < 5 && ok; ++i) {
            cudaPointerAttributes at;
            if (cudaPointerGetAttributes(&at, hp[i]) != cudaSuccess) {
                cudaGetLastError();
                ok = false;
            } else {
                ok = at.type == cudaMemoryTypeHost && at.devicePointer != nullptr;
                dp[i] = at.devicePointer;
            }
        }
        if (ok && ((uintptr_t)dp[0] & 15) == 0) {
            const size_t need = slot_bytes(n_rows);
            if (need > h->hbuf_bytes[0]) {
                if (h->hbuf[0]) {
                    FD_CUDA(h, cudaStreamSynchronize(h->hs[0]));
                    FD_CUDA(h, cudaFree(h->hbuf[0]));
                    h->hbuf[0] = nullptr;
                }
                FD_CUDA(h, cudaMalloc((void**)&h->hbuf[0], need));
                h->hbuf_bytes[0] = need;
                FD_CUDA(h, cudaMemsetAsync(h->hbuf[0], 0, 256, h->hs[0]));
            }
            cudaStream_t st = h->hs[0];
            RowArgs a;
            memset(&a, 0, sizeof(a));
            fill_slot(a, h->hbuf[0], n_rows);
            a.text = static_cast<const uint8_t*>(dp[0]);
            a.offs = static_cast<const int64_t*>(dp[1]);
            a.n_rows = n_rows;
            a.flags = flags;
            a.host_bytes = offs_host[n_rows] - offs_host[0];
            a.m = h->m;  // (after prepare_fused)
            a.pred = static_cast<double*>(const_cast<void*>(dp[2]));
            a.prob = static_cast<double*>(const_cast<void*>(dp[3]));
            a.raw = static_cast<double*>(const_cast<void*>(dp[4]));
            FD_CUDA(h, cudaMemsetAsync(a.ctrl, 0, CTRL_CLEAR_BYTES, st));
            FD_CUDA(h, launch_rows(h, mode, a, st));
            FD_CUDA(h, cudaMemcpyAsync(h->pinned_ctrl + 1, h->hbuf[0], sizeof(Ctrl), cudaMemcpyDeviceToHost, st));
            FD_CUDA(h, cudaStreamSynchronize(st));
            if (h->pinned_ctrl[1].status != FD_OK) {
                const int status = h->pinned_ctrl[1].status;
                cudaMemsetAsync(&reinterpret_cast<Ctrl*>(h->hbuf[0])->status, 0, sizeof(int), st);
                return set_err(h, status, "a dialogue exceeds the supported number of distinct hashed terms (%d)", (1 << BIG_LOG2) * 3 / 4);
            }
            return FD_OK;
        }
    }

    constexpr int64_t CHUNK_BYTES = 64 << 20;  // pinned copies reach their full rate (55.5 GB/s here) from about 64 MiB; 24 MiB chunks ran at ~50
    constexpr int64_t CHUNK_ROWS = 1 << 18;
    int64_t r0 = 0;
    int c = 0;
    bool used[3] = {false, false, false};
    while (r0 < n_rows) {
        // rows [r0, r1): at most CHUNK_ROWS rows and (unless a single dialogue is larger) CHUNK_BYTES bytes
        int64_t r1 = std::min(n_rows, r0 + CHUNK_ROWS);
        const int64_t lim = offs_host[r0] + CHUNK_BYTES;
        if (offs_host[r1] > lim) {
            r1 = std::upper_bound(offs_host + r0, offs_host + r1 + 1, lim) - offs_host - 1;
            if (r1 <= r0) r1 = r0 + 1;
        }
        const int64_t nr = r1 - r0, b0 = offs_host[r0], nbytes = offs_host[r1] - b0;
        const int s = c % 3;
        const size_t d = (size_t)(b0 & 15);
        const size_t off_text = slot_bytes(nr);
        const size_t off_offs = align_up(off_text + d + (size_t)nbytes + 32, 256);
        const size_t off_pred = align_up(off_offs + (size_t)(nr + 1) * 8, 256);
        const size_t off_prob = off_pred + (size_t)nr * 8;
        const size_t off_raw = off_prob + (size_t)nr * 16;
        const size_t need = off_raw + (size_t)nr * 16;
        if (need > h->hbuf_bytes[s]) {
            if (h->hbuf[s]) {
                FD_CUDA(h, cudaStreamSynchronize(h->hs[s]));
                FD_CUDA(h, cudaFree(h->hbuf[s]));
                h->hbuf[s] = nullptr;
            }
            const size_t sz = std::max(need, slot_bytes(CHUNK_ROWS) + (size_t)CHUNK_BYTES + (size_t)CHUNK_ROWS * 48 + 4096);
            FD_CUDA(h, cudaMalloc((void**)&h->hbuf[s], sz));
            h->hbuf_bytes[s] = sz;
            FD_CUDA(h, cudaMemsetAsync(h->hbuf[s], 0, 256, h->hs[s]));
        }
        char* base = h->hbuf[s];
        cudaStream_t st = h->hs[s];
        RowArgs a;
        memset(&a, 0, sizeof(a));
    a.host_bytes = -1;
        fill_slot(a, base, nr);
        if (nbytes > 0)
            FD_CUDA(h, cudaMemcpyAsync(base + off_text + d, text_host + b0, (size_t)nbytes, cudaMemcpyHostToDevice, st));
        FD_CUDA(h, cudaMemcpyAsync(base + off_offs, offs_host + r0, (size_t)(nr + 1) * 8, cudaMemcpyHostToDevice, st));
        a.text = reinterpret_cast<const uint8_t*>(base + off_text + d) - b0;  // so that text + offs[r] is the dialogue
        a.offs = reinterpret_cast<const int64_t*>(base + off_offs);
        a.n_rows = nr;
        a.flags = flags;
        a.host_bytes = nbytes;  // (the offsets are host memory here: the kernel shape is picked without a device round trip)
        a.m = h->m;  // (after prepare_fused)
        a.pred = reinterpret_cast<double*>(base + off_pred);
        a.prob = reinterpret_cast<double*>(base + off_prob);
        a.raw = reinterpret_cast<double*>(base + off_raw);
        FD_CUDA(h, cudaMemsetAsync(a.ctrl, 0, CTRL_CLEAR_BYTES, st));
        FD_CUDA(h, launch_rows(h, mode, a, st));
        FD_CUDA(h, cudaMemcpyAsync(out_pred_host + r0, a.pred, (size_t)nr * 8, cudaMemcpyDeviceToHost, st));
        FD_CUDA(h, cudaMemcpyAsync(out_prob_host + 2 * r0, a.prob, (size_t)nr * 16, cudaMemcpyDeviceToHost, st));
        FD_CUDA(h, cudaMemcpyAsync(out_raw_host + 2 * r0, a.raw, (size_t)nr * 16, cudaMemcpyDeviceToHost, st));
        used[s] = true;
        r0 = r1;
        ++c;
    }
    int status = FD_OK;
    for (int s = 0; s < 3; ++s) {
        if (!used[s]) continue;
        FD_CUDA(h, cudaMemcpyAsync(h->pinned_ctrl + 1 + s, h->hbuf[s], sizeof(Ctrl), cudaMemcpyDeviceToHost, h->hs[s]));
        FD_CUDA(h, cudaStreamSynchronize(h->hs[s]));
        if (h->pinned_ctrl[1 + s].status != FD_OK) {
            status = h->pinned_ctrl[1 + s].status;
            cudaMemsetAsync(&reinterpret_cast<Ctrl*>(h->hbuf[s])->status, 0, sizeof(int), h->hs[s]);
        }
    }
    if (status != FD_OK)
        return set_err(h, status, "a dialogue exceeds the supported number of distinct hashed terms (%d)", (1 << BIG_LOG2) * 3 / 4);
    return FD_OK;
}

}  // extern "C"
