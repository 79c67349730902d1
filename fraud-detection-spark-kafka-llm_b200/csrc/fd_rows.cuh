// fd_rows.cuh -- shared declarations of the dialogue kernels (fd_rows.cu) and their launchers.
#pragma once
#include "fd_common.cuh"

namespace fd {

// Two shapes of the dialogue kernel, chosen per launch:
//   TL = 8: 256-slot count tables, 16 warps per CTA -> 32 warps per SM; for large batches, where occupancy pays
//   TL = 9: 512-slot count tables,  8 warps per CTA -> 24 warps per SM; for micro-batches, where fewer dialogues
//           overflow into the wide-table kernel (which runs after the main one and would add to the latency)
__host__ __device__ constexpr int row_warps(int tl) { return tl == 8 ? 16 : 8; }
constexpr long long ROWS_FOR_SMALL_TABLES = 65536;  // batches at least this large may use TL = 8 ...
constexpr long long MEAN_BYTES_FOR_SMALL_TABLES = 1400;  // ... when their dialogues average at most this many bytes
                                                         // (longer ones would mostly overflow a 256-slot table)
constexpr int ROW_CH = 1024;                 // bytes per staged window (32 per lane)
constexpr int ROW_EXT = 64;                  // a dialogue's LAST window may run this much longer (one word per lane, lanes 0..15):
                                             // a short remainder is absorbed instead of costing a whole extra window
constexpr int ROW_WIN = ROW_CH + ROW_EXT;
constexpr int ROW_RAWBUF = ROW_WIN + 16;     // + slack for word over-reads at the end
// Dialogues that overflow the shared-memory table are deferred to wide tables in global memory, in two classes
// by size (a dialogue of b bytes has at most b/2 + 1 tokens, which bounds the table it can need):
//   medium: table <= 2^MED_LOG2 slots (32 KiB, stays in L2): 8 warps per CTA, MED_CTAS CTAs -> thousands of warps
//   large : table <= 2^BIG_LOG2 slots: one warp per CTA, BIG_WORKERS CTAs (rare: dialogues beyond ~6 KB)
constexpr int MED_LOG2 = 12;
constexpr int MED_WARPS = 8;
constexpr int MED_CTAS = 296;
constexpr int BIG_LOG2 = 16;
constexpr int BIG_WORKERS = 48;   // (the two variants run side by side in one launch: each has its own table pool)
// Micro-batch / long-dialogue shape (TL = 9): a warp whose shared-memory table fills up carries on in its OWN 2^MED_LOG2-slot
// table in global memory (the entries move over, the dialogue is not restarted); only dialogues that outgrow that one as
// well are deferred.  One table per warp of the grid, SPILL_WARPS at most.
constexpr int SPILL_WARPS = 160 * 3 * 8;
constexpr size_t SPILL_SLOTS = (size_t)SPILL_WARPS << MED_LOG2;
constexpr size_t med_pool_slots(int med_ctas) { return (size_t)med_ctas * MED_WARPS << MED_LOG2; }
constexpr size_t big_pool_slots(int big_workers) { return (size_t)big_workers << BIG_LOG2; }
// one pool for both when they run one after the other (fd_transform_noalloc), two disjoint pools when they run together
constexpr size_t wide_pool_slots(int med_ctas, int big_workers) {
    return med_pool_slots(med_ctas) > big_pool_slots(big_workers) ? med_pool_slots(med_ctas) : big_pool_slots(big_workers);
}

// the kernel shape for a large batch, from the mean dialogue length (device side: every CTA of both shapes evaluates it)
__host__ __device__ inline int shape_for(long long total_bytes, long long n_rows, float keep) {
    const long long mean = total_bytes / (n_rows > 0 ? n_rows : 1);
    return (float)mean * keep <= (float)MEAN_BYTES_FOR_SMALL_TABLES ? 8 : 9;
}

enum Mode { MODE_CSR = 0, MODE_LINEAR = 1, MODE_DT = 2, MODE_RF = 3, MODE_GBT = 4, MODE_FIT = 5 };

template <int TL>
struct __align__(16) WarpSmem {
    uint8_t raw[2][ROW_RAWBUF];        // TMA-staged windows (double-buffered)
    uint32_t lmask[36];                // letter bitmask of the current window (+ zero tail)
    uint16_t tok[ROW_WIN + 8];         // token start offsets of the current window; epilogue scratch
    unsigned long long mbar[2];        // one mbarrier per window buffer
    unsigned long long tab[1 << TL];   // count in the low word, bucket in the high word
};
static_assert(sizeof(WarpSmem<8>) % 16 == 0 && sizeof(WarpSmem<9>) % 16 == 0, "WarpSmem must keep 16-byte alignment");

struct RowArgs {
    const uint8_t* text;
    const int64_t* offs;
    long long n_rows;
    uint32_t flags;
    Model m;
    Ctrl* ctrl;
    long long* big_rows;
    long long big_cap;
    int pick_shape;  // 1: both main-kernel shapes are launched and each decides from the mean dialogue length whether it runs
    float pick_keep; // ... scaled by the fraction of a dialogue's terms that reach the count table (sparse LogisticRegression)
    long long host_bytes;  // total text bytes when the caller knows them on the host (the shape is then picked there), else -1
    int dyn_rows;    // micro-batches: after its first dialogue a warp takes the next one from Ctrl::next_item (a warp that
                     // holds a 20 KB dialogue then gets no second one while others sit idle); set by launch_main_rows
    unsigned long long* big_tabs;    // table pool of the medium wide-table variant
    unsigned long long* big2_tabs;   // ... of the large one (may equal big_tabs when the two variants run one after the other)
    unsigned long long* spill_tabs;  // SPILL_SLOTS slots, or nullptr (no spilling: overflowing dialogues are deferred)
    // featurize outputs (staging)
    int32_t* row_nnz;
    long long* row_tmpoff;
    unsigned long long* tmp;
    long long tmp_cap;
    // CountVectorizer.fit (MODE_FIT): global term table owned by the caller
    unsigned long long* fit_keys;   // [1 << fit_log2], ~0 = empty
    uint4* fit_info;                // per slot: x = term count, y = document frequency, z = pool offset, w = length
    int fit_log2;
    uint8_t* fit_pool;              // bytes of terms longer than 7 bytes
    long long fit_pool_cap;
    unsigned long long* fit_counters;  // [0] = entries, [1] = pool bytes used, [2] = overflow flag
    // score outputs
    double* pred;
    double* prob;
    double* raw;
};

cudaError_t launch_rows(fd_handle* h, int mode, const RowArgs& a, cudaStream_t st);
cudaError_t launch_main_rows(fd_handle* h, int mode, const RowArgs& a, cudaStream_t st);  // fd_main.cu
cudaError_t launch_rows_workers(fd_handle* h, int mode, const RowArgs& a, int med_ctas, int big_workers, bool merged, cudaStream_t st);
cudaError_t launch_csr_assembly(fd_handle* h, const int32_t* nnz, const long long* tmpoff, unsigned long long* tmp,
                                long long n_rows, long long* bsum, long long* rowptr, int32_t* out_idx, double* out_val,
                                bool scan_only, long long unsorted_rows, cudaStream_t st);
cudaError_t launch_score_csr(fd_handle* h, const int64_t* rowptr, const int32_t* idx, const double* val, long long n_rows,
                             double* pred, double* prob, double* raw, cudaStream_t st);

}  // namespace fd
