// fd_main.cu -- the main dialogue kernel: clean -> tokenise -> stop-word filter -> murmur3 HashingTF -> per-dialogue
// term counts -> CSR emission or a fused classifier epilogue, for every dialogue whose count table fits in shared
// memory (the others are deferred to the wide-table kernels of fd_rows.cu).
//
// One WARP owns one dialogue at a time.  Its UTF-8 bytes are staged from HBM into the warp's shared-memory windows
// (1 KiB, double-buffered) with TMA bulk copies (cp.async.bulk + mbarrier); the next window -- of this dialogue or of
// the warp's next one -- is in flight while the current one is hashed.  Per window:
//   pass A  each lane classifies 32 bytes with SWAR arithmetic (kept byte / separator / dropped) and folds case in
//           place; the per-byte flags of TWO words are gathered into eight mask bits by one multiply; then the list of
//           token start offsets is written (one entry after every separator);
//   pass B  one token per lane: the first 12 bytes come out of shared memory with four word loads and three funnel
//           shifts, the block masks from a 13-row table, MurmurHash3_x86_32(seed 42) runs straight-line, then
//             * a small decision tree (<= 63 split features, <= 5 levels) maps the bucket to one of its own counters
//               (perfect hash) and drops every other term after one compare;
//             * a classifier that sees few features (sparse LogisticRegression, forests) tests one bit of Model::filt
//               and drops the term, or inserts it into the warp's (bucket -> count) table;
//             * featurize (CSR) probes the stop-word table (perfect hash keyed by the same murmur3 value) and inserts.
//           Stop words never reach a classifier through a feature it cannot see, so on the classifier paths the stop-word
//           table is only probed for the rare bucket that both the classifier sees and a stop word maps to.
// End of dialogue: the small decision tree evaluates all its split nodes at once (one per lane, thresholds pre-quantised
// to term counts), walks the ballot and lanes 0..4 store the five result doubles; the other modes run the shared
// epilogues of fd_rows_dev.cuh.
//
// Semantics restated from Spark 3.5.5; reference call sites:
// utils/agent_api.py:143-144 (cleaning), fraud_detection_spark.py:50-53 (Tokenizer, StopWordsRemover, HashingTF/IDF
// stages), utils/agent_api.py:158-167 (transform + result columns).
#include <cstddef>
#include <type_traits>

#include "fd_rows_dev.cuh"

namespace fd {

struct __align__(16) MainWarp {
    uint8_t raw[2][ROW_RAWBUF];        // TMA-staged windows (double-buffered)
    uint32_t lmask[36];                // kept-byte bitmask of the current window (+ zero tail)
    uint16_t tok[ROW_WIN + 8];         // token start offsets of the current window; epilogue scratch
    unsigned long long mbar[2];        // one mbarrier per window buffer
};                                     // the warp's count table (or counters) follows
static_assert(sizeof(MainWarp) % 16 == 0, "MainWarp must keep 16-byte alignment");

constexpr int LUT_BYTES = 256;  // 13 rows of block masks

// flags at bit 7 of every byte of two words (other bits: anything) -> eight mask bits in the TOP byte of the result:
// bits 24..27 = bytes 0..3 of `lo`, bits 28..31 = bytes 0..3 of `hi`.  (lo & 0x80808080) >> 4 | (hi & 0x80808080) leaves
// the eight flags at bits 3, 7, ..., 31; the multiplier 2^21 + 2^14 + 2^7 + 1 moves them next to each other (no two
// partial products meet on one bit -- bit distances are multiples of 4, shift distances multiples of 7 -- so nothing
// carries).
__device__ __forceinline__ uint32_t gather8(uint32_t lo, uint32_t hi) {
    return (((lo & 0x80808080u) >> 4) | (hi & 0x80808080u)) * 0x00204081u;
}
__device__ __forceinline__ uint32_t top_bytes(uint32_t p0, uint32_t p1, uint32_t p2, uint32_t p3) {
    return __byte_perm(__byte_perm(p0, p1, 0x0073), __byte_perm(p2, p3, 0x0073), 0x5410);
}

// SWAR classification of four bytes; bit 7 of every byte of l / s says "kept byte" / "separator" (other bits: garbage);
// nx = the word with ASCII upper-case letters folded; H collects every byte (bit 7 set somewhere: non-ASCII present).
//   CLEAN : lower() + regexp_replace("[^a-zA-Z ]", "") + split("\\s"): letters are kept, ' ' separates, the rest vanishes
//   !CLEAN: Tokenizer on already-clean text: \t \n \v \f \r ' ' separate, every other byte is kept
template <bool CLEAN>
__device__ __forceinline__ void classify4(uint32_t x, uint32_t& l, uint32_t& s, uint32_t& nx) {
    const uint32_t z = x ^ 0x20202020u;
    const uint32_t sp = ~(((z & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | z);  // bit 7 <=> byte == ' '
    if (CLEAN) {
        const uint32_t v7 = (x | 0x20202020u) & 0x7F7F7F7Fu;
        l = (v7 + 0x1F1F1F1Fu) & ~(v7 + 0x05050505u) & ~x;  // bit 7 <=> 'a' <= (byte | 0x20) <= 'z', byte < 0x80
        s = sp;
        nx = x | ((l >> 2) & 0x20202020u);
    } else {
        const uint32_t t = x & 0x7F7F7F7Fu;
        s = ((t + 0x77777777u) & ~(t + 0x72727272u) & ~x) | sp;  // 0x09..0x0D or ' '
        l = ~s;
        const uint32_t up = (t + 0x3F3F3F3Fu) & ~(t + 0x25252525u) & ~x;  // 'A'..'Z'
        nx = x | ((up >> 2) & 0x20202020u);
    }
}

// long_stop_word_equal with the window given by 32-bit shared addresses: the callers' hot loops then never hold (or rebuild,
// at twenty instructions a time) the generic pointers that only this rare comparison needs
static __device__ __noinline__ bool long_stop_word_equal_s(const uint8_t* word, uint32_t rbA, uint32_t lmA, int s, int e) {
    return long_stop_word_equal(word, reinterpret_cast<const uint8_t*>(__cvta_shared_to_generic(rbA)),
                                reinterpret_cast<const uint32_t*>(__cvta_shared_to_generic(lmA)), s, e);
}

// is the in-window token a stop word?  h = murmur3 of its kept bytes, n = their number, b0..b2 = the first 12 of them;
// [s, e) locates it in the window (for the byte compare of words longer than 12 bytes).  cand: the lane holds a non-empty
// token at all.  Straight-line: the word record is loaded unconditionally, only long words call out.
__device__ __forceinline__ bool stop_probe(const Model& m, const uint32_t* swt, const uint4* sww, uint32_t h, uint32_t n, uint32_t b0,
                                           uint32_t b1, uint32_t b2, uint32_t rbA, uint32_t lmA, int s, int e, bool cand) {
    bool stop = false;
    if (m.sw_perfect) {  // one probe: 16-bit slot -> word id, then the record is compared
        const uint32_t id = reinterpret_cast<const uint16_t*>(swt)[(h * m.sw_mul) >> m.sw_shift];
        const uint4 w = sww[id ? id - 1u : 0u];
        const bool c = cand && id != 0u;
        stop = c && (w.x == n) && (w.y == b0 && w.z == b1 && w.w == b2);
        if (c && w.x == (n | 0x80000000u)) stop = long_stop_word_equal_s(m.sw_pool + w.y, rbA, lmA, s, e);
    } else if (m.sw_cap) {
        const uint32_t swmask = (uint32_t)m.sw_cap - 1u;
        const uint32_t e1 = swt[h & swmask], e2 = swt[(h >> 10) & swmask];
        const bool m1 = e1 != 0u && ((e1 ^ h) >> 16) == 0u, m2 = e2 != 0u && ((e2 ^ h) >> 16) == 0u;
        if (cand && (m1 || m2)) {
            const uint32_t ent = m1 ? e1 : e2;
            uint4 w = sww[(ent & 0xFFFFu) - 1u];
            stop = (w.x == n) && (w.y == b0 && w.z == b1 && w.w == b2);
            if ((w.x & 0x7FFFFFFFu) == n && (w.x & 0x80000000u)) stop = long_stop_word_equal_s(m.sw_pool + w.y, rbA, lmA, s, e);
            if (!stop && m1 && m2) {  // fingerprint collision in the first slot: look at the second
                w = sww[(e2 & 0xFFFFu) - 1u];
                stop = (w.x == n) && (w.y == b0 && w.z == b1 && w.w == b2);
                if ((w.x & 0x7FFFFFFFu) == n && (w.x & 0x80000000u)) stop = long_stop_word_equal_s(m.sw_pool + w.y, rbA, lmA, s, e);
            }
        }
    }
    return stop;
}

// U+0130 (C4 B0) -> 'i', U+212A (E2 84 AA) -> 'k': the only non-ASCII characters whose lower-case forms keep an ASCII
// letter.  bits = valid bytes of the group starting at buffer offset p0; returns the kept-byte bits to add.
__device__ __noinline__ uint32_t fix_non_ascii(uint8_t* rb, int p0, uint32_t bits, int nbits, int lbmin) {
    uint32_t add = 0;
    for (int j = 0; j < nbits; ++j) {
        const int p = p0 + j;
        if (!((bits >> j) & 1u)) continue;
        const uint8_t c = rb[p];
        if (c == 0xB0 && p - 1 >= lbmin && rb[p - 1] == 0xC4) {
            rb[p] = 'i';
            add |= 1u << j;
        } else if (c == 0xAA && p - 2 >= lbmin && rb[p - 1] == 0x84 && rb[p - 2] == 0xE2) {
            rb[p] = 'k';
            add |= 1u << j;
        }
    }
    return add;
}

// ---- shared-memory accesses by 32-bit shared address.  The hot loops address the warp's buffers through a few base
// addresses held in registers; going through generic pointers made the compiler rebuild the shared window base (three
// uniform-datapath instructions) at every use.
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ uint4 lds_v4(uint32_t a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts_u16(uint32_t a, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void red_add_u32(uint32_t a, uint32_t v) { asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ int bfind_u32(uint32_t x) {  // index of the highest set bit (x != 0)
    int b;
    asm("bfind.u32 %0, %1;" : "=r"(b) : "r"(x));
    return b;
}
// rotate left by a constant as ONE funnel shift (written in PTX so that "multiply, then rotate" is not turned into
// two multiplies and a LEA.HI)
template <int R>
__device__ __forceinline__ uint32_t rotl_c(uint32_t x) {
    uint32_t y;
    asm("shf.l.wrap.b32 %0, %1, %1, %2;" : "=r"(y) : "r"(x), "n"(R));
    return y;
}
__device__ __forceinline__ uint32_t mmk(uint32_t k) { return rotl_c<15>(k * 0xcc9e2d51u) * 0x1b873593u; }

// stop_probe with the tables addressed by 32-bit shared address (swtA: slots, swwA: 16-byte word records): the hot form
// for featurize and dense classifiers, where every token is probed.  The record of slot value `id` sits at
// swwA + 16 (id - 1); id = 0 (no word) reads the 16 bytes before the records -- valid shared memory -- and is masked out.
__device__ __forceinline__ bool stop_probe_s(const Model& m, uint32_t swtA, uint32_t swwA, uint32_t h, uint32_t n, uint32_t b0, uint32_t b1,
                                             uint32_t b2, uint32_t rbA, uint32_t lmA, int s, int e, bool cand) {
    bool stop = false;
    if (m.sw_perfect) {
        const uint32_t id = lds_u16(swtA + 2u * ((h * m.sw_mul) >> m.sw_shift));
        const uint4 w = lds_v4(swwA + 16u * id - 16u);
        const bool c = cand && id != 0u;
        stop = c && (w.x == n) && (w.y == b0 && w.z == b1 && w.w == b2);
        if (c && w.x == (n | 0x80000000u)) stop = long_stop_word_equal_s(m.sw_pool + w.y, rbA, lmA, s, e);
    } else if (m.sw_cap) {
        const uint32_t swmask = (uint32_t)m.sw_cap - 1u;
        const uint32_t e1 = lds_u32(swtA + 4u * (h & swmask)), e2 = lds_u32(swtA + 4u * ((h >> 10) & swmask));
        const bool m1 = e1 != 0u && ((e1 ^ h) >> 16) == 0u, m2 = e2 != 0u && ((e2 ^ h) >> 16) == 0u;
        if (cand && (m1 || m2)) {
            const uint32_t ent = m1 ? e1 : e2;
            uint4 w = lds_v4(swwA + 16u * (ent & 0xFFFFu) - 16u);
            stop = (w.x == n) && (w.y == b0 && w.z == b1 && w.w == b2);
            if ((w.x & 0x7FFFFFFFu) == n && (w.x & 0x80000000u)) stop = long_stop_word_equal_s(m.sw_pool + w.y, rbA, lmA, s, e);
            if (!stop && m1 && m2) {  // fingerprint collision in the first slot: look at the second
                w = lds_v4(swwA + 16u * (e2 & 0xFFFFu) - 16u);
                stop = (w.x == n) && (w.y == b0 && w.z == b1 && w.w == b2);
                if ((w.x & 0x7FFFFFFFu) == n && (w.x & 0x80000000u)) stop = long_stop_word_equal_s(m.sw_pool + w.y, rbA, lmA, s, e);
            }
        }
    }
    return stop;
}

__device__ __forceinline__ uint32_t atoms_add_u32(uint32_t a, uint32_t v) {  // returns the old value
    uint32_t old;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(a), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ void sts_v4(uint32_t a, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// featurize with a TOKEN LIST (Model::csr_list: hashing trick, numFeatures <= 2^22): pass B only appends the bucket of
// every surviving token to a list in shared memory; at the end of the dialogue the list is ordered by a counting sort
// (256 bins by the bucket's top bits -- hashed buckets are evenly spread --, a shared-memory atomic gives every token its
// arrival rank inside its bin, a scan the bins' start positions, tokens that share a bin are placed by counting the bin's
// smaller keys and earlier equals), runs of equal buckets become (bucket, count) entries, and the dialogue's row goes to
// the staging buffer already ORDERED (flagged in row_tmpoff), so that k_gather_csr only forms the values.
// listA: n keys (u32); binsA: 256 words + later the run starts (u16); outA: room for n keys.  Returns the row's nnz.
constexpr long long ROW_SORTED = 1LL << 62;  // flag in row_tmpoff
constexpr int CSR_LIST_MAX = 544;            // tokens a list may hold (the ordered copy lives in the token-list area)
template <bool POW2>
__device__ __forceinline__ uint32_t bucket_bin(const Model& m, uint32_t key, int bin_shift) {
    if (POW2) return key >> bin_shift;
    return m.bin_mul ? __umulhi(key, m.bin_mul) : key;
}
template <bool POW2>
__device__ int csr_sorted_emit(const RowArgs& a, uint32_t listA, int n, uint32_t binsA, uint32_t outA, long long row, int lane, int bin_shift) {
    constexpr uint32_t FULL = 0xFFFFFFFFu;
    const Model& m = a.m;
    const uint32_t below = (1u << lane) - 1u;
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    sts_v4(binsA + 32u * (uint32_t)lane, z);
    sts_v4(binsA + 32u * (uint32_t)lane + 16u, z);
    __syncwarp();
    for (int e = lane; e < n; e += 32) {  // arrival rank inside the bin, kept in the key's top 10 bits
        const uint32_t key = lds_u32(listA + 4u * (uint32_t)e);
        const uint32_t r = atoms_add_u32(binsA + 4u * bucket_bin<POW2>(m, key, bin_shift), 1u);
        sts_u32(listA + 4u * (uint32_t)e, key | (r << 22));
    }
    __syncwarp();
    {   // exclusive scan of the 256 bin sizes: 8 consecutive bins per lane
        const uint4 p = lds_v4(binsA + 32u * (uint32_t)lane), q = lds_v4(binsA + 32u * (uint32_t)lane + 16u);
        const uint32_t tot = p.x + p.y + p.z + p.w + q.x + q.y + q.z + q.w;
        uint32_t inc = tot;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(FULL, inc, d);
            if (lane >= d) inc += v;
        }
        uint32_t run = inc - tot;
        uint4 o0, o1;
        o0.x = run; run += p.x;
        o0.y = run; run += p.y;
        o0.z = run; run += p.z;
        o0.w = run; run += p.w;
        o1.x = run; run += q.x;
        o1.y = run; run += q.y;
        o1.z = run; run += q.z;
        o1.w = run;
        sts_v4(binsA + 32u * (uint32_t)lane, o0);
        sts_v4(binsA + 32u * (uint32_t)lane + 16u, o1);
    }
    __syncwarp();
    for (int e = lane; e < n; e += 32) {  // into the bins (arrival order inside a bin)
        const uint32_t w = lds_u32(listA + 4u * (uint32_t)e);
        const uint32_t key = w & 0x3FFFFFu;
        sts_u32(outA + 4u * (lds_u32(binsA + 4u * bucket_bin<POW2>(m, key, bin_shift)) + (w >> 22)), key);
    }
    __syncwarp();
    // A bin usually holds the repetitions of ONE bucket (then its tokens already sit together): bins with two or more
    // different buckets are marked (bit 31 of the bin's start), and only their tokens look for their final place
    for (int e = lane; e < n; e += 32) {
        const uint32_t key = lds_u32(outA + 4u * (uint32_t)e);
        const uint32_t ba = binsA + 4u * bucket_bin<POW2>(m, key, bin_shift);
        if (lds_u32(outA + 4u * (lds_u32(ba) & 0xFFFFu)) != key) asm volatile("red.shared.or.b32 [%0], %1;" ::"r"(ba), "r"(0x80000000u) : "memory");
    }
    __syncwarp();
    for (int e = lane; e < n; e += 32) {  // final place: smaller keys of the bin first, equal keys in position order
        const uint32_t key = lds_u32(outA + 4u * (uint32_t)e);
        const uint32_t bin = bucket_bin<POW2>(m, key, bin_shift);
        const uint32_t b0 = lds_u32(binsA + 4u * bin);
        uint32_t pos = (uint32_t)e;
        if (b0 >> 31) {
            const uint32_t s0 = b0 & 0xFFFFu, s1 = bin == 255u ? (uint32_t)n : (lds_u32(binsA + 4u * bin + 4u) & 0xFFFFu);
            pos = s0;
            for (uint32_t j = s0; j < s1; ++j) {
                const uint32_t kj = lds_u32(outA + 4u * j);
                pos += (kj < key || (kj == key && j < (uint32_t)e)) ? 1u : 0u;
            }
        }
        sts_u32(listA + 4u * pos, key);
    }
    __syncwarp();
    // runs of equal buckets: start[d] = first position of the d-th distinct bucket (u16, over the bins' area)
    uint32_t D = 0;
    for (int c = 0; c < n; c += 32) {
        const int e = c + lane;
        bool head = false;
        if (e < n) {
            const uint32_t key = lds_u32(listA + 4u * (uint32_t)e);
            head = e == 0 || lds_u32(listA + 4u * (uint32_t)e - 4u) != key;
        }
        const uint32_t bal = __ballot_sync(FULL, head);
        if (head) sts_u16(binsA + 2u * (D + (uint32_t)__popc(bal & below)), (uint32_t)e);
        D += (uint32_t)__popc(bal);
    }
    if (lane == 0) sts_u16(binsA + 2u * D, (uint32_t)n);
    __syncwarp();
    unsigned long long pos = 0;
    if (lane == 0) pos = atomicAdd(&a.ctrl->tmp_cursor, (unsigned long long)D);
    pos = __shfl_sync(FULL, pos, 0);
    const bool fits = (long long)(pos + D) <= a.tmp_cap;
    if (!fits && lane == 0) a.ctrl->tmp_overflow = 1u;
    unsigned long long* dst = a.tmp + pos;
    if (fits) {
        for (uint32_t d = (uint32_t)lane; d < D; d += 32u) {
            const uint32_t s0 = lds_u16(binsA + 2u * d), s1 = lds_u16(binsA + 2u * d + 2u);
            const uint32_t key = lds_u32(listA + 4u * s0);
            dst[d] = ((unsigned long long)key << 32) | (unsigned long long)(m.binary ? 1u : s1 - s0);
        }
    }
    if (lane == 0) {
        a.row_nnz[row] = (int32_t)D;
        a.row_tmpoff[row] = (long long)pos | ROW_SORTED;
    }
    return (int)D;
}

// MODE: what happens to the counted terms (fd_rows.cuh); CLEAN: FD_FLAG_CLEAN; TL: log2 of the count-table slots and the
// CTA shape (row_warps); DTS: small-decision-tree path (MODE_DT with Model::dt_heap_on, hashing trick) -- per-warp counters
// instead of a count table, nothing is ever deferred for size; 2 = some split feature is also a stop word's bucket;
// POW2: HashingTF with a power-of-two numFeatures (bucket = hash & mask).
template <int MODE, bool CLEAN, int TL, int VAR, bool POW2>
__global__ void __launch_bounds__(32 * row_warps(TL), TL == 8 ? 2 : 3) k_main(const RowArgs a, const int cta_bytes, const int warp_bytes) {
    constexpr int DTS = (VAR == 1 || VAR == 2) ? VAR : 0;  // VAR: 0 count table, 1 / 2 small decision tree, 3 token list (featurize)
    constexpr bool csr_list = VAR == 3;                    // featurize: token list instead of a count table
    static_assert(VAR != 3 || MODE == MODE_CSR, "the token list is featurize's");
    constexpr int ROW_WARPS = row_warps(TL);
    constexpr uint32_t FULL = 0xFFFFFFFFu;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const Model& m = a.m;
    if (a.pick_shape && shape_for(a.offs[a.n_rows] - a.offs[0], a.n_rows, a.pick_keep) != TL) return;  // the other shape runs this call

    // ---- CTA-shared tables
    uint4* lut = reinterpret_cast<uint4*>(smem_raw);
    uint32_t* swt = reinterpret_cast<uint32_t*>(smem_raw + LUT_BYTES);                       // !DTS: stop-word slots ...
    uint4* sww = reinterpret_cast<uint4*>(smem_raw + LUT_BYTES + m.sw_tab_bytes());          // ... and word records
    uint32_t* dtk = reinterpret_cast<uint32_t*>(smem_raw + LUT_BYTES);                       // DTS: bucket per counter slot
    double* dto = reinterpret_cast<double*>(smem_raw + LUT_BYTES + ((size_t)4 << m.dt_log2));  // DTS: 32 x 5 result doubles
    unsigned char* wmem = smem_raw + cta_bytes + (size_t)warp * warp_bytes;
    MainWarp& S = *reinterpret_cast<MainWarp*>(wmem);
    unsigned long long* stab = reinterpret_cast<unsigned long long*>(wmem + sizeof(MainWarp));  // !DTS: (count, bucket) slots
    uint32_t* cnt = reinterpret_cast<uint32_t*>(wmem + sizeof(MainWarp));                       // DTS: one counter per slot
    // the same places as shared addresses
    uint32_t lutA;  // (opaque to the compiler, so that it stays in a register instead of being rebuilt at every use)
    asm volatile("mov.u32 %0, %1;" : "=r"(lutA) : "r"(smem_u32(smem_raw)));
    const uint32_t dtkA = lutA + LUT_BYTES;
    const uint32_t swtA = lutA + LUT_BYTES, swwA = swtA + (uint32_t)m.sw_tab_bytes();  // !DTS: the stop-word tables
    const uint32_t rawA = smem_u32(wmem);
    const uint32_t tokA = rawA + (uint32_t)offsetof(MainWarp, tok);
    const uint32_t lmA = rawA + (uint32_t)offsetof(MainWarp, lmask);
    const uint32_t cntA = rawA + (uint32_t)sizeof(MainWarp);

    if (threadIdx.x < 13) {
        const int n = threadIdx.x;
        lut[n] = make_uint4(low_mask(8 * n), low_mask(8 * max(n - 4, 0)), low_mask(8 * max(n - 8, 0)), 0u);
    }
    if (DTS) {
        for (int i = threadIdx.x; i < (1 << m.dt_log2); i += blockDim.x) dtk[i] = m.dt_tab[i];
        for (int i = threadIdx.x; i < 160; i += blockDim.x) dto[i] = m.dt_out[i];
        for (int i = lane; i < (1 << m.dt_log2); i += 32) cnt[i] = 0u;
    } else {
        for (int i = threadIdx.x; i < (int)(m.sw_tab_bytes() / 4); i += blockDim.x) swt[i] = m.sw_tab[i];
        for (int i = threadIdx.x; i < m.n_sw; i += blockDim.x) sww[i] = m.sw_words[i];
        const uint4 ff = make_uint4(0u, ~0u, 0u, ~0u);  // count = 0, key = EMPTY
        for (int i = lane; i < (1 << TL) / 2; i += 32) reinterpret_cast<uint4*>(stab)[i] = ff;
    }
    if (lane == 0) {
        mbar_init(&S.mbar[0], 1);
        mbar_init(&S.mbar[1], 1);
        fence_mbar_init();
    }
    if (lane < 4) S.lmask[32 + lane] = 0u;
    __syncthreads();

    const long long n_items = a.n_rows;
    const long long stride = (long long)gridDim.x * ROW_WARPS;
    long long item = (long long)blockIdx.x * ROW_WARPS + warp;
    if (item >= n_items) return;

    const uint8_t* text = a.text;
    const bool binary = m.binary != 0;
    const uint32_t fmask = (uint32_t)(m.num_features - 1);
    const int bin_shift = max(__popc(fmask) - 8, 0);                // (POW2) bucket -> one of 256 bins
    const int list_cap = min(CSR_LIST_MAX, (8 << TL) / 4);          // keys the table area holds
    unsigned long long* tab = stab;  // TL == 9: may move to the warp's wide table in global memory
    int TLOG2 = TL;
    // DTS: this lane's split node (heap order) and the output column it stores
    int2 hn = make_int2(0, 0x7FFFFFFF);
    double* outp = nullptr;
    int ostride = 0;
    if (DTS) {
        if (lane < 31) hn = m.dt_heap[lane];
        outp = lane == 0 ? a.pred : (lane < 3 ? a.prob + (lane - 1) : a.raw + (lane - 3));
        ostride = lane == 0 ? 1 : 2;
    }

    // ---- window bookkeeping (all warp-uniform)
    long long row = item;
    long long row_beg = a.offs[row], row_end = a.offs[row + 1];
    long long wbase = row_beg & ~15LL;             // global byte offset of buffer byte 0
    int lo = (int)(row_beg - wbase);               // first valid byte of the window
    // a dialogue's last window may run ROW_EXT bytes past the 1 KiB window in the large-batch shape (dialogues of about one
    // window: a short remainder does not cost a whole extra window)
    constexpr int WIN = TL == 8 ? ROW_WIN : ROW_CH;
    bool last = (row_end - wbase) <= WIN;
    int hi = last ? (int)(row_end - wbase) : ROW_CH;
    int lbmin = lo;                                // lowest buffer offset a UTF-8 look-back may read
    int buf = 0;
    uint32_t phase0 = 0, phase1 = 0;
    if (lane == 0) {
        const uint32_t bytes = (uint32_t)((hi + 15) & ~15);
        if (bytes) {
            mbar_expect_tx(&S.mbar[0], bytes);
            tma_load_1d(S.raw[0], text + wbase, bytes, &S.mbar[0]);
        }
    }
    bool cur_loaded = hi > 0;

    // ---- per-dialogue state
    uint32_t D = 0;        // distinct buckets so far
    uint32_t pend = 0;     // empty tokens not yet known to be interior
    uint32_t row_sp = 0;   // separators seen
    bool row_letter = false;
    bool big = false;      // deferred to the wide-table kernels
    bool ext_dirty = false;  // lmask[32..33] hold the extension of an earlier window

    for (;;) {
        uint8_t* rb = S.raw[buf];
        const uint32_t rbA = rawA + (uint32_t)buf * ROW_RAWBUF;
        if (cur_loaded) {
            if (buf == 0) {
                mbar_wait(&S.mbar[0], phase0);
                phase0 ^= 1u;
            } else {
                mbar_wait(&S.mbar[1], phase1);
                phase1 ^= 1u;
            }
        }

        // ================= pass A: classify 32 bytes per lane =================
        uint32_t L, SP, vmask;
        {
            const uint4 q0 = *reinterpret_cast<const uint4*>(rb + 32 * lane);
            const uint4 q1 = *reinterpret_cast<const uint4*>(rb + 32 * lane + 16);
            uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
            uint32_t l[8], s[8];
            uint32_t H = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                H |= w[j];
                classify4<CLEAN>(w[j], l[j], s[j], w[j]);
            }
            const int lo_l = min(max(lo - 32 * lane, 0), 32), hi_l = min(max(hi - 32 * lane, 0), 32);
            vmask = hi_l > lo_l ? (low_mask(hi_l) & ~low_mask(lo_l)) : 0u;
            L = top_bytes(gather8(l[0], l[1]), gather8(l[2], l[3]), gather8(l[4], l[5]), gather8(l[6], l[7])) & vmask;
            SP = top_bytes(gather8(s[0], s[1]), gather8(s[2], s[3]), gather8(s[4], s[5]), gather8(s[6], s[7])) & vmask;
            *reinterpret_cast<uint4*>(rb + 32 * lane) = make_uint4(w[0], w[1], w[2], w[3]);
            *reinterpret_cast<uint4*>(rb + 32 * lane + 16) = make_uint4(w[4], w[5], w[6], w[7]);
            if (CLEAN && (H & 0x80808080u)) L |= fix_non_ascii(rb, 32 * lane, vmask, 32, lbmin);
        }
        S.lmask[lane] = L;
        // a window is "plain" when every valid byte is kept or a separator (already-clean text): tokens are then exactly
        // the gaps between separators and pass B skips the kept-byte analysis
        bool plain = __all_sync(FULL, (L | SP) == vmask);
        // token starts: one after every separator
        const int nsp = __popc(SP);
        int incl = nsp;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int v = __shfl_up_sync(FULL, incl, d);
            if (lane >= d) incl += v;
        }
        int win_sp = __shfl_sync(FULL, incl, 31);
        {
            // this lane's separators, highest first: entry [incl] down to [incl - nsp + 1] of the list
            uint32_t ta = tokA + 2u * (uint32_t)incl;
            uint32_t sp = SP;
            const uint32_t p1 = 32u * (uint32_t)lane + 1u;
            while (sp) {  // two separators per trip
                int b = bfind_u32(sp);
                sts_u16(ta, p1 + (uint32_t)b);
                sp ^= 1u << b;
                if (sp) {
                    b = bfind_u32(sp);
                    sts_u16(ta - 2u, p1 + (uint32_t)b);
                    sp ^= 1u << b;
                }
                ta -= 4u;
            }
        }
        const int my_ll = L ? (32 * lane + 31 - __clz(L)) : -1;
        int last_letter = __reduce_max_sync(FULL, my_ll);
        if (WIN > ROW_CH && hi > ROW_CH) {
            // ---- extension of a dialogue's last window: bytes [ROW_CH, hi), one word per lane on lanes 0..15
            const int p4 = ROW_CH + 4 * lane;
            uint32_t l4 = 0, s4 = 0, v4 = 0;
            if (lane < ROW_EXT / 4) {
                const uint32_t x = *reinterpret_cast<const uint32_t*>(rb + p4);
                uint32_t l80, s80, nx;
                classify4<CLEAN>(x, l80, s80, nx);
                v4 = low_mask(min(max(hi - p4, 0), 4));
                l4 = movemask4(l80 & 0x80808080u) & v4;
                s4 = movemask4(s80 & 0x80808080u) & v4;
                if (nx != x) *reinterpret_cast<uint32_t*>(rb + p4) = nx;
                if (CLEAN && (x & 0x80808080u)) l4 |= fix_non_ascii(rb, p4, v4, 4, lbmin);
            }
            plain = plain && __all_sync(FULL, (l4 | s4) == v4);
            const int sh = 4 * (lane & 7);
            const uint32_t L0 = __reduce_or_sync(FULL, lane < 8 ? l4 << sh : 0u);
            const uint32_t L1 = __reduce_or_sync(FULL, lane >= 8 ? l4 << sh : 0u);   // lanes >= 16 hold 0
            const uint32_t S0 = __reduce_or_sync(FULL, lane < 8 ? s4 << sh : 0u);
            const uint32_t S1 = __reduce_or_sync(FULL, lane >= 8 ? s4 << sh : 0u);
            if (lane < 2) S.lmask[32 + lane] = lane ? L1 : L0;
            // separators of the extension, appended to the list: lane l looks at bits l of both words
            const int n0 = __popc(S0);
            if ((S0 >> lane) & 1u) S.tok[1 + win_sp + __popc(S0 & low_mask(lane))] = (uint16_t)(ROW_CH + lane + 1);
            if ((S1 >> lane) & 1u) S.tok[1 + win_sp + n0 + __popc(S1 & low_mask(lane))] = (uint16_t)(ROW_CH + 32 + lane + 1);
            win_sp += n0 + __popc(S1);
            if (L1) last_letter = ROW_CH + 32 + 31 - __clz(L1);
            else if (L0) last_letter = ROW_CH + 31 - __clz(L0);
            ext_dirty = true;
        } else if (ext_dirty) {
            if (lane < 2) S.lmask[32 + lane] = 0u;
            ext_dirty = false;
        }
        const int T = win_sp + 1;
        if (lane == 0) {
            S.tok[0] = (uint16_t)lo;
            S.tok[T] = (uint16_t)(hi + 1);
        }
        __syncwarp();

        // ================= choose + prefetch the next window =================
        // tokens complete in this window: all of them when it is the dialogue's last window, otherwise all but the trailing
        // one, which restarts the next window.  A window without any separator holds a token of more than a KiB: such a
        // dialogue goes to the wide-table kernel, which carries the token as hash state.
        const bool giant = !last && T == 1;
        const int Tc = last ? T : T - 1;
        long long n_item = item, n_row = row, n_beg = row_beg, n_end = row_end, n_wbase;
        int n_lo, n_hi, n_lbmin;
        bool n_last, have_next = true;
        if (!last) {
            const long long cursor = giant ? wbase + hi : wbase + (long long)S.tok[T - 1];
            n_wbase = (cursor & ~15LL) - (giant ? 16 : 0);
            n_lo = (int)(cursor - n_wbase);
            n_lbmin = 0;
        } else {
            if (TL == 9 && a.dyn_rows) {  // (the 512-slot shape is the one micro-batches run in; the large-batch shape stays static)
                unsigned int nx = 0;
                if (lane == 0) nx = atomicAdd(&a.ctrl->next_item, 1u);
                n_item = stride + (long long)__shfl_sync(FULL, nx, 0);  // (the first `stride` dialogues were handed out by warp index)
            } else {
                n_item = item + stride;
            }
            have_next = n_item < n_items;
            if (have_next) {
                n_row = n_item;
                n_beg = a.offs[n_row];
                n_end = a.offs[n_row + 1];
            }
            n_wbase = n_beg & ~15LL;
            n_lo = (int)(n_beg - n_wbase);
            n_lbmin = n_lo;
        }
        n_last = (n_end - n_wbase) <= WIN;
        n_hi = n_last ? (int)(n_end - n_wbase) : ROW_CH;
        const bool n_loaded = have_next && n_hi > 0;
        if (n_loaded && lane == 0) {
            const uint32_t bytes = (uint32_t)((n_hi + 15) & ~15);
            fence_proxy_async();
            mbar_expect_tx(&S.mbar[buf ^ 1], bytes);
            tma_load_1d(S.raw[buf ^ 1], text + n_wbase, bytes, &S.mbar[buf ^ 1]);
        }

        // ================= pass B: one token per lane =================
        row_sp += (uint32_t)win_sp;
        if (giant) big = true;
        if (last_letter >= 0) {
            // a letter exists in this window: empties carried from earlier windows were interior
            if (pend && !big && !m.drop_empty) {
                if (DTS) {
                    // (the empty token's bucket is a split feature only by accident; it is never a stop word)
                    const uint32_t fi = m.empty_idx, slot = (fi * m.dt_mul) >> m.dt_shift;
                    if (lane == 0 && ((dtk[slot] ^ fi) & 0x7FFFFFFFu) == 0u) atomicAdd(&cnt[slot], pend);
                } else {
                    bool isnew = false;
                    const uint32_t fi = m.empty_idx;
                    const bool seen = MODE == MODE_CSR || !m.filt_on || ((__ldg(&m.filt[fi >> 5].x) >> (fi & 31)) & 1u);
                    if (csr_list) {
                        if (D + pend > (uint32_t)list_cap) {
                            big = true;
                        } else {
                            for (uint32_t i = (uint32_t)lane; i < pend; i += 32u) sts_u32(cntA + 4u * (D + i), fi);
                            D += pend;
                        }
                    } else {
                        if (lane == 0 && seen) isnew = tab_insert(tab, TLOG2, fi, pend, binary);
                        D += __popc(__ballot_sync(FULL, isnew));
                    }
                }
            }
            pend = 0;
            row_letter = true;
        }
        // The loop body exists twice: for plain windows (token = the bytes between two separators) and for windows with
        // dropped bytes (kept-byte analysis per token, byte-wise walk for tokens whose kept bytes are not one run).
        auto pass_b = [&](auto plain_tag) {
            constexpr bool PLAIN = decltype(plain_tag)::value;
            for (int t0 = 0; t0 < Tc; t0 += 32) {
                if (csr_list) {
                    if (D + 32u > (uint32_t)list_cap) {  // the list is full: the wide-table kernel takes the dialogue
                        big = true;
                        break;
                    }
                } else if (!DTS && D + 32u > (3u << TLOG2) / 4u) {  // keep the table at most 3/4 full
                    const long long gw = (long long)blockIdx.x * ROW_WARPS + warp;
                    if (TL == 9 && tab == stab && a.spill_tabs != nullptr && gw < SPILL_WARPS) {
                        // carry on in this warp's own wide table: clear it, move the entries over, switch
                        unsigned long long* wt = a.spill_tabs + ((size_t)gw << MED_LOG2);
                        const uint4 ff = make_uint4(0u, ~0u, 0u, ~0u);
                        for (int i = lane; i < (1 << MED_LOG2) / 2; i += 32) reinterpret_cast<uint4*>(wt)[i] = ff;
                        __syncwarp();
                        for (int i = lane; i < (1 << TL); i += 32) {
                            const unsigned long long e = stab[i];
                            if ((uint32_t)(e >> 32) != EMPTY_KEY) tab_insert(wt, MED_LOG2, (uint32_t)(e >> 32), (uint32_t)e, false);
                        }
                        __syncwarp();
                        tab = wt;
                        TLOG2 = MED_LOG2;
                    } else {
                        big = true;
                        break;
                    }
                }
                // ---- token extent and its kept-byte pattern
                const int ti = t0 + lane;
                const bool valid = ti < Tc;
                const uint32_t ta = tokA + 2u * (uint32_t)(valid ? ti : 0);
                int s = (int)lds_u16(ta);
                const int e = (int)lds_u16(ta + 2u) - 1;
                const int len = e - s;
                uint32_t n;      // kept bytes
                int p0, nn;      // where the straight-line hash reads, and how many bytes (0: nothing)
                bool empty, slow = false;
                if (PLAIN) {     // every byte between two separators is kept
                    n = (uint32_t)len;
                    p0 = s;
                    nn = len;
                    empty = len == 0;
                } else {
                    const uint32_t bits = __funnelshift_r(S.lmask[s >> 5], S.lmask[(s >> 5) + 1], s & 31) & low_mask(min(len, 32));
                    const int f = __ffs(bits) - 1;  // -1 when no kept byte
                    n = (uint32_t)__popc(bits);
                    const bool fast = len <= 32 && bits != 0u && (bits >> (f & 31)) == low_mask((int)n);  // one run of kept bytes
                    empty = len <= 32 && bits == 0u;
                    slow = valid && !fast && !empty;
                    s = fast ? s + f : s;  // fast tokens: s now points at the first kept byte
                    p0 = fast ? s : 0;
                    nn = fast ? (int)n : 0;
                }
                // ---- murmur3 of a contiguous token, first 12 bytes straight-line
                uint32_t b0, b1, b2, h;
                bool any_empty = false;  // (warp-uniform; plain windows only)
                {
                    const uint32_t pa = rbA + (uint32_t)p0;
                    const uint32_t wa = pa & ~3u;
                    const uint32_t sh8 = pa << 3;  // (the funnel shift takes the amount modulo 32; the buffer is 16-byte aligned)
                    const uint32_t w0 = lds_u32(wa), w1 = lds_u32(wa + 4u), w2 = lds_u32(wa + 8u), w3 = lds_u32(wa + 12u);
                    const uint4 mk = lds_v4(lutA + 16u * (uint32_t)min(nn, 12));
                    b0 = __funnelshift_r(w0, w1, sh8) & mk.x;
                    b1 = __funnelshift_r(w1, w2, sh8) & mk.y;
                    b2 = __funnelshift_r(w2, w3, sh8) & mk.z;
                    // a block past the token is all zero and mmk(0) = 0, so "h ^= mmk(block)" covers both the tail block
                    // and the blocks beyond it: one select per block (full block -> rotate/multiply step)
                    uint32_t t = 42u ^ mmk(b0);
                    h = nn >= 4 ? rotl_c<13>(t) * 5u + 0xe6546b64u : t;
                    t = h ^ mmk(b1);
                    h = nn >= 8 ? rotl_c<13>(t) * 5u + 0xe6546b64u : t;
                    t = h ^ mmk(b2);
                    h = nn >= 12 ? rotl_c<13>(t) * 5u + 0xe6546b64u : t;
                    // one vote covers both rare cases of a plain window: a token longer than 12 bytes, an empty token
                    if (__any_sync(FULL, nn > 12 || (PLAIN && valid && empty))) {
                        if (nn > 12) {
                            uint32_t prev = w3;
                            for (int i = 3; nn > 4 * i; ++i) {
                                const uint32_t nxt = lds_u32(wa + 4u * (uint32_t)(i + 1));
                                const uint32_t k = __funnelshift_r(prev, nxt, sh8);
                                prev = nxt;
                                const int r = nn - 4 * i;
                                h = r >= 4 ? mm_h(h, mm_k(k)) : h ^ mm_k(k & low_mask(8 * r));
                            }
                        }
                        if (PLAIN) any_empty = __any_sync(FULL, valid && empty);
                    }
                    h = mm_fmix(h, (uint32_t)nn);
                }
                if (!PLAIN && __any_sync(FULL, slow)) {
                    if (slow) {  // kept bytes are not contiguous (or the token is longer than 32 bytes): byte-wise walk
                        HashState hs;
                        hs.reset();
                        for (int p = s; p < e; ++p)
                            if (letter_at(S.lmask, p)) hs.put(rb[p]);
                        h = hs.finish();
                        n = hs.n;
                        b0 = hs.b0;
                        b1 = hs.b1;
                        b2 = hs.b2;
                        empty = n == 0u;
                    }
                }
                // ---- empty tokens count when a kept byte follows somewhere in the dialogue (Java's split drops trailing ones)
                bool ins = valid;
                if (PLAIN ? any_empty : __any_sync(FULL, valid && empty)) {
                    const bool interior = empty && s <= last_letter;
                    ins = valid && (empty ? (interior && !m.drop_empty) : true);
                    pend += __popc(__ballot_sync(FULL, valid && empty && !interior));
                }
                // ---- term index
                uint32_t fi;
                if (POW2) {
                    fi = h & fmask;  // (the empty token hashes to murmur3("") by itself: nn = 0)
                } else if (!m.vocab_cap) {
                    fi = empty ? m.empty_idx : feature_index(h, m);
                } else {
                    fi = empty ? m.empty_idx : vocab_index(m, h, n, b0, b1, b2, rb, S.lmask, s, e);
                    ins = ins && fi != NOT_IN_VOCAB;
                    fi = ins ? fi : 0u;
                }
                if (DTS) {
                    // small decision tree: bucket -> counter slot -> one compare -> one shared-memory add; nothing is ever "new"
                    // (a binary HashingTF counts too: the end of the dialogue clamps the counts to 1)
                    const uint32_t slot4 = ((fi * m.dt_mul) >> m.dt_shift) * 4u;
                    const uint32_t key = lds_u32(dtkA + slot4);
                    bool hit = ins && ((key ^ fi) & 0x7FFFFFFFu) == 0u;
                    if (DTS == 2 && __any_sync(FULL, hit && (int)key < 0)) {  // a stop word maps to this split feature: is the token one?
                        if (hit && (int)key < 0 && !empty) hit = !stop_probe(m, m.sw_tab, m.sw_words, h, n, b0, b1, b2, rbA, lmA, s, e, true);
                    }
                    if (hit) red_add_u32(cntA + slot4, 1u);
                } else {
                    if (MODE != MODE_CSR && m.filt_on) {
                        const uint2 fw = __ldg(&m.filt[fi >> 5]);
                        ins = ins && ((fw.x >> (fi & 31)) & 1u);           // can the classifier see this feature at all?
                        const bool chk = ins && !empty && ((fw.y >> (fi & 31)) & 1u);
                        if (__any_sync(FULL, chk)) {                       // ... and a stop word maps to it: is the token one?
                            if (chk) ins = !stop_probe_s(m, swtA, swwA, h, n, b0, b1, b2, rbA, lmA, s, e, true);
                        }
                    } else if (m.sw_cap) {
                        ins = ins && !stop_probe_s(m, swtA, swwA, h, n, b0, b1, b2, rbA, lmA, s, e, valid && !empty);
                    }
                    bool isnew = false;
                    if (csr_list) {  // append the bucket; the end of the dialogue orders and counts the list
                        const uint32_t bal = __ballot_sync(FULL, ins);
                        if (ins) sts_u32(cntA + 4u * (D + (uint32_t)__popc(bal & ((1u << lane) - 1u))), fi);
                        D += (uint32_t)__popc(bal);
                        continue;
                    }
                    if (TL == 8 || tab == stab) {  // (named so that the common case keeps shared-memory atomics)
                        if (ins) isnew = tab_insert(stab, TL, fi, 1u, binary);
                    } else {
                        if (ins) isnew = tab_insert(tab, TLOG2, fi, 1u, binary);
                    }
                    D += __popc(__ballot_sync(FULL, isnew));
                }
            }
        };
        if (!big) {
            if (plain) pass_b(std::true_type{});
            else pass_b(std::false_type{});
        }
        __syncwarp();

        // ================= end of dialogue =================
        if (last) {
            if (!big && row_sp == 0 && !row_letter && !m.drop_empty) {
                // no separator anywhere and nothing kept: Java's split returns [""] -- one empty token
                const uint32_t fi = m.empty_idx;
                if (DTS) {
                    const uint32_t slot = (fi * m.dt_mul) >> m.dt_shift;
                    if (lane == 0 && ((dtk[slot] ^ fi) & 0x7FFFFFFFu) == 0u) cnt[slot] = 1u;
                } else {
                    bool isnew = false;
                    const bool seen = MODE == MODE_CSR || !m.filt_on || ((__ldg(&m.filt[fi >> 5].x) >> (fi & 31)) & 1u);
                    if (csr_list) {
                        if (lane == 0) sts_u32(cntA + 4u * D, fi);  // (D == 0 here: nothing was kept)
                        D += 1u;
                    } else {
                        if (lane == 0 && seen) isnew = tab_insert(tab, TLOG2, fi, 1u, binary);
                        D += __popc(__ballot_sync(FULL, isnew));
                    }
                }
            }
            __syncwarp();
            if (big) {
                if (lane == 0) {
                    // size class by the table the dialogue can need at most
                    if (wide_table_log2(row_end - row_beg) <= MED_LOG2) {
                        const unsigned int pos = atomicAdd(&a.ctrl->big_count, 1u);
                        a.big_rows[pos] = row;
                    } else {
                        const unsigned int pos = atomicAdd(&a.ctrl->big2_count, 1u);
                        a.big_rows[a.big_cap - 1 - pos] = row;
                    }
                }
                if (DTS) {
                    __syncwarp();
                    for (int i = lane; i < (1 << m.dt_log2); i += 32) cnt[i] = 0u;
                }
            } else if (DTS) {
                // every split node at once: lane i holds heap node i; bit i of the ballot = "go right at node i"
                const uint32_t ca = cntA + 4u * (uint32_t)hn.x;
                uint32_t c = lds_u32(ca);
                if (binary) c = min(c, 1u);
                const uint32_t bits = __ballot_sync(FULL, (int)c > hn.y);
                int p = 0;
#pragma unroll
                for (int d = 0; d < 5; ++d) p = 2 * p + 1 + (int)((bits >> p) & 1u);
                if (lane < 5) outp[(long long)row * ostride] = dto[(p - 31) * 5 + lane];
                sts_u32(ca, 0u);  // (only split-feature counters are ever touched; each is some lane's)
            } else if (csr_list) {
                csr_sorted_emit<POW2>(a, cntA, (int)D, rbA, tokA, row, lane, bin_shift);
            } else if (TL == 8 || tab == stab) {
                row_epilogue<MODE>(a, reinterpret_cast<double*>(S.tok), reinterpret_cast<double*>(rb), true, stab, TL, row, (int)D, lane);
            } else {
                row_epilogue<MODE>(a, reinterpret_cast<double*>(S.tok), reinterpret_cast<double*>(rb), false, tab, TLOG2, row, (int)D, lane);
            }
            // reset the table and the per-dialogue state
            __syncwarp();
            if (!DTS) {
                if (TL == 9 && tab != stab) {  // back to the shared-memory table for the next dialogue
                    tab = stab;
                    TLOG2 = TL;
                }
                if (!csr_list) {
                    const uint4 ff = make_uint4(0u, ~0u, 0u, ~0u);  // count = 0, key = EMPTY
                    for (int i = lane; i < (1 << TL) / 2; i += 32) reinterpret_cast<uint4*>(stab)[i] = ff;
                }
            }
            D = 0;
            pend = 0;
            row_sp = 0;
            row_letter = false;
            big = false;
            __syncwarp();
        }
        if (!have_next) break;
        item = n_item;
        row = n_row;
        row_beg = n_beg;
        row_end = n_end;
        wbase = n_wbase;
        lo = n_lo;
        hi = n_hi;
        last = n_last;
        lbmin = n_lbmin;
        cur_loaded = n_loaded;
        buf ^= 1;
    }
}

// ------------------------------------------------------------------------------------------------
// launcher
// ------------------------------------------------------------------------------------------------
template <int MODE, bool CLEAN, int TL, int VAR, bool POW2>
static cudaError_t launch_k(fd_handle* h, const RowArgs& a, cudaStream_t st) {
    constexpr int DTS = (VAR == 1 || VAR == 2) ? VAR : 0;
    constexpr int NW = row_warps(TL);
    const size_t cta_bytes = LUT_BYTES + (DTS ? (((size_t)4 << a.m.dt_log2) + 160 * 8) : ((a.m.sw_smem_bytes() + 15) & ~(size_t)15));
    const size_t warp_bytes = sizeof(MainWarp) + (DTS ? ((size_t)4 << a.m.dt_log2) : ((size_t)8 << TL));
    const size_t smem = cta_bytes + (size_t)NW * warp_bytes;
    auto kern = k_main<MODE, CLEAN, TL, VAR, POW2>;
    // shared-memory opt-in and occupancy are looked up once per (kernel, size) and kept in the handle
    int occ = 0;
    for (const auto& c : h->launch_cfg)
        if (c.fn == (const void*)kern && c.smem == smem) occ = c.occ;
    if (!occ) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 32 * NW, smem);
        if (e != cudaSuccess) return e;
        if (occ < 1) occ = 1;
        h->launch_cfg.push_back({(const void*)kern, smem, occ});
    }
    long long grid = std::min<long long>((long long)h->sm_count * occ, (a.n_rows + NW - 1) / NW);
    if (grid < 1) grid = 1;
    kern<<<(unsigned)grid, 32 * NW, smem, st>>>(a, (int)cta_bytes, (int)warp_bytes);
    h->launches += 1;
    return cudaGetLastError();
}

// count-table kernel, or featurize's token-list variant when the model allows it (Model::csr_list)
template <int MODE, bool CLEAN, int TL, bool POW2>
static cudaError_t launch_generic(fd_handle* h, const RowArgs& a, cudaStream_t st) {
    if (MODE == MODE_CSR && a.m.csr_list) return launch_k<MODE, CLEAN, TL, MODE == MODE_CSR ? 3 : 0, POW2>(h, a, st);
    return launch_k<MODE, CLEAN, TL, 0, POW2>(h, a, st);
}

template <int MODE, bool CLEAN, bool POW2>
static cudaError_t launch_shapes(fd_handle* h, const RowArgs& a0, cudaStream_t st) {
    RowArgs a = a0;
    if (MODE == MODE_DT && a.m.dt_small && a.m.dt_heap_on && !a.m.vocab_cap) {
        a.pick_shape = 0;  // counters, not a table: one shape (16 warps per CTA) whatever the batch
        return a.m.dt_stop ? launch_k<MODE_DT, CLEAN, 8, 2, POW2>(h, a, st) : launch_k<MODE_DT, CLEAN, 8, 1, POW2>(h, a, st);
    }
    a.m.dt_small = 0;  // (the generic path below keeps a full count table; the wide-table kernels see the caller's model)
    a.pick_shape = 0;
    if (a.n_rows >= ROWS_FOR_SMALL_TABLES) {
        // large batches take the 256-slot shape when their dialogues are short.  A caller that holds the offsets on the host
        // says how many bytes the batch has and the shape is picked here; otherwise both shapes are launched and each CTA
        // finds out from the offsets whether its shape is the one that runs (the other exits at once)
        const float keep = (MODE == MODE_LINEAR && a.m.filt_on) ? a.m.lr_keep : 1.0f;
        if (a.host_bytes >= 0) {
            if (shape_for(a.host_bytes, a.n_rows, keep) == 8) return launch_generic<MODE, CLEAN, 8, POW2>(h, a, st);
        } else {
            a.pick_shape = 1;
            a.pick_keep = keep;
            const cudaError_t e = launch_generic<MODE, CLEAN, 8, POW2>(h, a, st);
            if (e != cudaSuccess) return e;
        }
    }
    return launch_generic<MODE, CLEAN, 9, POW2>(h, a, st);
}

// micro-batches take their dialogues dynamically: one atomic per dialogue on one address is nothing for a few thousand
// dialogues and would serialise a batch of a million
constexpr long long DYN_ROWS_MAX = 1 << 16;

cudaError_t launch_main_rows(fd_handle* h, int mode, const RowArgs& a0, cudaStream_t st) {
    RowArgs a = a0;
    a.dyn_rows = a.n_rows <= DYN_ROWS_MAX ? 1 : 0;
    const bool c = (a.flags & FD_FLAG_CLEAN) != 0;
    const bool p2 = a.m.mod_pow2 && !a.m.vocab_cap;
#define FD_DISPATCH(M)                                                                                                 \
    return c ? (p2 ? launch_shapes<M, true, true>(h, a, st) : launch_shapes<M, true, false>(h, a, st))                 \
             : (p2 ? launch_shapes<M, false, true>(h, a, st) : launch_shapes<M, false, false>(h, a, st))
    switch (mode) {
        case MODE_CSR: FD_DISPATCH(MODE_CSR);
        case MODE_LINEAR: FD_DISPATCH(MODE_LINEAR);
        case MODE_DT: FD_DISPATCH(MODE_DT);
        case MODE_RF: FD_DISPATCH(MODE_RF);
        case MODE_GBT: FD_DISPATCH(MODE_GBT);
    }
#undef FD_DISPATCH
    return cudaErrorInvalidValue;
}

}  // namespace fd
