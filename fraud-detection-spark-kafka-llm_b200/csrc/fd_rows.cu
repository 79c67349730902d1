// fd_rows.cu -- the dialogue kernels: clean -> tokenise -> stop-word filter -> murmur3 HashingTF ->
// per-dialogue term counts, followed by either CSR emission or a fused classifier epilogue.
//
// One WARP owns one dialogue at a time.  Its UTF-8 bytes are staged from HBM into the warp's shared
// memory windows (1 KiB, double-buffered) with TMA bulk copies (cp.async.bulk + mbarrier), so the
// next window -- of this dialogue or of the warp's next dialogue -- is in flight while the current
// one is hashed.  Per window:
//   pass A  each lane classifies 32 bytes with SWAR arithmetic (letter / separator / dropped),
//           lower-cases in place, and publishes a letter bitmask, a separator bitmask and the list
//           of token start offsets (a dialogue's last window may run 64 bytes longer: one word per lane);
//   pass B  one token per lane: MurmurHash3_x86_32(seed 42) straight out of shared memory,
//           stop-word probe (perfect-hash table in shared memory keyed by the same hash, record compared
//           byte for byte), nonNegativeMod, and an atomicCAS insert into the warp's open-addressed
//           (bucket -> count) table.
// At the end of a dialogue the table is either compacted + bitonic-sorted and appended to the CSR
// staging buffer (featurize), dotted with the logistic coefficients in ascending-bucket order, or
// probed by a lane-per-tree walk over the forest's two-level 32-byte records (DT / RF / GBT; forests first
// build a loop-free lookup front over the table, a small decision tree counts only its own split features).
//
// Semantics restated from Spark 3.5.5; reference call sites:
// utils/agent_api.py:143-144 (cleaning), fraud_detection_spark.py:50-53 (Tokenizer, StopWordsRemover,
// HashingTF/IDF stages), utils/agent_api.py:158-167 (transform + result columns).
#include <algorithm>

#include "fd_rows.cuh"

namespace fd {

// ------------------------------------------------------------------------------------------------
// PTX helpers: mbarrier + TMA bulk copy (global -> shared), sm_90+/sm_100a
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, unsigned long long* b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(b))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* b, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
            : "=r"(ok)
            : "r"(smem_u32(b)), "r"(parity)
            : "memory");
    } while (!ok);
}

// n low bits set (n clamped to [0, 32]) -- one BMSK instruction
__device__ __forceinline__ uint32_t low_mask(int n) {
    uint32_t r;
    asm("bmsk.clamp.b32 %0, 0, %1;" : "=r"(r) : "r"(n));
    return r;
}
// gather the four 0x80 flag bits of a SWAR word into a nibble (byte 0 -> bit 0)
__device__ __forceinline__ uint32_t movemask4(uint32_t f80) { return ((f80 >> 7) * 0x01020408u) >> 24; }

// nonNegativeMod(int32(h), nf)
__device__ __forceinline__ uint32_t feature_index(uint32_t h, const Model& m) {
    if (m.mod_pow2) return h & (uint32_t)(m.num_features - 1);
    const uint32_t nf = (uint32_t)m.num_features;
    uint32_t q = __umulhi(h, m.mod_magic);
    uint32_t r = h - q * nf;  // h mod nf, possibly + a few nf
    while (r >= nf) r -= nf;
    if ((int32_t)h < 0) r = (r >= m.mod_r32) ? r - m.mod_r32 : r + nf - m.mod_r32;  // value is h - 2^32
    return r;
}

constexpr uint32_t NOT_IN_VOCAB = 0xFFFFFFFFu;

// token bytes [s, e) filtered by the letter mask == word (same length already checked)
__device__ __noinline__ bool long_word_equal(const uint8_t* word, const uint8_t* rb, const uint32_t* lmask, int s, int e) {
    uint32_t i = 0;
    for (int p = s; p < e; ++p)
        if (((lmask[p >> 5] >> (p & 31)) & 1u) && rb[p] != word[i++]) return false;
    return true;
}

// CountVectorizerModel: term -> vocabulary index (NOT_IN_VOCAB when absent).  h = murmur3(term), n = its length,
// b0..b2 = its first 12 bytes; [s, e) locates it in the window for the byte compare of longer words (s < 0: a
// multi-window token, never matched against words longer than 12 bytes -- documented limitation).
__device__ uint32_t vocab_index(const Model& m, uint32_t h, uint32_t n, uint32_t b0, uint32_t b1, uint32_t b2, const uint8_t* rb,
                                const uint32_t* lmask, int s, int e) {
    const uint32_t mask = (uint32_t)m.vocab_cap - 1u;
    const uint32_t slots[2] = {h & mask, sw_key(n, b0, b1, b2) & mask};
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const uint2 ent = __ldg(m.vocab_tab + slots[i]);
        if (ent.y != 0u && ent.x == h) {
            const uint4 w = __ldg(m.vocab_words + (ent.y - 1u));
            if ((w.x & 0x7FFFFFFFu) == n) {
                if (!(w.x & 0x80000000u)) {
                    if (w.y == b0 && w.z == b1 && w.w == b2) return ent.y - 1u;
                } else if (s >= 0 && long_word_equal(m.vocab_pool + w.y, rb, lmask, s, e)) {
                    return ent.y - 1u;
                }
            }
        }
    }
    return NOT_IN_VOCAB;
}

// CountVectorizer.fit: find-or-create the term in the caller's global table and add `inc` to its term count; returns
// the slot (the term's identity for the per-document count table).  Terms of up to 7 bytes are identified by their
// bytes; longer ones by a 56-bit hash (murmur3 x cheap mix) -- a collision there would merge two terms' counts.
__device__ uint32_t fit_slot(const RowArgs& a, uint32_t h, uint32_t n, uint32_t b0, uint32_t b1, uint32_t b2, const uint8_t* rb,
                             const uint32_t* lmask, int s, int e, uint32_t inc) {
    unsigned long long key;
    if (n <= 7u) key = ((unsigned long long)n << 56) | ((unsigned long long)(b1 & 0x00FFFFFFu) << 32) | b0;
    else key = 0xFF00000000000000ULL | ((((unsigned long long)h << 24) ^ sw_key(n, b0, b1, b2)) & 0x00FFFFFFFFFFFFFFULL);
    if (key == ~0ULL) key ^= 1ULL;
    const uint32_t mask = (1u << a.fit_log2) - 1u;
    uint32_t slot = (uint32_t)((key * 0x9E3779B97F4A7C15ULL) >> (64 - a.fit_log2));
    for (uint32_t probes = 0; probes <= mask; ++probes) {
        const unsigned long long old = atomicCAS(a.fit_keys + slot, ~0ULL, key);
        if (old == ~0ULL) {  // this thread created the entry: record its bytes for the host
            atomicAdd(a.fit_counters, 1ULL);
            uint32_t off = 0xFFFFFFFFu;
            if (n > 7u && s >= 0) {
                const unsigned long long o = atomicAdd(a.fit_counters + 1, (unsigned long long)n);
                if ((long long)(o + n) <= a.fit_pool_cap) {
                    off = (uint32_t)o;
                    uint32_t i = 0;
                    for (int p = s; p < e; ++p)
                        if ((lmask[p >> 5] >> (p & 31)) & 1u) a.fit_pool[o + i++] = rb[p];
                } else {
                    a.fit_counters[2] = 1ULL;
                }
            }
            a.fit_info[slot].z = off;
            a.fit_info[slot].w = n;
        }
        if (old == ~0ULL || old == key) {
            atomicAdd(&a.fit_info[slot].x, inc);
            return slot;
        }
        slot = (slot + 1u) & mask;
    }
    a.fit_counters[2] = 1ULL;  // table full
    return 0u;
}

// term -> index of the per-document count table, by mode: hashing trick, vocabulary lookup, or (fit) global term slot
template <int MODE>
__device__ __forceinline__ uint32_t term_index(const RowArgs& a, uint32_t h, uint32_t n, uint32_t b0, uint32_t b1, uint32_t b2,
                                               const uint8_t* rb, const uint32_t* lmask, int s, int e) {
    if (MODE == MODE_FIT) return fit_slot(a, h, n, b0, b1, b2, rb, lmask, s, e, 1u);
    return a.m.vocab_cap ? vocab_index(a.m, h, n, b0, b1, b2, rb, lmask, s, e) : feature_index(h, a.m);
}
template <int MODE>
__device__ __forceinline__ uint32_t empty_index(const RowArgs& a, uint32_t inc) {
    if (MODE == MODE_FIT) return fit_slot(a, 142593372u, 0u, 0u, 0u, 0u, nullptr, nullptr, -1, -1, inc);  // murmur3("") = 142593372
    return a.m.empty_idx;
}
template <int MODE>
__device__ __forceinline__ bool count_term(const RowArgs& a, unsigned long long* tab, int log2cap, uint32_t fi, uint32_t inc, bool binary);

__device__ __forceinline__ uint32_t tab_slot(uint32_t idx, int log2cap) { return (idx * 0x9E3779B1u) >> (32 - log2cap); }

// (bucket -> count) insert; returns true when the bucket is new for this dialogue
__device__ __forceinline__ bool tab_insert(unsigned long long* tab, int log2cap, uint32_t idx, uint32_t inc, bool binary) {
    const uint32_t mask = (1u << log2cap) - 1u;
    uint32_t slot = tab_slot(idx, log2cap);
    for (;;) {
        unsigned int* cell = reinterpret_cast<unsigned int*>(tab + slot);  // [0] = count, [1] = key
        unsigned int old = atomicCAS(cell + 1, EMPTY_KEY, idx);
        if (old == EMPTY_KEY || old == idx) {
            if (binary)
                cell[0] = 1u;
            else
                atomicAdd(cell, inc);
            return old == EMPTY_KEY;
        }
        slot = (slot + 1u) & mask;
    }
}

template <int MODE>
__device__ __forceinline__ bool count_term(const RowArgs& a, unsigned long long* tab, int log2cap, uint32_t fi, uint32_t inc, bool binary) {
    if (MODE == MODE_DT && a.m.dt_small) {
        // only the tree's own split features are counted (Model::dt_small): one compare drops every other term, and
        // the dialogue never outgrows its table (nothing is "new", so nothing is ever deferred)
        const uint32_t slot = (fi * a.m.dt_mul) >> a.m.dt_shift;
        if (__ldg(a.m.dt_tab + slot) == fi) {
            unsigned int* c = reinterpret_cast<unsigned int*>(tab) + slot;
            if (binary)
                *c = 1u;
            else
                atomicAdd(c, inc);
        }
        return false;
    }
    if (MODE == MODE_LINEAR && a.m.lr_sparse && !((__ldg(a.m.lr_nz + (fi >> 5)) >> (fi & 31)) & 1u)) return false;  // zero coefficient
    // (not in the 256-slot shape: large batches of short dialogues rarely outgrow their table, and there the bitmap probe
    // was measured to cost more than the inserts it saves; it pays where dialogues are long -- micro-batches, wide tables)
    if ((MODE == MODE_DT || MODE == MODE_RF || MODE == MODE_GBT) && a.m.used_on && log2cap != 8 &&
        !((__ldg(a.m.used_nz + (fi >> 5)) >> (fi & 31)) & 1u))
        return false;  // no node of the forest splits on this feature
    const bool isnew = tab_insert(tab, log2cap, fi, inc, binary);
    if (MODE == MODE_FIT && isnew) atomicAdd(&a.fit_info[fi].y, 1u);  // first occurrence in this document
    return isnew;
}

__device__ __forceinline__ uint32_t tab_lookup(const unsigned long long* tab, int log2cap, uint32_t idx) {
    const uint32_t mask = (1u << log2cap) - 1u;
    uint32_t slot = tab_slot(idx, log2cap);
    for (;;) {
        unsigned long long e = tab[slot];
        uint32_t k = (uint32_t)(e >> 32);
        if (k == idx) return (uint32_t)e;
        if (k == EMPTY_KEY) return 0u;
        slot = (slot + 1u) & mask;
    }
}

// Forest walks probe the count table ~20 times per tree, mostly for features the dialogue does not have, and a
// linear-probe loop costs the warp its slowest lane every time.  For forests the epilogue therefore builds a
// loop-free front: fidx[g(f)] = 1 + table slot of THE present feature that maps to g, 0 when none does (the
// common case: the feature is absent, no table access at all), FIDX_MANY when several do (probe the table).
constexpr int FIDX_LOG2 = 10;               // 1,024 16-bit entries = 2 KiB (the token-list area, free in the epilogue)
constexpr uint32_t FIDX_MANY = 0xFFFFu;
__device__ __forceinline__ uint32_t fidx_slot(uint32_t f) { return (f * 0x85EBCA6Bu) >> (32 - FIDX_LOG2); }

__device__ __forceinline__ void build_fidx(const unsigned long long* tab, int log2cap, uint16_t* fidx, int lane) {
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (int i = lane; i < (2 << FIDX_LOG2) / 16; i += 32) reinterpret_cast<uint4*>(fidx)[i] = z;
    __syncwarp();
    const int cap = 1 << log2cap;
    for (int i = lane; i < cap; i += 32) {
        const uint32_t k = (uint32_t)(tab[i] >> 32);
        if (k != EMPTY_KEY) fidx[fidx_slot(k)] = (uint16_t)(i + 1);
    }
    __syncwarp();
    for (int i = lane; i < cap; i += 32) {  // whoever lost a slot marks it shared
        const uint32_t k = (uint32_t)(tab[i] >> 32);
        if (k != EMPTY_KEY && fidx[fidx_slot(k)] != (uint16_t)(i + 1)) fidx[fidx_slot(k)] = (uint16_t)FIDX_MANY;
    }
    __syncwarp();
}

__device__ __forceinline__ uint32_t tab_lookup(const unsigned long long* tab, int log2cap, const uint16_t* fidx, uint32_t idx) {
    if (fidx == nullptr) return tab_lookup(tab, log2cap, idx);
    const uint32_t j = fidx[fidx_slot(idx)];
    if (j == 0u) return 0u;
    if (j == FIDX_MANY) return tab_lookup(tab, log2cap, idx);
    const unsigned long long e = tab[j - 1u];
    return (uint32_t)(e >> 32) == idx ? (uint32_t)e : 0u;
}

// resumable murmur3 state for tokens whose kept bytes are not contiguous (or span windows)
struct HashState {
    uint32_t h, k, n, b0, b1, b2;
    __device__ __forceinline__ void reset() {
        h = 42u;
        k = 0u;
        n = 0u;
        b0 = b1 = b2 = 0u;
    }
    __device__ __forceinline__ void put(uint32_t c) {
        k |= c << (8u * (n & 3u));
        if ((n & 3u) == 3u) {
            if (n < 4u) b0 = k;
            else if (n < 8u) b1 = k;
            else if (n < 12u) b2 = k;
            h = mm_h(h, mm_k(k));
            k = 0u;
        }
        ++n;
    }
    __device__ __forceinline__ uint32_t finish() {
        uint32_t hh = h;
        if (n & 3u) {
            if (n < 4u) b0 = k;
            else if (n < 8u) b1 = k;
            else if (n < 12u) b2 = k;
            hh ^= mm_k(k);
        }
        return mm_fmix(hh, n);
    }
};

// SWAR classification of four bytes: bit 7 of each byte of l80 / s80 says "kept letter" / "separator"; nx is the word with
// its ASCII upper-case letters folded; H collects the high bits (non-ASCII bytes present).
//   CLEAN : lower() + regexp_replace("[^a-zA-Z ]", "") + split("\\s"): letters are kept, ' ' separates, the rest vanishes
//   !CLEAN: Tokenizer on already-clean text: \t \n \v \f \r ' ' separate, every other byte is kept
template <bool CLEAN>
__device__ __forceinline__ void classify_word(uint32_t x, uint32_t& l80, uint32_t& s80, uint32_t& nx, uint32_t& H) {
    if (CLEAN) {
        const uint32_t v7 = (x | 0x20202020u) & 0x7F7F7F7Fu;
        const uint32_t ge_a = v7 + 0x1F1F1F1Fu;  // bit7 <=> byte >= 0x61
        const uint32_t gt_z = v7 + 0x05050505u;  // bit7 <=> byte >= 0x7B
        l80 = ge_a & ~gt_z & ~x & 0x80808080u;
        const uint32_t z = x ^ 0x20202020u;
        s80 = ~(((z & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | z) & 0x80808080u;
        nx = x | (l80 >> 2);
        H |= x & 0x80808080u;
    } else {
        const uint32_t t = x & 0x7F7F7F7Fu;
        const uint32_t ge9 = t + 0x77777777u;   // >= 0x09
        const uint32_t geE = t + 0x72727272u;   // >= 0x0E
        const uint32_t z = x ^ 0x20202020u;
        const uint32_t sp = ~(((z & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | z) & 0x80808080u;
        s80 = ((ge9 & ~geE & ~x) | sp) & 0x80808080u;
        l80 = ~s80 & 0x80808080u;
        const uint32_t geA = t + 0x3F3F3F3Fu;   // >= 'A'
        const uint32_t gtZ = t + 0x25252525u;   // >= '['
        const uint32_t up80 = geA & ~gtZ & ~x & 0x80808080u;
        nx = x | (up80 >> 2);
    }
}

__device__ __forceinline__ bool letter_at(const uint32_t* lmask, int p) { return (lmask[p >> 5] >> (p & 31)) & 1u; }

// token bytes [s, e) filtered by the letter mask == word (same length already checked)
__device__ __noinline__ bool long_stop_word_equal(const uint8_t* word, const uint8_t* rb, const uint32_t* lmask, int s, int e) {
    uint32_t i = 0;
    for (int p = s; p < e; ++p)
        if (letter_at(lmask, p) && rb[p] != word[i++]) return false;
    return true;
}

// is the multi-window token (hash h, n kept bytes, first 12 bytes b0..b2) a stop word?  (words longer than 12
// bytes are not matched on this rare path -- documented limitation)
__device__ bool is_stop_word(const Model& m, const uint32_t* swt, const uint4* sww, uint32_t h, uint32_t n, uint32_t b0,
                             uint32_t b1, uint32_t b2) {
    if (m.sw_perfect) {
        const uint32_t id = reinterpret_cast<const uint16_t*>(swt)[(h * m.sw_mul) >> m.sw_shift];
        if (!id) return false;
        const uint4 w = sww[id - 1u];
        return w.x == n && w.y == b0 && w.z == b1 && w.w == b2;
    }
    const uint32_t mask = (uint32_t)m.sw_cap - 1u;
    const uint32_t cand[2] = {swt[h & mask], swt[(h >> 10) & mask]};
    for (int i = 0; i < 2; ++i) {
        const uint32_t e = cand[i];
        if (e != 0u && ((e ^ h) >> 16) == 0u) {
            const uint4 w = sww[(e & 0xFFFFu) - 1u];
            if (w.x == n && w.y == b0 && w.z == b1 && w.w == b2) return true;
        }
    }
    return false;
}

// wide-table size for a dialogue of `bytes` bytes: room for every possible token at <= 3/4 load, 2^9 .. 2^BIG_LOG2
__device__ __forceinline__ int wide_table_log2(long long bytes) {
    const long long need = (bytes / 2 + 2 + 32) * 2;
    const int tl = need >= (1LL << BIG_LOG2) ? BIG_LOG2 : 32 - __clz((int)need - 1);  // ceil(log2(need))
    return max(tl, 9);
}

// ------------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------------
// BIG = 0: main kernel (count table in shared memory); 1: medium wide tables (8 warps per CTA); 2: large wide tables
template <int MODE, bool CLEAN, int BIG, int TL>
__global__ void __launch_bounds__(BIG == 2 ? 32 : (BIG == 1 ? 32 * MED_WARPS : 32 * row_warps(TL)), BIG ? 1 : (TL == 8 ? 2 : 3))
k_rows(const RowArgs a) {
    constexpr int ROW_WARPS = BIG == 2 ? 1 : (BIG == 1 ? MED_WARPS : row_warps(TL));
    constexpr int TAB_LOG2 = TL;
    using WarpSmemT = WarpSmem<TL>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* swt = reinterpret_cast<uint32_t*>(smem_raw);
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const Model& m = a.m;
    uint4* sww = reinterpret_cast<uint4*>(smem_raw + m.sw_tab_bytes());
    WarpSmemT& S = *reinterpret_cast<WarpSmemT*>(smem_raw + m.sw_smem_bytes() + (size_t)warp * sizeof(WarpSmemT));

    if (BIG && (long long)blockIdx.x * ROW_WARPS >= (long long)(BIG == 1 ? a.ctrl->big_count : a.ctrl->big2_count))
        return;  // nothing deferred to this CTA
    if (!BIG && a.pick_shape && a.ctrl->shape != TL) return;  // the other shape was picked for this call
    for (int i = threadIdx.x; i < (int)(m.sw_tab_bytes() / 4); i += blockDim.x) swt[i] = m.sw_tab[i];
    for (int i = threadIdx.x; i < m.n_sw; i += blockDim.x) sww[i] = m.sw_words[i];
    if (lane == 0) {
        mbar_init(&S.mbar[0], 1);
        mbar_init(&S.mbar[1], 1);
        fence_mbar_init();
    }
    if (lane < 4) S.lmask[32 + lane] = 0u;

    unsigned long long* tab = BIG == 1 ? a.big_tabs + ((size_t)(blockIdx.x * MED_WARPS + warp) << MED_LOG2)
                            : (BIG == 2 ? a.big_tabs + ((size_t)blockIdx.x << BIG_LOG2) : S.tab);
    // table size: fixed in shared memory; in the wide-table kernel it follows the dialogue (a dialogue of b bytes
    // has at most b/2 + 1 tokens), so clearing it costs in proportion to the dialogue, not to the 2^16 maximum
    int TLOG2 = TAB_LOG2;
    const long long n_items = BIG == 1 ? (long long)a.ctrl->big_count : (BIG == 2 ? (long long)a.ctrl->big2_count : a.n_rows);
    const long long stride = (long long)gridDim.x * ROW_WARPS;
    long long item = (long long)blockIdx.x * ROW_WARPS + warp;
    // deferred lists: medium dialogues from the front of big_rows, large ones from its back
    auto deferred_row = [&](long long i) { return BIG == 1 ? a.big_rows[i] : a.big_rows[a.big_cap - 1 - i]; };
    if (BIG && item < n_items) {
        const long long r0 = deferred_row(item);
        TLOG2 = wide_table_log2(a.offs[r0 + 1] - a.offs[r0]);
    }
    const bool dts = MODE == MODE_DT && m.dt_small;  // the table is 2^dt_log2 plain counters (see count_term)
    const bool dts_fast = dts && m.mod_pow2 && !m.vocab_cap;  // ... and buckets are h & (numFeatures - 1)
    if (dts) {
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        for (int i = lane; i < (1 << m.dt_log2) / 4; i += 32) reinterpret_cast<uint4*>(tab)[i] = z;
    } else {
        uint4 ff = make_uint4(0u, ~0u, 0u, ~0u);  // count = 0, key = EMPTY
        for (int i = lane; i < (1 << TLOG2) / 2; i += 32) reinterpret_cast<uint4*>(tab)[i] = ff;
    }
    __syncthreads();
    if (item >= n_items) return;

    const uint8_t* text = a.text;
    const bool binary = m.binary != 0;

    // ---- window bookkeeping (all warp-uniform) ----
    long long row = BIG ? deferred_row(item) : item;
    long long row_beg = a.offs[row], row_end = a.offs[row + 1];
    long long wbase = row_beg & ~15LL;             // global byte offset of buffer byte 0
    int lo = (int)(row_beg - wbase);               // first valid byte of the window
    // the 64-byte extension of a dialogue's last window pays for dialogues of about one window -- the large-batch
    // shape; the other shapes (long dialogues, micro-batches, wide tables) keep plain 1 KiB windows and leaner code
    constexpr int WIN = (TL == 8 && BIG == 0) ? ROW_WIN : ROW_CH;
    bool last = (row_end - wbase) <= WIN;
    int hi = last ? (int)(row_end - wbase) : ROW_CH;
    int lbmin = lo;                                // lowest buffer offset a UTF-8 look-back may read
    int buf = 0;
    uint32_t phase0 = 0, phase1 = 0;
    if (lane == 0) {
        uint32_t bytes = (uint32_t)((hi + 15) & ~15);
        if (bytes) {
            mbar_expect_tx(&S.mbar[0], bytes);
            tma_load_1d(S.raw[0], text + wbase, bytes, &S.mbar[0]);
        }
    }
    bool cur_loaded = hi > 0;

    // ---- per-dialogue state ----
    uint32_t D = 0;        // distinct buckets so far
    uint32_t pend = 0;     // empty tokens not yet known to be interior
    uint32_t row_sp = 0;   // separators seen
    bool row_letter = false;
    bool big = false;      // deferred to the wide-table kernel (or over the limit there)
    bool ext_dirty = false;  // lmask[32..33] hold the extension of an earlier window
    bool gt_active = false;  // first token of the window continues a token from the previous window
    HashState gt;            // its state (lane 0 only)
    gt.reset();

    for (;;) {
        uint8_t* rb = S.raw[buf];
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
        uint32_t L = 0, SP = 0, vmask_all = 0;
        {
            uint4 q0 = *reinterpret_cast<const uint4*>(rb + 32 * lane);
            uint4 q1 = *reinterpret_cast<const uint4*>(rb + 32 * lane + 16);
            uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
            const int lo_l = min(max(lo - 32 * lane, 0), 32), hi_l = min(max(hi - 32 * lane, 0), 32);
            const uint32_t vmask = hi_l > lo_l ? (low_mask(hi_l) & ~low_mask(lo_l)) : 0u;
            vmask_all = vmask;
            uint32_t H = 0;
            bool changed = false;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t x = w[j];
                uint32_t l80, s80, nx;
                classify_word<CLEAN>(x, l80, s80, nx, H);
                L |= movemask4(l80) << (4 * j);
                SP |= movemask4(s80) << (4 * j);
                changed |= (nx != x);
                w[j] = nx;
            }
            L &= vmask;
            SP &= vmask;
            if (changed) {
                *reinterpret_cast<uint4*>(rb + 32 * lane) = make_uint4(w[0], w[1], w[2], w[3]);
                *reinterpret_cast<uint4*>(rb + 32 * lane + 16) = make_uint4(w[4], w[5], w[6], w[7]);
            }
            if (CLEAN && H) {
                // U+0130 (C4 B0) -> 'i', U+212A (E2 84 AA) -> 'k': their lower-case forms keep an ASCII letter
                for (int j = 0; j < 32; ++j) {
                    const int p = 32 * lane + j;
                    if (!((vmask >> j) & 1u)) continue;
                    const uint8_t c = rb[p];
                    if (c == 0xB0 && p - 1 >= lbmin && rb[p - 1] == 0xC4) {
                        rb[p] = 'i';
                        L |= 1u << j;
                    } else if (c == 0xAA && p - 2 >= lbmin && rb[p - 1] == 0x84 && rb[p - 2] == 0xE2) {
                        rb[p] = 'k';
                        L |= 1u << j;
                    }
                }
            }
        }
        S.lmask[lane] = L;
        // a window is "plain" when every valid byte is kept or a separator (already-clean text): tokens are then exactly
        // the gaps between separators and pass B skips the kept-byte analysis
        bool plain = __all_sync(0xFFFFFFFFu, (L | SP) == vmask_all);
        // token starts: one after every separator
        const int nsp = __popc(SP);
        int incl = nsp;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= d) incl += v;
        }
        int win_sp = __shfl_sync(0xFFFFFFFFu, incl, 31);
        {
            int base = 1 + incl - nsp;
            uint32_t sp = SP;
            while (sp) {
                const int b = __ffs(sp) - 1;
                sp &= sp - 1u;
                S.tok[base++] = (uint16_t)(32 * lane + b + 1);
            }
        }
        const int my_ll = L ? (32 * lane + 31 - __clz(L)) : -1;
        int last_letter = __reduce_max_sync(0xFFFFFFFFu, my_ll);
        if (WIN > ROW_CH && hi > ROW_CH) {
            // ---- extension of a dialogue's last window: bytes [ROW_CH, hi), one word per lane on lanes 0..15
            const int p4 = ROW_CH + 4 * lane;
            uint32_t l4 = 0, s4 = 0, v4 = 0;
            if (lane < ROW_EXT / 4) {
                const uint32_t x = *reinterpret_cast<const uint32_t*>(rb + p4);
                uint32_t l80, s80, nx, H = 0;
                classify_word<CLEAN>(x, l80, s80, nx, H);
                v4 = low_mask(min(max(hi - p4, 0), 4));
                l4 = movemask4(l80) & v4;
                s4 = movemask4(s80) & v4;
                if (nx != x) *reinterpret_cast<uint32_t*>(rb + p4) = nx;
                if (CLEAN && H) {
                    for (int j = 0; j < 4; ++j) {
                        const int p = p4 + j;
                        if (!((v4 >> j) & 1u)) continue;
                        const uint8_t c = rb[p];
                        if (c == 0xB0 && p - 1 >= lbmin && rb[p - 1] == 0xC4) {
                            rb[p] = 'i';
                            l4 |= 1u << j;
                        } else if (c == 0xAA && p - 2 >= lbmin && rb[p - 1] == 0x84 && rb[p - 2] == 0xE2) {
                            rb[p] = 'k';
                            l4 |= 1u << j;
                        }
                    }
                }
            }
            plain = plain && __all_sync(0xFFFFFFFFu, (l4 | s4) == v4);
            const int sh = 4 * (lane & 7);
            const uint32_t L0 = __reduce_or_sync(0xFFFFFFFFu, lane < 8 ? l4 << sh : 0u);
            const uint32_t L1 = __reduce_or_sync(0xFFFFFFFFu, lane >= 8 ? l4 << sh : 0u);   // lanes >= 16 hold 0
            const uint32_t S0 = __reduce_or_sync(0xFFFFFFFFu, lane < 8 ? s4 << sh : 0u);
            const uint32_t S1 = __reduce_or_sync(0xFFFFFFFFu, lane >= 8 ? s4 << sh : 0u);
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
        // tokens complete in this window: all of them when it is the dialogue's last window, otherwise
        // all but the trailing one (which restarts the next window; a window without any separator is a
        // "giant" token that is carried as hash state instead).
        const bool giant = !last && T == 1;
        const int Tc = last ? T : T - 1;
        long long n_item = item, n_row = row, n_beg = row_beg, n_end = row_end, n_wbase;
        int n_lo, n_hi, n_lbmin;
        bool n_last, have_next = true;
        if (!last) {
            long long cursor = giant ? wbase + hi : wbase + (long long)S.tok[T - 1];
            n_wbase = (cursor & ~15LL) - (giant ? 16 : 0);
            n_lo = (int)(cursor - n_wbase);
            n_lbmin = 0;
        } else {
            n_item = item + stride;
            have_next = n_item < n_items;
            if (have_next) {
                n_row = BIG ? deferred_row(n_item) : n_item;
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
        if (last_letter >= 0) {
            // a letter exists in this window: empties carried from earlier windows were interior
            if (pend && !big && !m.drop_empty) {
                bool isnew = false;
                if (lane == 0) isnew = count_term<MODE>(a, tab, TLOG2, empty_index<MODE>(a, pend), pend, binary);
                D += __popc(__ballot_sync(0xFFFFFFFFu, isnew));
            }
            pend = 0;
            row_letter = true;
        }
        if (!big) {
            if (giant || gt_active) {
                // token 0 of this window continues / starts a multi-window token: lane 0 feeds it serially
                const int s0 = S.tok[0], e0 = giant ? hi : (int)S.tok[1] - 1;
                if (lane == 0)
                    for (int p = s0; p < e0; ++p)
                        if (letter_at(S.lmask, p)) gt.put(rb[p]);
                if (!giant) {  // it ends here (separator or end of dialogue)
                    bool isnew = false, was_empty = false;
                    if (lane == 0) {
                        const uint32_t h = gt.finish();
                        if (gt.n == 0) {
                            was_empty = true;
                        } else if (!(m.sw_cap && is_stop_word(m, swt, sww, h, gt.n, gt.b0, gt.b1, gt.b2))) {
                            const uint32_t fi = term_index<MODE>(a, h, gt.n, gt.b0, gt.b1, gt.b2, rb, S.lmask, -1, -1);
                            if (fi != NOT_IN_VOCAB) isnew = count_term<MODE>(a, tab, TLOG2, fi, 1u, binary);
                        }
                        gt.reset();
                    }
                    D += __popc(__ballot_sync(0xFFFFFFFFu, isnew));
                    was_empty = __shfl_sync(0xFFFFFFFFu, (int)was_empty, 0);
                    if (was_empty) {
                        // empty multi-window token: interior iff a letter follows in this window
                        if (S.tok[0] <= last_letter) {
                            bool nw2 = false;
                            if (lane == 0 && !m.drop_empty) nw2 = count_term<MODE>(a, tab, TLOG2, empty_index<MODE>(a, 1u), 1u, binary);
                            D += __popc(__ballot_sync(0xFFFFFFFFu, nw2));
                        } else {
                            pend += 1;
                        }
                    }
                }
            }
            const int t_first = (giant || gt_active) ? 1 : 0;
            const uint32_t swmask = (uint32_t)m.sw_cap - 1u;
            constexpr int TPL = 1;  // tokens per lane per iteration (2 was measured slower: more idle lanes in the last iteration, 79 registers)
            for (int t0 = t_first; t0 < Tc; t0 += 32 * TPL) {
                if (D + 32u * TPL > (3u << TLOG2) / 4u) {  // keep the table at most 3/4 full
                    const long long gw = (long long)blockIdx.x * ROW_WARPS + warp;
                    if (TL == 9 && BIG == 0 && tab == S.tab && a.spill_tabs != nullptr && gw < SPILL_WARPS) {
                        // carry on in this warp's own wide table: clear it, move the entries over, switch
                        unsigned long long* wt = a.spill_tabs + ((size_t)gw << MED_LOG2);
                        const uint4 ff = make_uint4(0u, ~0u, 0u, ~0u);
                        for (int i = lane; i < (1 << MED_LOG2) / 2; i += 32) reinterpret_cast<uint4*>(wt)[i] = ff;
                        __syncwarp();
                        for (int i = lane; i < (1 << TAB_LOG2); i += 32) {
                            const unsigned long long e = S.tab[i];
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
                // The body is written in warp-uniform phases (boundaries -> hash -> stop-word lookup -> insert)
                // with per-lane predicates instead of nested branches, so the 64 tokens stay in lock step.
                bool valid[TPL], fast[TPL], empty[TPL], slow[TPL];
                int s[TPL], e[TPL];
                uint32_t n[TPL], h[TPL], b0[TPL], b1[TPL], b2[TPL];
                // ---- phase 1: token extent and its kept-byte pattern
#pragma unroll
                for (int u = 0; u < TPL; ++u) {
                    const int t = t0 + 32 * u + lane;
                    valid[u] = t < Tc;
                    const int tt = valid[u] ? t : t_first;
                    s[u] = S.tok[tt];
                    e[u] = (int)S.tok[tt + 1] - 1;
                    const int len = e[u] - s[u];
                    if (plain) {  // every byte between two separators is kept
                        n[u] = (uint32_t)len;
                        fast[u] = valid[u] && len > 0;
                        empty[u] = valid[u] && len == 0;
                        slow[u] = false;
                    } else {
                        const uint32_t bits = __funnelshift_r(S.lmask[s[u] >> 5], S.lmask[(s[u] >> 5) + 1], s[u] & 31) & low_mask(min(len, 32));
                        const int f = __ffs(bits) - 1;  // -1 when no kept byte
                        n[u] = (uint32_t)__popc(bits);
                        fast[u] = valid[u] && len <= 32 && bits != 0u && (bits >> (f & 31)) == low_mask((int)n[u]);
                        empty[u] = valid[u] && len <= 32 && bits == 0u;
                        slow[u] = valid[u] && !fast[u] && !empty[u];
                        s[u] = fast[u] ? s[u] + f : s[u];  // fast tokens: s now points at the first kept byte
                    }
                }
                // ---- phase 2: murmur3 of a contiguous token, first 12 bytes straight-line
                int nn[TPL];
                bool any_long = false;
#pragma unroll
                for (int u = 0; u < TPL; ++u) {
                    const int p0 = fast[u] ? s[u] : 0;
                    nn[u] = fast[u] ? (int)n[u] : 0;
                    const uint32_t* wp = reinterpret_cast<const uint32_t*>(rb + (p0 & ~3));
                    const int sh8 = (p0 & 3) * 8;
                    const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3];
                    b0[u] = __funnelshift_r(w0, w1, sh8) & low_mask(8 * nn[u]);
                    b1[u] = __funnelshift_r(w1, w2, sh8) & low_mask(8 * max(nn[u] - 4, 0));
                    b2[u] = __funnelshift_r(w2, w3, sh8) & low_mask(8 * max(nn[u] - 8, 0));
                    // a block past the token is all zero and mm_k(0) = 0, so "h ^= mm_k(block)" covers both the tail block
                    // and the blocks beyond it: one select per block (full block -> rotate/multiply step)
                    uint32_t hh = 42u;
                    uint32_t t = hh ^ mm_k(b0[u]);
                    hh = nn[u] >= 4 ? rotl32(t, 13) * 5u + 0xe6546b64u : t;
                    t = hh ^ mm_k(b1[u]);
                    hh = nn[u] >= 8 ? rotl32(t, 13) * 5u + 0xe6546b64u : t;
                    t = hh ^ mm_k(b2[u]);
                    hh = nn[u] >= 12 ? rotl32(t, 13) * 5u + 0xe6546b64u : t;
                    h[u] = hh;
                    any_long |= nn[u] > 12;
                }
                if (__any_sync(0xFFFFFFFFu, any_long)) {
#pragma unroll
                    for (int u = 0; u < TPL; ++u) {
                        if (nn[u] > 12) {
                            const uint32_t* wp = reinterpret_cast<const uint32_t*>(rb + (s[u] & ~3));
                            const int sh8 = (s[u] & 3) * 8;
                            uint32_t prev = wp[3];
                            for (int i = 3; nn[u] > 4 * i; ++i) {
                                const uint32_t nxt = wp[i + 1];
                                const uint32_t k = __funnelshift_r(prev, nxt, sh8);
                                prev = nxt;
                                const int r = nn[u] - 4 * i;
                                h[u] = r >= 4 ? mm_h(h[u], mm_k(k)) : h[u] ^ mm_k(k & low_mask(8 * r));
                            }
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < TPL; ++u) {
                    h[u] = mm_fmix(h[u], (uint32_t)nn[u]);
                    if (slow[u]) {  // kept bytes are not contiguous (or the token is longer than 32 bytes): byte-wise walk
                        HashState hs;
                        hs.reset();
                        for (int p = s[u]; p < e[u]; ++p)
                            if (letter_at(S.lmask, p)) hs.put(rb[p]);
                        h[u] = hs.finish();
                        n[u] = hs.n;
                        b0[u] = hs.b0;
                        b1[u] = hs.b1;
                        b2[u] = hs.b2;
                        empty[u] = (n[u] == 0u);
                    }
                }
                // ---- phase 3: stop-word lookup: perfect-hash slot (or two cuckoo slots with 16-bit fingerprints), then
                //      the word record is compared (no loop; the long-word compare is a rare call)
                bool stop[TPL];
#pragma unroll
                for (int u = 0; u < TPL; ++u) {
                    stop[u] = false;
                    if (m.sw_perfect) {  // one probe: 16-bit slot -> word id, then the record is compared
                        // straight-line: the record load is unconditional (id 0 reads record 0 and is masked out), only
                        // the rare words longer than 12 bytes branch
                        const uint32_t id = reinterpret_cast<const uint16_t*>(swt)[(h[u] * m.sw_mul) >> m.sw_shift];
                        const uint4 w = sww[id ? id - 1u : 0u];
                        const bool cand = valid[u] && !empty[u] && id != 0u;
                        stop[u] = cand && (w.x == n[u]) && (w.y == b0[u] && w.z == b1[u] && w.w == b2[u]);
                        if (cand && w.x == (n[u] | 0x80000000u)) stop[u] = long_stop_word_equal(m.sw_pool + w.y, rb, S.lmask, s[u], e[u]);
                    } else if (m.sw_cap) {
                        const uint32_t e1 = swt[h[u] & swmask], e2 = swt[(h[u] >> 10) & swmask];
                        const bool m1 = e1 != 0u && ((e1 ^ h[u]) >> 16) == 0u, m2 = e2 != 0u && ((e2 ^ h[u]) >> 16) == 0u;
                        if (valid[u] && !empty[u] && (m1 || m2)) {
                            const uint32_t ent = m1 ? e1 : e2;
                            uint4 w = sww[(ent & 0xFFFFu) - 1u];
                            stop[u] = (w.x == n[u]) && (w.y == b0[u] && w.z == b1[u] && w.w == b2[u]);
                            if ((w.x & 0x7FFFFFFFu) == n[u] && (w.x & 0x80000000u))
                                stop[u] = long_stop_word_equal(m.sw_pool + w.y, rb, S.lmask, s[u], e[u]);
                            if (!stop[u] && m1 && m2) {  // fingerprint collision in the first slot: look at the second
                                w = sww[(e2 & 0xFFFFu) - 1u];
                                stop[u] = (w.x == n[u]) && (w.y == b0[u] && w.z == b1[u] && w.w == b2[u]);
                                if ((w.x & 0x7FFFFFFFu) == n[u] && (w.x & 0x80000000u))
                                    stop[u] = long_stop_word_equal(m.sw_pool + w.y, rb, S.lmask, s[u], e[u]);
                            }
                        }
                    }
                }
                // ---- phase 4: count the surviving tokens
#pragma unroll
                for (int u = 0; u < TPL; ++u) {
                    const bool interior_empty = empty[u] && s[u] <= last_letter;
                    bool ins = valid[u] && (empty[u] ? (interior_empty && !m.drop_empty) : !stop[u]);
                    if (MODE == MODE_DT && dts_fast) {
                        // small decision tree, power-of-two HashingTF: bucket -> counter slot -> one compare -> one atomic;
                        // nothing is ever "new", so there is no distinct count to maintain
                        const uint32_t fi = empty[u] ? m.empty_idx : (h[u] & (uint32_t)(m.num_features - 1));
                        const uint32_t slot = (fi * m.dt_mul) >> m.dt_shift;
                        if (ins && __ldg(m.dt_tab + slot) == fi) {
                            unsigned int* c = reinterpret_cast<unsigned int*>(tab) + slot;
                            if (binary)
                                *c = 1u;
                            else
                                atomicAdd(c, 1u);
                        }
                        pend += __popc(__ballot_sync(0xFFFFFFFFu, empty[u] && !interior_empty));
                        continue;
                    }
                    uint32_t fi = 0u;
                    if (MODE != MODE_FIT && !m.vocab_cap && m.mod_pow2) {  // hashing trick, power-of-two numFeatures: one AND, computed for every lane
                        fi = empty[u] ? m.empty_idx : feature_index(h[u], m);
                    } else if (ins) {
                        fi = empty[u] ? empty_index<MODE>(a, 1u) : term_index<MODE>(a, h[u], n[u], b0[u], b1[u], b2[u], rb, S.lmask, s[u], e[u]);
                        ins = fi != NOT_IN_VOCAB;
                    }
                    bool isnew = false;
                    if (BIG == 0 && tab == S.tab) {  // (named so that the common case keeps shared-memory atomics)
                        if (ins) isnew = count_term<MODE>(a, S.tab, TAB_LOG2, fi, 1u, binary);
                    } else {
                        if (ins) isnew = count_term<MODE>(a, tab, TLOG2, fi, 1u, binary);
                    }
                    D += __popc(__ballot_sync(0xFFFFFFFFu, isnew));
                    pend += __popc(__ballot_sync(0xFFFFFFFFu, empty[u] && !interior_empty));
                }
            }
        }
        gt_active = giant;
        __syncwarp();

        // ================= end of dialogue =================
        if (last) {
            if (!big && row_sp == 0 && !row_letter && !m.drop_empty) {
                // no separator anywhere and nothing kept: Java's split returns [""] -- one empty token
                bool isnew = false;
                if (lane == 0) isnew = count_term<MODE>(a, tab, TLOG2, empty_index<MODE>(a, 1u), 1u, binary);
                D += __popc(__ballot_sync(0xFFFFFFFFu, isnew));
            }
            __syncwarp();
            if (big) {
                if (BIG) {
                    if (lane == 0) a.ctrl->status = FD_ERR_LIMIT;
                } else if (lane == 0) {
                    // size class by the table the dialogue can need at most
                    if (wide_table_log2(row_end - row_beg) <= MED_LOG2) {
                        const unsigned int pos = atomicAdd(&a.ctrl->big_count, 1u);
                        a.big_rows[pos] = row;
                    } else {
                        const unsigned int pos = atomicAdd(&a.ctrl->big2_count, 1u);
                        a.big_rows[a.big_cap - 1 - pos] = row;
                    }
                }
            } else {
                if (BIG == 0 && tab == S.tab)
                    row_epilogue<MODE>(a, reinterpret_cast<double*>(S.tok), reinterpret_cast<double*>(rb), true, S.tab, TAB_LOG2, row, (int)D, lane);
                else
                    row_epilogue<MODE>(a, reinterpret_cast<double*>(S.tok), reinterpret_cast<double*>(rb), false, tab, TLOG2, row, (int)D, lane);
            }
            // reset the table and the per-dialogue state
            __syncwarp();
            if (BIG && have_next) TLOG2 = wide_table_log2(n_end - n_beg);
            if (TL == 9 && BIG == 0 && tab != S.tab) {  // back to the shared-memory table for the next dialogue
                tab = S.tab;
                TLOG2 = TAB_LOG2;
            }
            if (dts) {
                const uint4 z = make_uint4(0u, 0u, 0u, 0u);
                for (int i = lane; i < (1 << m.dt_log2) / 4; i += 32) reinterpret_cast<uint4*>(tab)[i] = z;
            } else {
                uint4 ff = make_uint4(0u, ~0u, 0u, ~0u);  // count = 0, key = EMPTY
                for (int i = lane; i < (1 << TLOG2) / 2; i += 32) reinterpret_cast<uint4*>(tab)[i] = ff;
            }
            D = 0;
            pend = 0;
            row_sp = 0;
            row_letter = false;
            big = false;
            gt_active = false;
            if (lane == 0) gt.reset();
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
// end-of-dialogue epilogues
// ------------------------------------------------------------------------------------------------
// compact the table in place to entries [0, n) and bitonic-sort them ascending by bucket
// bitonic sort of 32*K packed keys held K per lane (element e = lane*K + r), ascending
template <int K>
__device__ __forceinline__ void bitonic_regs(uint32_t (&v)[K], int lane) {
    constexpr int P = 32 * K;
#pragma unroll
    for (int k = 2; k <= P; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j >= K) {  // partner element lives in lane ^ (j / K), same r
#pragma unroll
                for (int r = 0; r < K; ++r) {
                    const uint32_t o = __shfl_xor_sync(0xFFFFFFFFu, v[r], j / K);
                    const int e = lane * K + r;
                    const bool keep_min = ((e & j) == 0) == ((e & k) == 0);
                    v[r] = keep_min ? min(v[r], o) : max(v[r], o);
                }
            } else {  // partner in the same lane
#pragma unroll
                for (int r = 0; r < K; ++r) {
                    if ((r ^ j) > r) {
                        const int e = lane * K + r;
                        const bool asc = (e & k) == 0;
                        const uint32_t x = v[r], y = v[r ^ j];
                        v[r] = asc ? min(x, y) : max(x, y);
                        v[r ^ j] = asc ? max(x, y) : min(x, y);
                    }
                }
            }
        }
    }
}

template <int K>
__device__ __noinline__ void sort_perm_k(const unsigned long long* tab, int n, uint16_t* perm, int lane) {
    uint32_t v[K];
#pragma unroll
    for (int r = 0; r < K; ++r) {
        const int e = lane * K + r;
        v[r] = e < n ? (((uint32_t)(tab[e] >> 32) << 9) | (uint32_t)e) : 0xFFFFFFFFu;  // (bucket, position)
    }
    bitonic_regs<K>(v, lane);
#pragma unroll
    for (int r = 0; r < K; ++r) {
        const int e = lane * K + r;
        if (e < n) perm[e] = (uint16_t)(v[r] & 511u);
    }
}

// compact the table in place to entries [0, n); then order them ascending by bucket.  Small tables (n <= 512, buckets
// below 2^23) are sorted in REGISTERS -- packed (bucket << 9 | position) keys, one shuffle per compare-exchange -- and
// the order is returned through perm[] (entry i of the sorted sequence is tab[perm[i]]); otherwise the entries
// themselves are bitonic-sorted in place and perm is the identity (returned as nullptr).
__device__ int compact_and_sort(unsigned long long* tab, int log2cap, int lane, bool small_keys, uint16_t* perm_buf,
                                const uint16_t** perm_out) {
    const int cap = 1 << log2cap;
    int w = 0;
    for (int base = 0; base < cap; base += 32) {
        const unsigned long long e = tab[base + lane];
        const bool v = (uint32_t)(e >> 32) != EMPTY_KEY;
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, v);
        __syncwarp();
        if (v) tab[w + __popc(bal & ((1u << lane) - 1u))] = e;
        w += __popc(bal);
        __syncwarp();
    }
    const int n = w;
    *perm_out = nullptr;
    if (n <= 1) return n;
    if (small_keys && n <= 512) {
        if (n <= 32) sort_perm_k<1>(tab, n, perm_buf, lane);
        else if (n <= 64) sort_perm_k<2>(tab, n, perm_buf, lane);
        else if (n <= 128) sort_perm_k<4>(tab, n, perm_buf, lane);
        else if (n <= 256) sort_perm_k<8>(tab, n, perm_buf, lane);
        else sort_perm_k<16>(tab, n, perm_buf, lane);
        __syncwarp();
        *perm_out = perm_buf;
        return n;
    }
    int P = 1;
    while (P < n) P <<= 1;
    for (int i = n + lane; i < P; i += 32) tab[i] = ~0ULL;
    __syncwarp();
    for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int c = lane; c < (P >> 1); c += 32) {
                const int i = ((c & ~(j - 1)) << 1) | (c & (j - 1));
                const int ixj = i | j;
                const unsigned long long x = tab[i], y = tab[ixj];
                const bool asc = (i & k) == 0;
                if ((x > y) == asc) {
                    tab[i] = y;
                    tab[ixj] = x;
                }
            }
            __syncwarp();
        }
    }
    return n;
}

__device__ __forceinline__ double feat_value(const Model& m, uint32_t idx, uint32_t cnt) {
    const double c = (double)cnt;
    return m.idf ? __dmul_rn(c, __ldg(m.idf + idx)) : c;
}

__device__ __forceinline__ void write_result(const RowArgs& a, long long row, int lane, double pred, double p0, double p1,
                                             double r0, double r1) {
    if (lane == 0) a.pred[row] = pred;
    else if (lane == 1) a.prob[2 * row] = p0;
    else if (lane == 2) a.prob[2 * row + 1] = p1;
    else if (lane == 3) a.raw[2 * row] = r0;
    else if (lane == 4) a.raw[2 * row + 1] = r1;
}

// walk one tree with the dialogue's (bucket -> count) table; returns the absolute leaf index
// one 32-byte record per lane in ONE load (LDG.256): a divergent fetch costs the L1 one wavefront per lane and
// instruction, so two 16-byte loads of the same sector would double the cost of the step that bounds forest walks
__device__ __forceinline__ void ldg_256(const uint4* p, uint4& a, uint4& b) {
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
                 : "l"(p));
}

// returns the node id of the leaf the dialogue's count table reaches from `root`
__device__ __forceinline__ int walk_tree(const Model& m, const unsigned long long* tab, int log2cap, const uint16_t* fidx, int root,
                                         const uint32_t* dtc = nullptr) {  // dtc: the small-DT counters (records hold their slots)
    int n = root;
    if (m.int_walk) {
        // two levels per 32-byte record, thresholds pre-quantised to term counts (prepare_fused): a fetch, two table
        // probes and two integer compares decide two levels -- no idf load and no fp64 on the dependent chain
        for (;;) {
            uint4 q0, q1;
            ldg_256((dtc ? m.nodes_dt : m.nodes_s) + 2 * (size_t)n, q0, q1);
            const bool r1 = (int)(dtc ? dtc[q0.x] : tab_lookup(tab, log2cap, fidx, q0.x)) > (int)q1.y;
            const uint32_t cf = r1 ? q0.z : q0.y, cc = r1 ? q1.x : q0.w;
            const int ccs = (int)(r1 ? q1.w : q1.z);
            if ((cf & 0x7FFFFFFFu) == 0x7FFFFFFFu) return (int)cc;  // the child is a leaf
            const bool r2 = (int)(dtc ? dtc[cf & 0x7FFFFFFFu] : tab_lookup(tab, log2cap, fidx, cf & 0x7FFFFFFFu)) > ccs;
            n = (int)(cc & 0x7FFFFFFFu) + (r2 ? 1 : 0);
            if ((r2 ? cc : cf) >> 31) return n;                     // the grandchild is a leaf
        }
    }
    for (;;) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(m.nodes + n));
        Node nd;
        nd.thr = __hiloint2double((int)q.y, (int)q.x);
        nd.feat = (int)q.z;
        nd.child = (int)q.w;
        if (nd.feat < 0) return n;
        const uint32_t cnt = tab_lookup(tab, log2cap, (uint32_t)nd.feat);
        const double x = cnt ? feat_value(m, (uint32_t)nd.feat, cnt) : 0.0;
        n = nd.child + ((x <= nd.thr) ? 0 : 1);
    }
}

// acc + src[0] + src[1] + ... + src[cnt-1], strictly left to right (the order Spark's loops add in); src is 16-byte
// aligned shared memory, read two doubles per load
__device__ __forceinline__ double ordered_sum(double acc, const double* src, int cnt) {
    int l = 0;
#pragma unroll 4
    for (; l + 1 < cnt; l += 2) {
        const double2 v = *reinterpret_cast<const double2*>(src + l);
        acc = __dadd_rn(acc, v.x);
        acc = __dadd_rn(acc, v.y);
    }
    if (l < cnt) acc = __dadd_rn(acc, src[l]);
    return acc;
}

template <int MODE>
__device__ void row_epilogue(const RowArgs& a, double* scratch, double* scratch2, bool smem_tab, unsigned long long* tab, int log2cap,
                             long long row, int D, int lane) {
    const Model& m = a.m;  // scratch: the token-list area (2 KiB+), scratch2: the finished window buffer (1 KiB+); both free here
    const uint16_t* fidx = nullptr;
    if ((MODE == MODE_RF || MODE == MODE_GBT) && smem_tab && m.int_walk && m.n_trees >= 8) {
        build_fidx(tab, log2cap, reinterpret_cast<uint16_t*>(scratch), lane);
        fidx = reinterpret_cast<const uint16_t*>(scratch);
        scratch = scratch2;
    }
    // sorted order of the compacted entries: tab[perm[i]] (register sort) or tab[i] (in-place sort)
    uint16_t* perm_buf = reinterpret_cast<uint16_t*>(scratch) + 512;  // second KiB of the token-list area
    const uint16_t* perm = nullptr;
    const bool small_keys = m.num_features <= (1 << 23);
    if (MODE == MODE_CSR) {
        D = compact_and_sort(tab, log2cap, lane, small_keys, perm_buf, &perm);
        unsigned long long pos = 0;
        if (lane == 0) pos = atomicAdd(&a.ctrl->tmp_cursor, (unsigned long long)D);
        pos = __shfl_sync(0xFFFFFFFFu, pos, 0);
        if ((long long)(pos + D) <= a.tmp_cap) {
            for (int i = lane; i < D; i += 32) a.tmp[pos + i] = tab[perm ? perm[i] : i];
        } else if (lane == 0) {
            a.ctrl->tmp_overflow = 1u;
        }
        if (lane == 0) {
            a.row_nnz[row] = D;
            a.row_tmpoff[row] = (long long)pos;
        }
    } else if (MODE == MODE_LINEAR) {
        D = compact_and_sort(tab, log2cap, lane, small_keys, perm_buf, &perm);
        double sum = 0.0;
        for (int b = 0; b < D; b += 32) {
            const int i = b + lane;
            if (i < D) {
                const unsigned long long e = tab[perm ? perm[i] : i];
                const uint32_t idx = (uint32_t)(e >> 32);
                scratch[lane] = __dmul_rn(feat_value(m, idx, (uint32_t)e), __ldg(m.coef + idx));
            }
            __syncwarp();
            if (lane == 0) sum = ordered_sum(sum, scratch, min(32, D - b));
            __syncwarp();
        }
        sum = __shfl_sync(0xFFFFFFFFu, sum, 0);
        const double mg = __dadd_rn(sum, m.intercept);
        const double p0 = 1.0 / (1.0 + exp(mg));
        write_result(a, row, lane, mg > m.raw_threshold ? 1.0 : 0.0, p0, 1.0 - p0, -mg, mg);
    } else if (MODE == MODE_DT) {
        const int n = walk_tree(m, tab, log2cap, nullptr, __ldg(m.roots), m.dt_small ? reinterpret_cast<const uint32_t*>(tab) : nullptr);
        const double2 st = __ldg(m.leaf2 + n), pr = __ldg(m.leaf_prob + n);
        write_result(a, row, lane, st.y > st.x ? 1.0 : 0.0, pr.x, pr.y, st.x, st.y);
    } else if (MODE == MODE_RF) {
        double v = 0.0;  // lane 0: class 0 votes, lane 1: class 1 votes
        for (int b = 0; b < m.n_trees; b += 32) {
            const int t = b + lane;
            if (t < m.n_trees) {
                const int n = walk_tree(m, tab, log2cap, fidx, __ldg(m.roots + t));
                const double2 st = __ldg(m.leaf2 + n);
                scratch[lane] = st.x;
                scratch[32 + lane] = st.y;
            }
            __syncwarp();
            if (lane < 2) v = ordered_sum(v, scratch + 32 * lane, min(32, m.n_trees - b));
            __syncwarp();
        }
        const double v0 = __shfl_sync(0xFFFFFFFFu, v, 0), v1 = __shfl_sync(0xFFFFFFFFu, v, 1);
        const double tot = __dadd_rn(v0, v1);
        const double p0 = tot != 0.0 ? v0 / tot : v0, p1 = tot != 0.0 ? v1 / tot : v1;
        write_result(a, row, lane, v1 > v0 ? 1.0 : 0.0, p0, p1, v0, v1);
    } else if (MODE == MODE_GBT) {
        double mg = 0.0;
        for (int b = 0; b < m.n_trees; b += 32) {
            const int t = b + lane;
            if (t < m.n_trees) {
                const int n = walk_tree(m, tab, log2cap, fidx, __ldg(m.roots + t));
                scratch[lane] = __dmul_rn(__ldg(&m.nodes[n].thr), __ldg(m.tree_w + t));
            }
            __syncwarp();
            if (lane == 0) mg = ordered_sum(mg, scratch, min(32, m.n_trees - b));
            __syncwarp();
        }
        mg = __shfl_sync(0xFFFFFFFFu, mg, 0);
        const double r0 = -mg;
        const double p0 = 1.0 / (1.0 + exp(-2.0 * r0));
        write_result(a, row, lane, mg > r0 ? 1.0 : 0.0, p0, 1.0 - p0, r0, mg);
    }
}

// ------------------------------------------------------------------------------------------------
// CSR assembly: exclusive scan of per-dialogue nnz, then gather staging -> (idx, tf * idf)
// ------------------------------------------------------------------------------------------------
constexpr int SCAN_BLOCK = 1024;

__global__ void k_scan_block_sums(const int32_t* __restrict__ nnz, long long n, long long* __restrict__ bsum) {
    __shared__ long long sh[32];
    const long long i = (long long)blockIdx.x * SCAN_BLOCK + threadIdx.x;
    long long v = i < n ? nnz[i] : 0;
    for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, d);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        v = sh[threadIdx.x];
        for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, d);
        if (threadIdx.x == 0) bsum[blockIdx.x] = v;
    }
}

// single block: exclusive scan of the block sums in place; total goes to bsum[n_blocks]
__global__ void k_scan_top(long long* __restrict__ bsum, int n_blocks) {
    __shared__ long long sh[SCAN_BLOCK];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n_blocks; base += SCAN_BLOCK) {
        const int i = base + threadIdx.x;
        const long long v = i < n_blocks ? bsum[i] : 0;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int d = 1; d < SCAN_BLOCK; d <<= 1) {
            long long t = threadIdx.x >= d ? sh[threadIdx.x - d] : 0;
            __syncthreads();
            sh[threadIdx.x] += t;
            __syncthreads();
        }
        if (i < n_blocks) bsum[i] = carry + sh[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == 0) carry += sh[SCAN_BLOCK - 1];
        __syncthreads();
    }
    if (threadIdx.x == 0) bsum[n_blocks] = carry;
}

__global__ void k_scan_rowptr(const int32_t* __restrict__ nnz, long long n, const long long* __restrict__ bsum,
                              long long* __restrict__ rowptr) {
    __shared__ long long sh[SCAN_BLOCK];
    const long long i = (long long)blockIdx.x * SCAN_BLOCK + threadIdx.x;
    const long long v = i < n ? nnz[i] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int d = 1; d < SCAN_BLOCK; d <<= 1) {
        long long t = threadIdx.x >= d ? sh[threadIdx.x - d] : 0;
        __syncthreads();
        sh[threadIdx.x] += t;
        __syncthreads();
    }
    if (i < n) rowptr[i] = bsum[blockIdx.x] + sh[threadIdx.x] - v;
    if (i == n - 1) rowptr[n] = bsum[blockIdx.x] + sh[threadIdx.x];
}

__global__ void k_gather_csr(const unsigned long long* __restrict__ tmp, const long long* __restrict__ tmpoff,
                             const long long* __restrict__ rowptr, long long n_rows, const double* __restrict__ idf,
                             int32_t* __restrict__ out_idx, double* __restrict__ out_val) {
    const int lane = threadIdx.x & 31;
    const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long r = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < n_rows; r += nwarps) {
        const long long o = rowptr[r], n = rowptr[r + 1] - o, t = tmpoff[r];
        for (long long i = lane; i < n; i += 32) {
            const unsigned long long e = tmp[t + i];
            const uint32_t idx = (uint32_t)(e >> 32);
            const double c = (double)(uint32_t)e;
            out_idx[o + i] = (int32_t)idx;
            out_val[o + i] = idf ? __dmul_rn(c, __ldg(idf + idx)) : c;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
// large batches: pick the kernel shape from the mean dialogue length, on the device (no host synchronisation)
// (keep: the fraction of a dialogue's terms that reach the count table -- below 1 for a sparse LogisticRegression)
__global__ void k_pick_shape(const int64_t* __restrict__ offs, long long n_rows, float keep, Ctrl* ctrl) {
    const long long mean = (offs[n_rows] - offs[0]) / (n_rows > 0 ? n_rows : 1);
    ctrl->shape = (float)mean * keep <= (float)MEAN_BYTES_FOR_SMALL_TABLES ? 8 : 9;
}

template <int MODE, bool CLEAN, int TL>
static cudaError_t launch_main(fd_handle* h, const RowArgs& a, cudaStream_t st) {
    constexpr int NW = row_warps(TL);
    const size_t smem = a.m.sw_smem_bytes() + (size_t)NW * sizeof(WarpSmem<TL>);
    auto kern = k_rows<MODE, CLEAN, 0, TL>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 32 * NW, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
    long long grid = std::min<long long>((long long)h->sm_count * occ, (a.n_rows + NW - 1) / NW);
    if (grid < 1) grid = 1;
    kern<<<(unsigned)grid, 32 * NW, smem, st>>>(a);
    h->launches += 1;
    return cudaGetLastError();
}

template <int MODE, bool CLEAN>
static cudaError_t launch_wide(fd_handle* h, const RowArgs& a, int med_ctas, int big_workers, cudaStream_t st) {
    const size_t smem_med = a.m.sw_smem_bytes() + (size_t)MED_WARPS * sizeof(WarpSmem<9>);
    const size_t smem_big = a.m.sw_smem_bytes() + sizeof(WarpSmem<9>);
    auto kmed = k_rows<MODE, CLEAN, 1, 9>;
    auto kbig = k_rows<MODE, CLEAN, 2, 9>;
    cudaError_t e = cudaFuncSetAttribute(kmed, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_med);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(kbig, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_big);
    if (e != cudaSuccess) return e;
    kmed<<<med_ctas, 32 * MED_WARPS, smem_med, st>>>(a);
    kbig<<<big_workers, 32, smem_big, st>>>(a);
    h->launches += 2;
    return cudaGetLastError();
}

// Micro-batches always take the 512-slot shape (fewer deferrals, lower latency).  Large batches let the device
// choose: both shapes are launched and the one that was not picked exits at once.
template <int MODE, bool CLEAN>
static cudaError_t launch_mode_workers(fd_handle* h, const RowArgs& a0, int med_ctas, int big_workers, cudaStream_t st) {
    RowArgs a = a0;
    cudaError_t e;
    if (a.n_rows >= ROWS_FOR_SMALL_TABLES) {
        a.pick_shape = 1;
        k_pick_shape<<<1, 1, 0, st>>>(a.offs, a.n_rows, (MODE == MODE_LINEAR && a.m.lr_sparse) ? a.m.lr_keep : 1.0f, a.ctrl);
        h->launches += 1;
        e = launch_main<MODE, CLEAN, 8>(h, a, st);
        if (e != cudaSuccess) return e;
    }
    e = launch_main<MODE, CLEAN, 9>(h, a, st);
    if (e != cudaSuccess) return e;
    return launch_wide<MODE, CLEAN>(h, a, med_ctas, big_workers, st);
}

// the caller provides a.big_tabs with wide_pool_slots(med_ctas, big_workers) slots
cudaError_t launch_rows_workers(fd_handle* h, int mode, const RowArgs& a, int med_ctas, int big_workers, cudaStream_t st) {
    const bool c = (a.flags & FD_FLAG_CLEAN) != 0;
#define FD_DISPATCH(M) return c ? launch_mode_workers<M, true>(h, a, med_ctas, big_workers, st) : launch_mode_workers<M, false>(h, a, med_ctas, big_workers, st)
    switch (mode) {
        case MODE_CSR: FD_DISPATCH(MODE_CSR);
        case MODE_LINEAR: FD_DISPATCH(MODE_LINEAR);
        case MODE_DT: FD_DISPATCH(MODE_DT);
        case MODE_RF: FD_DISPATCH(MODE_RF);
        case MODE_GBT: FD_DISPATCH(MODE_GBT);
        case MODE_FIT: FD_DISPATCH(MODE_FIT);
    }
#undef FD_DISPATCH
    return cudaErrorInvalidValue;
}

cudaError_t launch_rows(fd_handle* h, int mode, const RowArgs& a, cudaStream_t st) {
    return launch_rows_workers(h, mode, a, MED_CTAS, BIG_WORKERS, st);
}

cudaError_t launch_csr_assembly(fd_handle* h, const int32_t* nnz, const long long* tmpoff, const unsigned long long* tmp,
                                long long n_rows, long long* bsum, long long* rowptr, int32_t* out_idx, double* out_val,
                                bool scan_only, cudaStream_t st) {
    const int nb = (int)((n_rows + SCAN_BLOCK - 1) / SCAN_BLOCK);
    if (scan_only) {
        k_scan_block_sums<<<nb, SCAN_BLOCK, 0, st>>>(nnz, n_rows, bsum);
        k_scan_top<<<1, SCAN_BLOCK, 0, st>>>(bsum, nb);
        h->launches += 2;
        return cudaGetLastError();
    }
    k_scan_rowptr<<<nb, SCAN_BLOCK, 0, st>>>(nnz, n_rows, bsum, rowptr);
    long long g = (n_rows + 7) / 8;
    if (g > (long long)h->sm_count * 16) g = (long long)h->sm_count * 16;
    k_gather_csr<<<(unsigned)g, 256, 0, st>>>(tmp, tmpoff, rowptr, n_rows, h->m.idf, out_idx, out_val);
    h->launches += 2;
    return cudaGetLastError();
}

}  // namespace fd
