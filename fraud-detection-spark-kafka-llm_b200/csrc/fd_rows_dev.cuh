// fd_rows_dev.cuh -- device-side pieces shared by the dialogue kernels (fd_rows.cu: wide-table and fit kernels,
// fd_main.cu: the main kernel): TMA/mbarrier PTX wrappers, byte classification, murmur3 state, stop-word and vocabulary
// lookups, the per-dialogue count table, and the end-of-dialogue epilogues (CSR emission, LR dot, tree walks).
#pragma once
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
static __device__ __noinline__ bool long_word_equal(const uint8_t* word, const uint8_t* rb, const uint32_t* lmask, int s, int e) {
    uint32_t i = 0;
    for (int p = s; p < e; ++p)
        if (((lmask[p >> 5] >> (p & 31)) & 1u) && rb[p] != word[i++]) return false;
    return true;
}

// CountVectorizerModel: term -> vocabulary index (NOT_IN_VOCAB when absent).  h = murmur3(term), n = its length,
// b0..b2 = its first 12 bytes; [s, e) locates it in the window for the byte compare of longer words (s < 0: a
// multi-window token, never matched against words longer than 12 bytes -- documented limitation).
static __device__ uint32_t vocab_index(const Model& m, uint32_t h, uint32_t n, uint32_t b0, uint32_t b1, uint32_t b2, const uint8_t* rb,
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
static __device__ uint32_t fit_slot(const RowArgs& a, uint32_t h, uint32_t n, uint32_t b0, uint32_t b1, uint32_t b2, const uint8_t* rb,
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
        if ((__ldg(a.m.dt_tab + slot) & 0x7FFFFFFFu) == fi) {  // (bit 31: stop-word hint of the main kernel)
            unsigned int* c = reinterpret_cast<unsigned int*>(tab) + slot;
            if (binary)
                *c = 1u;
            else
                atomicAdd(c, inc);
        }
        return false;
    }
    // terms the classifier cannot see (zero LogisticRegression coefficient, feature no node splits on) are dropped here
    if ((MODE == MODE_LINEAR || MODE == MODE_DT || MODE == MODE_RF || MODE == MODE_GBT) && a.m.filt_on &&
        !((__ldg(&a.m.filt[fi >> 5].x) >> (fi & 31)) & 1u))
        return false;
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

// dense byte counts over forest-local feature ids (Model::dense_on): zeroed, then every entry of the dialogue's count table
// whose feature the forest splits on stores min(count, 255) at its local id
__device__ __forceinline__ void build_dense(const unsigned long long* tab, int log2cap, uint8_t* dense, const uint16_t* f_local, int lane) {
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (int i = lane; i < DENSE_MAX / 16; i += 32) reinterpret_cast<uint4*>(dense)[i] = z;
    __syncwarp();
    const int cap = 1 << log2cap;
    for (int i = lane; i < cap; i += 32) {
        const unsigned long long e = tab[i];
        const uint32_t k = (uint32_t)(e >> 32);
        if (k != EMPTY_KEY) {
            const uint32_t loc = __ldg(f_local + k);
            if (loc != 0xFFFFu) dense[loc] = (uint8_t)min((uint32_t)e, 255u);
        }
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
static __device__ __noinline__ bool long_stop_word_equal(const uint8_t* word, const uint8_t* rb, const uint32_t* lmask, int s, int e) {
    uint32_t i = 0;
    for (int p = s; p < e; ++p)
        if (letter_at(lmask, p) && rb[p] != word[i++]) return false;
    return true;
}

// is the multi-window token (hash h, n kept bytes, first 12 bytes b0..b2) a stop word?  (words longer than 12
// bytes are not matched on this rare path -- documented limitation)
static __device__ bool is_stop_word(const Model& m, const uint32_t* swt, const uint4* sww, uint32_t h, uint32_t n, uint32_t b0,
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
static __device__ int compact_and_sort(unsigned long long* tab, int log2cap, int lane, bool small_keys, uint16_t* perm_buf,
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

// the same two-levels-per-record walk over nodes_l with the dense byte counts: a node visit is one byte load
__device__ __forceinline__ int walk_tree_dense(const Model& m, const uint8_t* dense, int root) {
    int n = root;
    for (;;) {
        uint4 q0, q1;
        ldg_256(m.nodes_l + 2 * (size_t)n, q0, q1);
        const bool r1 = (int)dense[q0.x] > (int)q1.y;
        const uint32_t cf = r1 ? q0.z : q0.y, cc = r1 ? q1.x : q0.w;
        const int ccs = (int)(r1 ? q1.w : q1.z);
        if ((cf & 0x7FFFFFFFu) == 0x7FFFFFFFu) return (int)cc;  // the child is a leaf
        const bool r2 = (int)dense[cf & 0x7FFFFFFFu] > ccs;
        n = (int)(cc & 0x7FFFFFFFu) + (r2 ? 1 : 0);
        if ((r2 ? cc : cf) >> 31) return n;                     // the grandchild is a leaf
    }
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
    const uint8_t* dense = nullptr;
    if ((MODE == MODE_RF || MODE == MODE_GBT) && smem_tab && m.int_walk && m.n_trees >= 8) {
        if (m.dense_on) {
            build_dense(tab, log2cap, reinterpret_cast<uint8_t*>(scratch), m.f_local, lane);
            dense = reinterpret_cast<const uint8_t*>(scratch);
        } else {
            build_fidx(tab, log2cap, reinterpret_cast<uint16_t*>(scratch), lane);
            fidx = reinterpret_cast<const uint16_t*>(scratch);
        }
        scratch = scratch2;
    }
    // sorted order of the compacted entries: tab[perm[i]] (register sort) or tab[i] (in-place sort)
    uint16_t* perm_buf = reinterpret_cast<uint16_t*>(scratch) + 512;  // second KiB of the token-list area
    const uint16_t* perm = nullptr;
    const bool small_keys = m.num_features <= (1 << 23);
    if (MODE == MODE_CSR) {
        // The table's entries go to the staging buffer in slot order, unsorted: the gather kernel (k_gather_csr) orders each
        // dialogue's entries by bucket while it forms the values.  D = the number of occupied slots (counted at insertion).
        unsigned long long pos = 0;
        if (lane == 0) pos = atomicAdd(&a.ctrl->tmp_cursor, (unsigned long long)D);
        pos = __shfl_sync(0xFFFFFFFFu, pos, 0);
        const bool fits = (long long)(pos + D) <= a.tmp_cap;
        if (!fits && lane == 0) a.ctrl->tmp_overflow = 1u;
        unsigned long long* dst = a.tmp + pos;
        const int cap = 1 << log2cap;
        const uint32_t below = (1u << lane) - 1u;
        uint32_t w = 0;
        for (int base = 0; base < cap; base += 32) {
            const unsigned long long e = tab[base + lane];
            const bool v = (uint32_t)(e >> 32) != EMPTY_KEY;
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, v);
            if (v && fits) dst[w + __popc(bal & below)] = e;
            w += __popc(bal);
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
                const int n = dense ? walk_tree_dense(m, dense, __ldg(m.roots + t)) : walk_tree(m, tab, log2cap, fidx, __ldg(m.roots + t));
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
                const int n = dense ? walk_tree_dense(m, dense, __ldg(m.roots + t)) : walk_tree(m, tab, log2cap, fidx, __ldg(m.roots + t));
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

}  // namespace fd
