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
#include "fd_rows_dev.cuh"

namespace fd {

// ------------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------------
// BIG = 0: main kernel (count table in shared memory); 1: medium wide tables (8 warps per CTA); 2: large wide tables
// (the body is a device function: k_rows runs one variant per launch, k_wide runs the two wide-table variants side by side;
//  bid / nblocks = this CTA's index and the CTA count WITHIN its variant)
template <int MODE, bool CLEAN, int BIG, int TL>
__device__ __forceinline__ void rows_body(const RowArgs& a, const int bid, const int nblocks) {
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

    if (BIG && (long long)bid * ROW_WARPS >= (long long)(BIG == 1 ? a.ctrl->big_count : a.ctrl->big2_count))
        return;  // nothing deferred to this CTA
    if (!BIG && a.pick_shape && shape_for(a.offs[a.n_rows] - a.offs[0], a.n_rows, a.pick_keep) != TL) return;  // the other shape runs this call
    for (int i = threadIdx.x; i < (int)(m.sw_tab_bytes() / 4); i += 32 * ROW_WARPS) swt[i] = m.sw_tab[i];
    for (int i = threadIdx.x; i < m.n_sw; i += 32 * ROW_WARPS) sww[i] = m.sw_words[i];
    if (lane == 0) {
        mbar_init(&S.mbar[0], 1);
        mbar_init(&S.mbar[1], 1);
        fence_mbar_init();
    }
    if (lane < 4) S.lmask[32 + lane] = 0u;

    unsigned long long* tab = BIG == 1 ? a.big_tabs + ((size_t)(bid * MED_WARPS + warp) << MED_LOG2)
                            : (BIG == 2 ? a.big2_tabs + ((size_t)bid << BIG_LOG2) : S.tab);
    // table size: fixed in shared memory; in the wide-table kernel it follows the dialogue (a dialogue of b bytes
    // has at most b/2 + 1 tokens), so clearing it costs in proportion to the dialogue, not to the 2^16 maximum
    int TLOG2 = TAB_LOG2;
    const long long n_items = BIG == 1 ? (long long)a.ctrl->big_count : (BIG == 2 ? (long long)a.ctrl->big2_count : a.n_rows);
    const long long stride = (long long)nblocks * ROW_WARPS;
    long long item = (long long)bid * ROW_WARPS + warp;
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
    if (BIG == 2) __syncwarp();  // (one warp per CTA of this variant; inside k_wide the CTA's other warps have left)
    else __syncthreads();
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
                    const long long gw = (long long)bid * ROW_WARPS + warp;
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
                        if (ins && (__ldg(m.dt_tab + slot) & 0x7FFFFFFFu) == fi) {
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


template <int MODE, bool CLEAN, int BIG, int TL>
__global__ void __launch_bounds__(BIG == 2 ? 32 : (BIG == 1 ? 32 * MED_WARPS : 32 * row_warps(TL)), BIG ? 1 : (TL == 8 ? 2 : 3))
k_rows(const RowArgs a) {
    rows_body<MODE, CLEAN, BIG, TL>(a, (int)blockIdx.x, (int)gridDim.x);
}

// Both wide-table variants in ONE launch: CTAs [0, med_ctas) take the medium dialogues (8 warps each), the following
// big_workers CTAs the large ones (their first warp only).  With nothing deferred every CTA leaves after one load.
template <int MODE, bool CLEAN>
__global__ void __launch_bounds__(32 * MED_WARPS, 1) k_wide(const RowArgs a, const int med_ctas, const int big_workers) {
    if ((int)blockIdx.x < med_ctas) {
        rows_body<MODE, CLEAN, 1, 9>(a, (int)blockIdx.x, med_ctas);
    } else {
        if (threadIdx.x >= 32) return;
        rows_body<MODE, CLEAN, 2, 9>(a, (int)blockIdx.x - med_ctas, big_workers);
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

// One dialogue's staged (count, bucket) entries -> its CSR row, ascending by bucket.  Up to 512 entries with buckets below
// 2^23 are ordered in REGISTERS: the entries are read coalesced into shared memory as packed (bucket << 9 | position) keys
// plus their counts, each lane then takes K consecutive keys, a bitonic network orders them (one shuffle per compare-exchange
// across lanes), and the ordered keys go back through shared memory so that the row is written coalesced.
constexpr int GATHER_WARPS = 8;
template <int K>
__device__ __forceinline__ void gather_sorted_k(const unsigned long long* __restrict__ src, int n, uint32_t* keys, uint32_t* cnts, int lane,
                                                const double* __restrict__ idf, int32_t* __restrict__ oi, double* __restrict__ ov) {
#pragma unroll
    for (int r = 0; r < K; ++r) {
        const int e = r * 32 + lane;
        uint32_t key = 0xFFFFFFFFu;
        if (e < n) {
            const unsigned long long x = src[e];
            key = ((uint32_t)(x >> 32) << 9) | (uint32_t)e;
            cnts[e] = (uint32_t)x;
        }
        keys[e] = key;
    }
    __syncwarp();
    uint32_t v[K];
#pragma unroll
    for (int r = 0; r < K; ++r) v[r] = keys[lane * K + r];
    __syncwarp();
    bitonic_regs<K>(v, lane);
#pragma unroll
    for (int r = 0; r < K; ++r) keys[lane * K + r] = v[r];
    __syncwarp();
#pragma unroll
    for (int r = 0; r < K; ++r) {
        const int e = r * 32 + lane;
        if (e < n) {
            const uint32_t key = keys[e];
            const uint32_t idx = key >> 9;
            const double c = (double)cnts[key & 511u];
            oi[e] = (int32_t)idx;
            ov[e] = idf ? __dmul_rn(c, __ldg(idf + idx)) : c;
        }
    }
    __syncwarp();
}

// Hashed buckets are spread evenly over [0, numFeatures), so a COUNTING SORT orders a dialogue's entries in a few dozen
// instructions: bin = the bucket's top bits (256 bins, monotone in the bucket), a shared-memory atomic gives every entry its
// arrival rank inside its bin, a scan turns the bin sizes into start positions, and entries that share a bin (a handful per
// dialogue) are ordered by counting the bin's smaller keys.  bin_mul = floor(2^40 / numFeatures), 0 when numFeatures <= 256
// (bin = bucket).  
template <int K>
__device__ __forceinline__ void gather_counted_k(const unsigned long long* __restrict__ src, int n, uint32_t* keys, uint32_t* fkeys,
                                                 uint32_t* cnts, uint32_t* bins, uint32_t bin_mul, int lane, const double* __restrict__ idf,
                                                 int32_t* __restrict__ oi, double* __restrict__ ov) {
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    reinterpret_cast<uint4*>(bins)[2 * lane] = z;
    reinterpret_cast<uint4*>(bins)[2 * lane + 1] = z;
    __syncwarp();
    uint32_t key[K], rb[K], cn[K];  // rb = arrival rank << 8 | bin
#pragma unroll
    for (int r = 0; r < K; ++r) {
        const int e = r * 32 + lane;
        key[r] = 0xFFFFFFFFu;
        rb[r] = 0u;
        cn[r] = 0u;
        if (e < n) {
            const unsigned long long x = src[e];
            const uint32_t b = (uint32_t)(x >> 32);
            key[r] = b;
            cn[r] = (uint32_t)x;
            const uint32_t bin = bin_mul ? __umulhi(b, bin_mul) : b;
            rb[r] = (atomicAdd(&bins[bin], 1u) << 8) | bin;
        }
    }
    __syncwarp();
    {   // exclusive scan of the 256 bin sizes: 8 consecutive bins per lane
        uint4 p = reinterpret_cast<uint4*>(bins)[2 * lane], q = reinterpret_cast<uint4*>(bins)[2 * lane + 1];
        const uint32_t tot = p.x + p.y + p.z + p.w + q.x + q.y + q.z + q.w;
        uint32_t inc = tot;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, inc, d);
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
        __syncwarp();
        reinterpret_cast<uint4*>(bins)[2 * lane] = o0;
        reinterpret_cast<uint4*>(bins)[2 * lane + 1] = o1;
    }
    __syncwarp();
    uint32_t start[K];
#pragma unroll
    for (int r = 0; r < K; ++r) {
        start[r] = bins[rb[r] & 255u];
        if (r * 32 + lane < n) keys[start[r] + (rb[r] >> 8)] = key[r];
    }
    __syncwarp();
#pragma unroll
    for (int r = 0; r < K; ++r) {
        if (r * 32 + lane < n) {
            const uint32_t bin = rb[r] & 255u;
            const uint32_t end = bin == 255u ? (uint32_t)n : bins[bin + 1u];
            uint32_t pos = start[r];
            for (uint32_t j = start[r]; j < end; ++j) pos += keys[j] < key[r] ? 1u : 0u;  // (keys of one dialogue are distinct)
            fkeys[pos] = key[r];
            cnts[pos] = cn[r];
        }
    }
    __syncwarp();
    // the ordered row leaves through shared memory so that the global stores are coalesced
#pragma unroll
    for (int r = 0; r < K; ++r) {
        const int e = r * 32 + lane;
        if (e < n) {
            const uint32_t idx = fkeys[e];
            const double c = (double)cnts[e];
            oi[e] = (int32_t)idx;
            ov[e] = idf ? __dmul_rn(c, __ldg(idf + idx)) : c;
        }
    }
    __syncwarp();
}

// rows the dialogue kernel staged already ordered (token-list path, flagged in tmpoff): only the values are formed.
// No shared memory and few registers, so 64 warps per SM hide the two dependent loads (entry, idf) of a row.
__global__ void __launch_bounds__(256) k_gather_copy(const unsigned long long* __restrict__ tmp, const long long* __restrict__ tmpoff,
                                                     const long long* __restrict__ rowptr, long long n_rows, const double* __restrict__ idf,
                                                     int32_t* __restrict__ out_idx, double* __restrict__ out_val) {
    const int lane = threadIdx.x & 31;
    const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long r = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < n_rows; r += nwarps) {
        const long long toff = tmpoff[r];
        if (!(toff & (1LL << 62))) continue;
        const long long o = rowptr[r], n = rowptr[r + 1] - o, t = toff & ~(1LL << 62);
        for (long long i = lane; i < n; i += 32) {
            const unsigned long long e = tmp[t + i];
            const uint32_t idx = (uint32_t)(e >> 32);
            const double c = (double)(uint32_t)e;
            out_idx[o + i] = (int32_t)idx;
            out_val[o + i] = idf ? __dmul_rn(c, __ldg(idf + idx)) : c;
        }
    }
}

__global__ void __launch_bounds__(32 * GATHER_WARPS, 4) k_gather_csr(unsigned long long* __restrict__ tmp, const long long* __restrict__ tmpoff,
                                                                     const long long* __restrict__ rowptr, long long n_rows,
                                                                     const double* __restrict__ idf, int small_keys, int even_keys,
                                                                     uint32_t bin_mul, int rpw, int32_t* __restrict__ out_idx,
                                                                     double* __restrict__ out_val) {
    __shared__ __align__(16) uint32_t key_sh[GATHER_WARPS][512];
    __shared__ __align__(16) uint32_t cnt_sh[GATHER_WARPS][512];
    __shared__ __align__(16) uint32_t bin_sh[GATHER_WARPS][256];
    __shared__ __align__(16) uint32_t fkey_sh[GATHER_WARPS][256];
    const int lane = threadIdx.x & 31;
    uint32_t* keys = key_sh[threadIdx.x >> 5];
    uint32_t* cnts = cnt_sh[threadIdx.x >> 5];
    uint32_t* bins = bin_sh[threadIdx.x >> 5];
    uint32_t* fkeys = fkey_sh[threadIdx.x >> 5];
    // a warp takes rpw (<= 32) consecutive dialogues at a time: their row pointers and staging offsets are read by the 32 lanes at
    // once (one coalesced round trip instead of three dependent ones per dialogue) and handed out by shuffles
    const long long nwarps = (long long)gridDim.x * GATHER_WARPS;
    for (long long r0 = ((long long)blockIdx.x * GATHER_WARPS + (threadIdx.x >> 5)) * rpw; r0 < n_rows; r0 += nwarps * rpw) {
      const long long mine = r0 + lane < n_rows ? r0 + lane : n_rows - 1;
      const long long my_o = rowptr[mine], my_e = rowptr[mine + 1], my_t = tmpoff[mine];
      const int nr = (int)(n_rows - r0 < rpw ? n_rows - r0 : rpw);
      for (int k = 0; k < nr; ++k) {
        const long long o = __shfl_sync(0xFFFFFFFFu, my_o, k), nl = __shfl_sync(0xFFFFFFFFu, my_e, k) - o;
        const long long toff = __shfl_sync(0xFFFFFFFFu, my_t, k);
        unsigned long long* src = tmp + (toff & ~(1LL << 62));
        int32_t* oi = out_idx + o;
        double* ov = out_val + o;
        if (toff & (1LL << 62)) continue;  // staged already ordered: k_gather_copy formed this row
        if (even_keys && nl <= 256) {
            const int n = (int)nl;
            if (n <= 32) gather_counted_k<1>(src, n, keys, fkeys, cnts, bins, bin_mul, lane, idf, oi, ov);
            else if (n <= 64) gather_counted_k<2>(src, n, keys, fkeys, cnts, bins, bin_mul, lane, idf, oi, ov);
            else if (n <= 128) gather_counted_k<4>(src, n, keys, fkeys, cnts, bins, bin_mul, lane, idf, oi, ov);
            else gather_counted_k<8>(src, n, keys, fkeys, cnts, bins, bin_mul, lane, idf, oi, ov);
            continue;
        }
        if (small_keys && nl <= 512) {
            const int n = (int)nl;
            if (n <= 32) gather_sorted_k<1>(src, n, keys, cnts, lane, idf, oi, ov);
            else if (n <= 64) gather_sorted_k<2>(src, n, keys, cnts, lane, idf, oi, ov);
            else if (n <= 128) gather_sorted_k<4>(src, n, keys, cnts, lane, idf, oi, ov);
            else if (n <= 256) gather_sorted_k<8>(src, n, keys, cnts, lane, idf, oi, ov);
            else gather_sorted_k<16>(src, n, keys, cnts, lane, idf, oi, ov);
            continue;
        }
        // long rows (or buckets of 2^23 and more): the 64-bit entries (bucket in the high word) are sorted in place in the
        // staging buffer.  Normalised bitonic network -- every comparator ascending, the first step of a merge pairs i with
        // its mirror i ^ (k - 1) -- so the positions from n up to the next power of two act as +infinity and are skipped.
        long long P = 1;
        while (P < nl) P <<= 1;
        for (long long k = 2; k <= P; k <<= 1) {
            for (long long j = k >> 1; j > 0; j >>= 1) {
                const bool mirror = j == (k >> 1);
                for (long long c = lane; c < (P >> 1); c += 32) {
                    const long long i = ((c & ~(j - 1)) << 1) | (c & (j - 1));
                    const long long q = mirror ? (i ^ (k - 1)) : (i | j);
                    if (q < nl) {
                        const unsigned long long x = src[i], y = src[q];
                        if (x > y) {
                            src[i] = y;
                            src[q] = x;
                        }
                    }
                }
                __syncwarp();
            }
        }
        for (long long i = lane; i < nl; i += 32) {
            const unsigned long long e = src[i];
            const uint32_t idx = (uint32_t)(e >> 32);
            const double c = (double)(uint32_t)e;
            oi[i] = (int32_t)idx;
            ov[i] = idf ? __dmul_rn(c, __ldg(idf + idx)) : c;
        }
      }
    }
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
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

static bool func_ready(fd_handle* h, const void* fn, size_t smem) {  // shared-memory opt-in done for (kernel, size)?
    for (const auto& c : h->launch_cfg)
        if (c.fn == fn && c.smem == smem) return true;
    return false;
}

// merged: both variants in one launch (their table pools must not overlap: RowArgs::big_tabs / big2_tabs);
// otherwise one launch each, medium first (the pools may then be the same memory -- fd_transform_noalloc's fixed array)
template <int MODE, bool CLEAN>
static cudaError_t launch_wide(fd_handle* h, const RowArgs& a, int med_ctas, int big_workers, bool merged, cudaStream_t st) {
    const size_t smem_med = a.m.sw_smem_bytes() + (size_t)MED_WARPS * sizeof(WarpSmem<9>);
    const size_t smem_big = a.m.sw_smem_bytes() + sizeof(WarpSmem<9>);
    if (merged) {
        auto kern = k_wide<MODE, CLEAN>;
        if (!func_ready(h, (const void*)kern, smem_med)) {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_med);
            if (e != cudaSuccess) return e;
            h->launch_cfg.push_back({(const void*)kern, smem_med, 1});
        }
        kern<<<med_ctas + big_workers, 32 * MED_WARPS, smem_med, st>>>(a, med_ctas, big_workers);
        h->launches += 1;
        return cudaGetLastError();
    }
    auto kmed = k_rows<MODE, CLEAN, 1, 9>;
    auto kbig = k_rows<MODE, CLEAN, 2, 9>;
    if (!func_ready(h, (const void*)kmed, smem_med)) {
        cudaError_t e = cudaFuncSetAttribute(kmed, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_med);
        if (e != cudaSuccess) return e;
        h->launch_cfg.push_back({(const void*)kmed, smem_med, 1});
    }
    if (!func_ready(h, (const void*)kbig, smem_big)) {
        cudaError_t e = cudaFuncSetAttribute(kbig, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_big);
        if (e != cudaSuccess) return e;
        h->launch_cfg.push_back({(const void*)kbig, smem_big, 1});
    }
    kmed<<<med_ctas, 32 * MED_WARPS, smem_med, st>>>(a);
    kbig<<<big_workers, 32, smem_big, st>>>(a);
    h->launches += 2;
    return cudaGetLastError();
}

// Main kernel first (fd_main.cu; CountVectorizer.fit keeps the kernel of this file, whose global term table must see
// every token exactly once), then the two wide-table kernels over the dialogues it deferred.
template <int MODE, bool CLEAN>
static cudaError_t launch_mode_workers(fd_handle* h, const RowArgs& a0, int med_ctas, int big_workers, bool merged, cudaStream_t st) {
    cudaError_t e;
    if (MODE == MODE_FIT) {
        // micro-batches always take the 512-slot shape; large batches let the device choose: both shapes are launched
        // and the one that was not picked exits at once
        RowArgs a = a0;
        if (a.n_rows >= ROWS_FOR_SMALL_TABLES) {
            a.pick_shape = 1;
            a.pick_keep = 1.0f;
            e = launch_main<MODE_FIT, CLEAN, 8>(h, a, st);
            if (e != cudaSuccess) return e;
        }
        e = launch_main<MODE_FIT, CLEAN, 9>(h, a, st);
        if (e != cudaSuccess) return e;
        return launch_wide<MODE_FIT, CLEAN>(h, a, med_ctas, big_workers, merged, st);
    }
    e = launch_main_rows(h, MODE, a0, st);
    if (e != cudaSuccess) return e;
    return launch_wide<MODE, CLEAN>(h, a0, med_ctas, big_workers, merged, st);
}

// the caller provides a.big_tabs with wide_pool_slots(med_ctas, big_workers) slots
cudaError_t launch_rows_workers(fd_handle* h, int mode, const RowArgs& a, int med_ctas, int big_workers, bool merged, cudaStream_t st) {
    const bool c = (a.flags & FD_FLAG_CLEAN) != 0;
#define FD_DISPATCH(M) return c ? launch_mode_workers<M, true>(h, a, med_ctas, big_workers, merged, st) : launch_mode_workers<M, false>(h, a, med_ctas, big_workers, merged, st)
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
    return launch_rows_workers(h, mode, a, MED_CTAS, BIG_WORKERS, true, st);
}

cudaError_t launch_csr_assembly(fd_handle* h, const int32_t* nnz, const long long* tmpoff, unsigned long long* tmp,
                                long long n_rows, long long* bsum, long long* rowptr, int32_t* out_idx, double* out_val,
                                bool scan_only, long long unsorted_rows, cudaStream_t st) {
    const int nb = (int)((n_rows + SCAN_BLOCK - 1) / SCAN_BLOCK);
    if (scan_only) {
        k_scan_block_sums<<<nb, SCAN_BLOCK, 0, st>>>(nnz, n_rows, bsum);
        k_scan_top<<<1, SCAN_BLOCK, 0, st>>>(bsum, nb);
        h->launches += 2;
        return cudaGetLastError();
    }
    k_scan_rowptr<<<nb, SCAN_BLOCK, 0, st>>>(nnz, n_rows, bsum, rowptr);
    // rows staged in order (token-list path) are copied; the others (count-table path, deferred dialogues) are ordered here
    if (h->m.csr_list) {
        long long gc = (n_rows + 7) / 8;
        if (gc > (long long)h->sm_count * 8) gc = (long long)h->sm_count * 8;
        k_gather_copy<<<(unsigned)gc, 256, 0, st>>>(tmp, tmpoff, rowptr, n_rows, h->m.idf, out_idx, out_val);
        h->launches += 1;
    }
    if (!h->m.csr_list || unsorted_rows > 0) {
        const int rpw = (int)std::min<long long>(32, std::max<long long>(1, n_rows / 8192));  // dialogues per warp and round
        long long g = (n_rows + rpw * GATHER_WARPS - 1) / (rpw * GATHER_WARPS);
        if (g > (long long)h->sm_count * 32) g = (long long)h->sm_count * 32;
        // hashed buckets are evenly spread (counting sort); vocabulary indices are not (frequent terms have the low ones)
        const int even_keys = h->m.vocab_cap ? 0 : 1;
        const uint32_t bin_mul = h->m.num_features > 256 ? (uint32_t)((1ULL << 40) / (uint64_t)h->m.num_features) : 0u;
        k_gather_csr<<<(unsigned)g, 32 * GATHER_WARPS, 0, st>>>(tmp, tmpoff, rowptr, n_rows, h->m.idf, h->m.num_features <= (1 << 23) ? 1 : 0, even_keys,
                                                           bin_mul, rpw, out_idx, out_val);
        h->launches += 1;
    }
    h->launches += 1;  // (the row-pointer scan)
    return cudaGetLastError();
}

}  // namespace fd
