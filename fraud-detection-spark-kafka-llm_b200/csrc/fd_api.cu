// fd_api.cu -- the extern "C" boundary (include/fdb200.h): handle lifetime, model upload, and the
// host-side orchestration of the kernels in fd_rows.cu / fd_score.cu / fd_train.cu.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cmath>
#include <algorithm>
#include <string>
#include <unordered_map>
#include <vector>

#include "fd_rows.cuh"

namespace fd {

int set_err(fd_handle* h, int code, const char* fmt, ...) {
    if (h) {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof(buf), fmt, ap);
        va_end(ap);
        h->err = buf;
    }
    return code;
}

int cuda_fail(fd_handle* h, cudaError_t e, const char* what) {
    return set_err(h, FD_ERR_CUDA, "CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
}

// grow the handle-owned workspace (contents are lost when it moves, except the sticky device-side status, which is
// carried over: an earlier asynchronous call may have reported FD_ERR_LIMIT there and nobody has read it yet)
int ensure_ws(fd_handle* h, size_t bytes) {
    if (bytes <= h->ws_bytes) return FD_OK;
    int carried = FD_OK;
    if (h->ws) {
        FD_CUDA(h, cudaDeviceSynchronize());
        FD_CUDA(h, cudaMemcpy(&carried, &reinterpret_cast<Ctrl*>(h->ws)->status, sizeof(int), cudaMemcpyDeviceToHost));
        FD_CUDA(h, cudaFree(h->ws));
        h->ws = nullptr;
        h->ws_bytes = 0;
    }
    bytes = (bytes + (size_t)(1 << 20) - 1) & ~(size_t)((1 << 20) - 1);
    FD_CUDA(h, cudaMalloc((void**)&h->ws, bytes));
    h->ws_bytes = bytes;
    h->feat_rows = -1;
    FD_CUDA(h, cudaMemset(h->ws, 0, 256));
    if (carried != FD_OK)
        FD_CUDA(h, cudaMemcpy(&reinterpret_cast<Ctrl*>(h->ws)->status, &carried, sizeof(int), cudaMemcpyHostToDevice));
    return FD_OK;
}

template <typename T>
static int upload(fd_handle* h, T** dst, const T* src, size_t n);

// The fused kernels never form the tf-idf value of a split feature: x = count * idf[f] is monotone in the integer
// count (idf >= 0), so "x <= threshold" is the same as "count <= c*" with c* = the largest count whose PRODUCT, rounded
// exactly as the device would round it, is still <= threshold.  c* is found here once per (forest, idf) pair.  Any
// negative or NaN idf entry disables the shortcut.
// The quantised forest is laid out two levels per 32-byte record (Model::nodes_s): a walk is bound by the L2 -> L1
// sector traffic of its node fetches (one 32-byte sector per fetch whatever the node size), so a record carries a
// split node AND both its children and one fetch decides two levels.
// features the stop words map to under the current featuriser (hashing trick or vocabulary)
static std::vector<uint32_t> stop_word_features(const fd_handle* h) {
    std::vector<uint32_t> out;
    const Model& m = h->m;
    if (m.num_features <= 0) return out;
    std::unordered_map<std::string, uint32_t> vocab;
    for (size_t i = 0; i < h->host_vocab.size(); ++i)
        vocab.emplace(std::string(h->host_vocab[i].begin(), h->host_vocab[i].end()), (uint32_t)i);
    for (const auto& w : h->host_stop) {
        if (!h->host_vocab.empty()) {
            auto it = vocab.find(std::string(w.begin(), w.end()));
            if (it != vocab.end()) out.push_back(it->second);
        } else {
            out.push_back(nonneg_mod_host(murmur3_host(w.data(), (int)w.size()), m.num_features));
        }
    }
    std::sort(out.begin(), out.end());
    out.erase(std::unique(out.begin(), out.end()), out.end());
    return out;
}

int prepare_fused(fd_handle* h) {
    if (!h->fused_dirty) return FD_OK;
    Model& m = h->m;
    h->fused_dirty = false;
    m.int_walk = 0;
    m.dense_on = 0;
    m.dt_small = 0;
    m.dt_heap_on = 0;
    m.filt_on = 0;
    const std::vector<uint32_t> stop_feats = stop_word_features(h);
    // (1) what the classifier can see, and which of those features a stop word maps to
    const size_t nwords = ((size_t)std::max(m.num_features, 1) + 31) / 32;
    if (m.classifier != 0 && h->filt_want && h->host_seen.size() == nwords) {
        std::vector<uint2> filt(nwords);
        for (size_t i = 0; i < nwords; ++i) filt[i] = make_uint2(h->host_seen[i], 0u);
        for (uint32_t f : stop_feats)
            if ((size_t)(f >> 5) < nwords) filt[f >> 5].y |= filt[f >> 5].x & (1u << (f & 31));
        int rcf = upload(h, &m.filt, filt.data(), filt.size());
        if (rcf) return rcf;
        m.filt_on = 1;
    }
    // (2) forests: thresholds quantised to term counts, two levels per record
    if (m.classifier != 2 || h->host_nodes.empty()) return FD_OK;
    const bool has_idf = !h->host_idf.empty();
    const std::vector<Node>& q = h->host_nodes;
    std::vector<int32_t> cs(q.size(), 0);
    std::vector<char> never(q.size(), 0);  // the node's test can never send a dialogue right (weight 0)
    for (const auto& nd : q) {
        if (nd.feat < 0) continue;
        const double w = has_idf ? h->host_idf[(size_t)nd.feat] : 1.0;
        const double thr = nd.thr;
        if (!(w >= 0.0)) return FD_OK;  // negative or NaN weight: keep the floating-point walk
        long long c;
        if (!(thr >= 0.0)) c = -1;                       // also NaN: "x <= NaN" is never true
        else if (w == 0.0) { c = 0x7FFFFFFF; never[(size_t)(&nd - q.data())] = 1; }  // x is 0.0 whatever the count
        else {
            const double g = std::floor(thr / w);
            c = g >= 2147483646.0 ? 2147483646LL : (long long)g;
            volatile double prod;                        // one IEEE multiply, as __dmul_rn on the device
            for (;;) {
                prod = (double)(c + 1) * w;
                if (c + 1 <= 2147483646LL && prod <= thr) ++c; else break;
            }
            for (;;) {
                prod = (double)c * w;
                if (c >= 0 && prod > thr) --c; else break;
            }
        }
        cs[(size_t)(&nd - q.data())] = (int32_t)c;
    }
    constexpr uint32_t LEAF = 0x7FFFFFFFu;
    std::vector<uint32_t> sec(8 * q.size(), 0u);
    auto is_leaf = [&](int n) { return q[(size_t)n].feat < 0; };
    for (size_t n = 0; n < q.size(); ++n) {
        uint32_t* w = &sec[8 * n];
        if (is_leaf((int)n)) {  // only ever fetched for a tree that is a single leaf: both "children" answer with this node
            w[0] = 0u;
            w[5] = 0x7FFFFFFFu;
            w[1] = w[2] = LEAF;
            w[3] = w[4] = (uint32_t)n;
            continue;
        }
        w[0] = (uint32_t)q[n].feat;
        w[5] = (uint32_t)cs[n];
        for (int side = 0; side < 2; ++side) {
            const int c = q[n].child + side;
            if (is_leaf(c)) {
                w[1 + side] = LEAF;
                w[3 + side] = (uint32_t)c;
                w[6 + side] = 0u;
            } else {
                const int g = q[(size_t)c].child;
                w[1 + side] = (uint32_t)q[(size_t)c].feat | (is_leaf(g) ? 0x80000000u : 0u);
                w[3 + side] = (uint32_t)g | (is_leaf(g + 1) ? 0x80000000u : 0u);
                w[6 + side] = (uint32_t)cs[(size_t)c];
            }
        }
    }
    int rc = upload(h, reinterpret_cast<uint32_t**>(&m.nodes_s), sec.data(), sec.size());
    if (rc) return rc;
    m.int_walk = 1;
    // dense walks (Model::dense_on): forest-local feature ids, byte counts
    if (h->dense_want && m.kind != FD_KIND_DT && m.num_features <= (1 << 22)) {
        std::vector<uint32_t> feats;
        bool fits = true;
        for (size_t n = 0; n < q.size(); ++n) {
            if (q[n].feat < 0) continue;
            feats.push_back((uint32_t)q[n].feat);
            if (!never[n] && cs[n] > 254) fits = false;  // a count saturated at 255 must still decide every test
        }
        std::sort(feats.begin(), feats.end());
        feats.erase(std::unique(feats.begin(), feats.end()), feats.end());
        if (fits && !feats.empty() && feats.size() <= (size_t)DENSE_MAX) {
            std::vector<uint16_t> local((size_t)m.num_features, (uint16_t)0xFFFFu);
            for (size_t i = 0; i < feats.size(); ++i) local[feats[i]] = (uint16_t)i;
            std::vector<uint32_t> secl(sec);
            for (size_t n = 0; n < q.size(); ++n) {
                if (is_leaf((int)n)) continue;
                uint32_t* w = &secl[8 * n];
                w[0] = local[(size_t)q[n].feat];
                if (never[n]) w[5] = 255u;
                for (int side = 0; side < 2; ++side) {
                    const int c = q[n].child + side;
                    if (is_leaf(c)) continue;
                    w[1 + side] = (uint32_t)local[(size_t)q[(size_t)c].feat] | (w[1 + side] & 0x80000000u);
                    if (never[(size_t)c]) w[6 + side] = 255u;
                }
            }
            rc = upload(h, reinterpret_cast<uint32_t**>(&m.nodes_l), secl.data(), secl.size());
            if (!rc) rc = upload(h, &m.f_local, local.data(), local.size());
            if (rc) return rc;
            m.dense_on = 1;
        }
    }
    // decision tree: count only the split features (Model::dt_small)
    if (m.kind == FD_KIND_DT) {
        std::vector<uint32_t> feats;
        for (const auto& nd : q)
            if (nd.feat >= 0 && std::find(feats.begin(), feats.end(), (uint32_t)nd.feat) == feats.end()) feats.push_back((uint32_t)nd.feat);
        bool found = false;
        uint32_t mul = 1u;
        int lg = 0;
        if (feats.size() <= 63) {
            uint64_t x = 0xD1B54A32D192ED03ULL;
            for (lg = 7; lg <= 9 && !found; ++lg) {
                std::vector<uint8_t> used((size_t)1 << lg);
                for (int trial = 0; trial < 200000 && !found; ++trial) {
                    x ^= x >> 12; x ^= x << 25; x ^= x >> 27;  // xorshift64*
                    mul = (uint32_t)((x * 0x2545F4914F6CDD1DULL) >> 32) | 1u;
                    std::fill(used.begin(), used.end(), (uint8_t)0);
                    found = true;
                    for (uint32_t f : feats) {
                        const uint32_t slot = (f * mul) >> (32 - lg);
                        if (used[slot]) {
                            found = false;
                            break;
                        }
                        used[slot] = 1;
                    }
                }
                if (found) break;
            }
        }
        // the main kernel's ballot walk needs the tree in heap order, at most 5 levels of split nodes
        std::vector<int2> heap(32, make_int2(0, 0x7FFFFFFF));
        std::vector<double> outv(32 * 5, 0.0);
        bool shallow = found;
        if (found) {
            std::vector<int> at(63, -1);  // heap position -> node index
            at[0] = h->host_root0;
            for (int p = 0; p < 63 && shallow; ++p) {
                const int n = at[p];
                const bool leaf = is_leaf(n);
                if (p >= 31) {
                    if (!leaf) shallow = false;
                    continue;
                }
                if (leaf) {
                    at[2 * p + 1] = at[2 * p + 2] = n;  // positions below a leaf repeat it ("go left" from an absent node)
                } else {
                    heap[(size_t)p] = make_int2((int)(((uint32_t)q[(size_t)n].feat * mul) >> (32 - lg)), cs[(size_t)n]);
                    at[2 * p + 1] = q[(size_t)n].child;
                    at[2 * p + 2] = q[(size_t)n].child + 1;
                }
            }
            if (shallow) {
                for (int p = 31; p < 63; ++p) {
                    const size_t n = (size_t)at[p];
                    const double2 st = h->host_leaf2[n], pr = h->host_leafp[n];
                    double* o = &outv[(size_t)(p - 31) * 5];
                    o[0] = st.y > st.x ? 1.0 : 0.0;
                    o[1] = pr.x;
                    o[2] = pr.y;
                    o[3] = st.x;
                    o[4] = st.y;
                }
            }
        }
        if (found) {
            std::vector<uint32_t> dtab((size_t)1 << lg, 0xFFFFFFFFu);
            for (uint32_t f : feats)
                dtab[(f * mul) >> (32 - lg)] = f | (std::binary_search(stop_feats.begin(), stop_feats.end(), f) ? 0x80000000u : 0u);
            std::vector<uint32_t> sec_dt = sec;
            for (size_t n = 0; n < q.size(); ++n) {
                uint32_t* w = &sec_dt[8 * n];
                if (is_leaf((int)n)) continue;  // (a single-leaf tree reads counter 0 and goes left whatever it holds)
                w[0] = (w[0] * mul) >> (32 - lg);
                for (int side = 0; side < 2; ++side)
                    if ((w[1 + side] & 0x7FFFFFFFu) != LEAF)
                        w[1 + side] = (w[1 + side] & 0x80000000u) | (((w[1 + side] & 0x7FFFFFFFu) * mul) >> (32 - lg));
            }
            rc = upload(h, &m.dt_tab, dtab.data(), dtab.size());
            if (!rc) rc = upload(h, reinterpret_cast<uint32_t**>(&m.nodes_dt), sec_dt.data(), sec_dt.size());
            if (!rc && shallow) rc = upload(h, &m.dt_heap, heap.data(), heap.size());
            if (!rc && shallow) rc = upload(h, &m.dt_out, outv.data(), outv.size());
            if (rc) return rc;
            m.dt_mul = mul;
            m.dt_log2 = lg;
            m.dt_shift = 32 - lg;
            m.dt_small = 1;
            m.dt_heap_on = shallow ? 1 : 0;
            m.dt_stop = 0;
            for (uint32_t e : dtab)
                if (e != 0xFFFFFFFFu && (e >> 31)) m.dt_stop = 1;
        }
    }
    return FD_OK;
}

template <typename T>
static int upload(fd_handle* h, T** dst, const T* src, size_t n) {
    if (*dst) {
        FD_CUDA(h, cudaFree(*dst));
        *dst = nullptr;
    }
    if (n == 0) return FD_OK;
    FD_CUDA(h, cudaMalloc((void**)dst, n * sizeof(T)));
    FD_CUDA(h, cudaMemcpy(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice));
    return FD_OK;
}

}  // namespace fd

using namespace fd;

extern "C" {

int fd_version(void) { return 100; }

int fd_create(int device, fd_handle** out) {
    if (!out) return FD_ERR_INVALID;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) return FD_ERR_CUDA;
    fd_handle* h = new fd_handle();
    h->device = device;
    if (cudaSetDevice(device) != cudaSuccess) {
        delete h;
        return FD_ERR_CUDA;
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) {
        h->sm_count = prop.multiProcessorCount;
        if (!prop.canMapHostMemory || !prop.unifiedAddressing) h->direct_host = 0;
    }
    const char* dw = getenv("FDB200_DENSE_WALK");
    if (dw && dw[0] == '0') h->dense_want = 0;
    const char* e = getenv("FDB200_DIRECT_HOST");
    if (e && e[0] == '0') h->direct_host = 0;
    *out = h;
    return FD_OK;
}

int fd_destroy(fd_handle* h) {
    if (!h) return FD_OK;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    Model& m = h->m;
    cudaFree(m.sw_tab);
    cudaFree(m.sw_words);
    cudaFree(m.sw_pool);
    cudaFree(m.vocab_tab);
    cudaFree(m.vocab_words);
    cudaFree(m.vocab_pool);
    cudaFree(m.idf);
    cudaFree(m.coef);
    cudaFree(m.filt);
    cudaFree(m.dt_heap);
    cudaFree(m.dt_out);
    cudaFree(m.nodes);
    cudaFree(m.nodes_s);
    cudaFree(m.nodes_l);
    cudaFree(m.f_local);
    cudaFree(m.nodes_dt);
    cudaFree(m.dt_tab);
    cudaFree(m.roots);
    cudaFree(m.leaf2);
    cudaFree(m.leaf_prob);
    cudaFree(m.tree_w);
    cudaFree(h->ws);
    for (int i = 0; i < 3; ++i) {
        if (h->hbuf[i]) cudaFree(h->hbuf[i]);
        if (h->hs[i]) cudaStreamDestroy(h->hs[i]);
        if (h->hev[i]) cudaEventDestroy(h->hev[i]);
    }
    if (h->pinned_ctrl) cudaFreeHost(h->pinned_ctrl);
    delete h;
    return FD_OK;
}

const char* fd_last_error(fd_handle* h) { return h ? h->err.c_str() : "null handle"; }

int64_t fd_launch_count(fd_handle* h) { return h ? h->launches : 0; }

int fd_set_stopwords(fd_handle* h, const uint8_t* words, const int32_t* offs, int32_t n_words, int32_t case_sensitive) {
    if (!h || n_words < 0 || (n_words > 0 && (!words || !offs))) return set_err(h, FD_ERR_INVALID, "bad stop-word arguments");
    FD_CUDA(h, cudaSetDevice(h->device));
    Model& m = h->m;
    m.empty_is_stop = 0;
    m.drop_empty = m.empty_absent;
    h->host_stop.clear();
    h->fused_dirty = true;
    if (n_words == 0) {
        if (m.sw_tab) cudaFree(m.sw_tab);
        if (m.sw_words) cudaFree(m.sw_words);
        if (m.sw_pool) cudaFree(m.sw_pool);
        m.sw_tab = nullptr;
        m.sw_words = nullptr;
        m.sw_pool = nullptr;
        m.sw_cap = 0;
        m.n_sw = 0;
        m.sw_perfect = 0;
        return prepare_fused(h);
    }
    if (n_words > 2048) return set_err(h, FD_ERR_LIMIT, "at most 2048 stop words are supported (got %d)", n_words);
    // word records first (duplicates collapse), then a 2-choice cuckoo table over them: every word sits in one of
    // its two slots (h & mask, (h >> 10) & mask), so a device lookup is two loads and no loop
    std::vector<uint4> wordsv;
    std::vector<uint32_t> hashes;
    std::vector<uint8_t> pool;
    std::vector<std::vector<uint8_t>> seen;
    for (int i = 0; i < n_words; ++i) {
        const int len = offs[i + 1] - offs[i];
        if (len < 0) return set_err(h, FD_ERR_INVALID, "stop-word offsets not ascending");
        std::vector<uint8_t> w(words + offs[i], words + offs[i] + len);
        if (!case_sensitive)
            for (auto& c : w)
                if (c >= 'A' && c <= 'Z') c |= 0x20;
        if (len == 0) {
            m.empty_is_stop = 1;
            m.drop_empty = 1;
            continue;
        }
        bool dup = false;
        for (auto& s0 : seen)
            if (s0 == w) dup = true;
        if (dup) continue;
        seen.push_back(w);
        uint4 s;
        if (len <= 12) {
            uint32_t b[3] = {0, 0, 0};
            for (int k = 0; k < len; ++k) b[k >> 2] |= (uint32_t)w[k] << (8 * (k & 3));
            s = make_uint4((uint32_t)len, b[0], b[1], b[2]);
        } else {
            s = make_uint4((uint32_t)len | 0x80000000u, (uint32_t)pool.size(), 0, 0);
            pool.insert(pool.end(), w.begin(), w.end());
        }
        wordsv.push_back(s);
        hashes.push_back(murmur3_host(w.data(), len));
    }
    // first choice: a perfect hash (one probe on the device) -- look for a multiplier that spreads the words' murmur
    // hashes over 2^11 or 2^12 sixteen-bit slots without any collision; feasible for a few hundred words
    std::vector<uint32_t> tab;
    int cap = 0;
    m.sw_perfect = 0;
    for (int lg = 11; lg <= 12 && !m.sw_perfect && !wordsv.empty() && wordsv.size() <= 400; ++lg) {
        std::vector<uint16_t> t16((size_t)1 << lg);
        uint64_t x = 0x9E3779B97F4A7C15ULL;
        for (int trial = 0; trial < (lg == 11 ? 60000 : 400000) && !m.sw_perfect; ++trial) {
            x ^= x >> 12; x ^= x << 25; x ^= x >> 27;            // xorshift64*
            const uint32_t mul = (uint32_t)((x * 0x2545F4914F6CDD1DULL) >> 32) | 1u;
            std::fill(t16.begin(), t16.end(), (uint16_t)0);
            bool ok = true;
            for (size_t i = 0; i < wordsv.size() && ok; ++i) {
                const uint32_t slot = (hashes[i] * mul) >> (32 - lg);
                ok = t16[slot] == 0;
                t16[slot] = (uint16_t)(i + 1);
            }
            if (ok) {
                m.sw_perfect = 1;
                m.sw_mul = mul;
                m.sw_shift = 32 - lg;
                cap = 1 << lg;
                tab.assign((size_t)cap / 2, 0u);
                memcpy(tab.data(), t16.data(), (size_t)cap * 2);
            }
        }
    }
    if (!m.sw_perfect) {
        cap = 64;
        while (cap < 4 * (int)wordsv.size()) cap <<= 1;
        for (;; cap <<= 1) {
            if (cap > (1 << 16)) return set_err(h, FD_ERR_LIMIT, "could not place the stop words in the lookup table");
            tab.assign(cap, 0u);
            const uint32_t mask = (uint32_t)cap - 1u;
            bool ok = true;
            for (size_t i = 0; i < wordsv.size() && ok; ++i) {
                uint32_t id = (uint32_t)i;
                uint32_t slot = hashes[id] & mask;
                ok = false;
                for (int kick = 0; kick < 512; ++kick) {
                    if (tab[slot] == 0u) {
                        tab[slot] = (hashes[id] & 0xFFFF0000u) | (id + 1u);
                        ok = true;
                        break;
                    }
                    const uint32_t other = (tab[slot] & 0xFFFFu) - 1u;  // evict the resident to its alternate slot
                    tab[slot] = (hashes[id] & 0xFFFF0000u) | (id + 1u);
                    id = other;
                    const uint32_t s1 = hashes[id] & mask, s2 = (hashes[id] >> 10) & mask;
                    slot = (slot == s1) ? s2 : s1;
                }
            }
            if (ok) break;
        }
    }
    int rc = upload(h, &m.sw_tab, tab.data(), tab.size());
    if (rc) return rc;
    rc = upload(h, &m.sw_words, wordsv.data(), wordsv.size());
    if (rc) return rc;
    rc = upload(h, &m.sw_pool, pool.data(), pool.size());
    if (rc) return rc;
    m.sw_cap = wordsv.empty() ? 0 : cap;
    m.n_sw = (int)wordsv.size();
    if (wordsv.empty()) m.sw_perfect = 0;
    h->host_stop = seen;
    return prepare_fused(h);
}

int fd_hash_terms(const uint8_t* terms, const int32_t* offs, int32_t n, int32_t num_features, int32_t* out) {
    if (n < 0 || num_features <= 0 || (n > 0 && (!offs || !out))) return FD_ERR_INVALID;
    for (int i = 0; i < n; ++i) {
        const int len = offs[i + 1] - offs[i];
        if (len < 0) return FD_ERR_INVALID;
        out[i] = (int32_t)nonneg_mod_host(murmur3_host(terms + offs[i], len), num_features);
    }
    return FD_OK;
}

// The IDF vector, the coefficients and the forest are indexed by feature: when the feature space changes they are
// dropped (the compute calls answer FD_ERR_STATE until the IDF stage and the classifier are set again).
static void drop_feature_space(fd_handle* h, int32_t new_num_features, bool meaning_changed) {
    Model& m = h->m;
    if (m.num_features == new_num_features && !meaning_changed) return;
    if (m.idf) cudaFree(m.idf);
    m.idf = nullptr;
    h->host_idf.clear();
    m.classifier = 0;
    m.int_walk = 0;
    m.dt_small = 0;
    m.dt_heap_on = 0;
    m.filt_on = 0;
    h->filt_want = false;
    h->host_seen.clear();
    h->host_nodes.clear();
}

int fd_set_hashing(fd_handle* h, int32_t num_features, int32_t binary) {
    if (!h || num_features <= 0) return set_err(h, FD_ERR_INVALID, "numFeatures must be > 0");
    FD_CUDA(h, cudaSetDevice(h->device));
    Model& m = h->m;
    drop_feature_space(h, num_features, !h->host_vocab.empty());  // (same size but hashed, not looked up: other features)
    m.num_features = num_features;
    m.binary = binary ? 1 : 0;
    h->host_vocab.clear();
    h->fused_dirty = true;
    m.mod_pow2 = (num_features & (num_features - 1)) == 0;
    m.mod_magic = m.mod_pow2 ? 0u : (uint32_t)((1ULL << 32) / (uint64_t)num_features);
    m.mod_r32 = (uint32_t)((1ULL << 32) % (uint64_t)num_features);
    m.empty_idx = nonneg_mod_host(murmur3_host(nullptr, 0), num_features);
    m.csr_list = num_features <= (1 << 22) ? 1 : 0;
    m.bin_mul = num_features > 256 ? (uint32_t)((1ULL << 40) / (uint64_t)num_features) : 0u;
    if (m.vocab_tab) cudaFree(m.vocab_tab);
    if (m.vocab_words) cudaFree(m.vocab_words);
    if (m.vocab_pool) cudaFree(m.vocab_pool);
    m.vocab_tab = nullptr;
    m.vocab_words = nullptr;
    m.vocab_pool = nullptr;
    m.vocab_cap = 0;
    m.empty_absent = 0;
    m.drop_empty = m.empty_is_stop;
    return prepare_fused(h);
}

int fd_set_vocabulary(fd_handle* h, const uint8_t* words, const int32_t* offs, int32_t n_words, int32_t binary) {
    if (!h || n_words <= 0 || !words || !offs) return set_err(h, FD_ERR_INVALID, "bad vocabulary arguments");
    FD_CUDA(h, cudaSetDevice(h->device));
    Model& m = h->m;
    std::vector<uint4> recs((size_t)n_words);
    std::vector<uint32_t> h1((size_t)n_words), h2((size_t)n_words);
    std::vector<uint8_t> pool;
    int empty_id = -1;
    for (int i = 0; i < n_words; ++i) {
        const int len = offs[i + 1] - offs[i];
        if (len < 0) return set_err(h, FD_ERR_INVALID, "vocabulary offsets not ascending");
        const uint8_t* w = words + offs[i];
        uint32_t b[3] = {0, 0, 0};
        for (int k = 0; k < len && k < 12; ++k) b[k >> 2] |= (uint32_t)w[k] << (8 * (k & 3));
        if (len <= 12) {
            recs[i] = make_uint4((uint32_t)len, b[0], b[1], b[2]);
        } else {
            recs[i] = make_uint4((uint32_t)len | 0x80000000u, (uint32_t)pool.size(), 0, 0);
            pool.insert(pool.end(), w, w + len);
        }
        h1[i] = murmur3_host(w, len);
        h2[i] = sw_key((uint32_t)len, b[0], b[1], b[2]);
        if (len == 0) empty_id = i;
    }
    int cap = 64;
    while (cap < 4 * n_words) cap <<= 1;
    std::vector<uint2> tab;
    for (int attempt = 0;; ++attempt, cap <<= 1) {
        if (attempt > 3) return set_err(h, FD_ERR_INVALID, "could not index the vocabulary (duplicate terms?)");
        tab.assign((size_t)cap, make_uint2(0u, 0u));
        const uint32_t mask = (uint32_t)cap - 1u;
        bool ok = true;
        for (int i = 0; i < n_words && ok; ++i) {
            uint32_t id = (uint32_t)i;
            uint32_t slot = h1[id] & mask;
            ok = false;
            for (int kick = 0; kick < 1024; ++kick) {
                if (tab[slot].y == 0u) {
                    tab[slot] = make_uint2(h1[id], id + 1u);
                    ok = true;
                    break;
                }
                const uint32_t other = tab[slot].y - 1u;
                tab[slot] = make_uint2(h1[id], id + 1u);
                id = other;
                const uint32_t s1 = h1[id] & mask, s2 = h2[id] & mask;
                slot = (slot == s1) ? s2 : s1;
            }
        }
        if (ok) break;
    }
    int rc = upload(h, &m.vocab_tab, tab.data(), tab.size());
    if (!rc) rc = upload(h, &m.vocab_words, recs.data(), recs.size());
    if (!rc) rc = upload(h, &m.vocab_pool, pool.data(), pool.size());
    if (rc) return rc;
    m.vocab_cap = cap;
    m.csr_list = 0;  // (vocabulary indices are not evenly spread: the count-table path)
    m.bin_mul = 0;
    drop_feature_space(h, n_words, true);
    m.num_features = n_words;
    m.binary = binary ? 1 : 0;
    m.mod_pow2 = 0;
    m.empty_absent = empty_id < 0;
    m.empty_idx = empty_id < 0 ? 0u : (uint32_t)empty_id;
    m.drop_empty = m.empty_is_stop || m.empty_absent;
    h->host_vocab.assign((size_t)n_words, std::vector<uint8_t>());
    for (int i = 0; i < n_words; ++i) h->host_vocab[(size_t)i].assign(words + offs[i], words + offs[i + 1]);
    h->fused_dirty = true;
    return prepare_fused(h);
}

int fd_set_idf(fd_handle* h, const double* idf, int32_t num_features) {
    if (!h) return FD_ERR_INVALID;
    FD_CUDA(h, cudaSetDevice(h->device));
    h->fused_dirty = true;
    if (!idf) {
        h->host_idf.clear();
        int rc0 = upload(h, &h->m.idf, (const double*)nullptr, 0);
        return rc0 ? rc0 : prepare_fused(h);
    }
    if (num_features != h->m.num_features)
        return set_err(h, FD_ERR_INVALID, "idf size %d != numFeatures %d", num_features, h->m.num_features);
    h->host_idf.assign(idf, idf + num_features);
    int rc = upload(h, &h->m.idf, idf, (size_t)num_features);
    return rc ? rc : prepare_fused(h);  // model upload is where allocation belongs, not the compute calls
}

int fd_set_linear(fd_handle* h, const double* coef_dense, int32_t num_features, double intercept, double threshold) {
    if (!h || !coef_dense) return set_err(h, FD_ERR_INVALID, "null coefficients");
    if (num_features != h->m.num_features)
        return set_err(h, FD_ERR_INVALID, "coefficient size %d != numFeatures %d", num_features, h->m.num_features);
    FD_CUDA(h, cudaSetDevice(h->device));
    int rc = upload(h, &h->m.coef, coef_dense, (size_t)num_features);
    if (rc) return rc;
    {
        std::vector<uint32_t> nz(((size_t)num_features + 31) / 32, 0u);
        size_t kept = 0;
        for (int32_t i = 0; i < num_features; ++i)
            if (coef_dense[i] != 0.0) {  // (NaN counts as non-zero and is kept)
                nz[(size_t)i >> 5] |= 1u << (i & 31);
                ++kept;
            }
        h->host_seen = nz;
        h->m.lr_keep = (float)kept / (float)num_features;
        h->filt_want = kept * 10 <= (size_t)num_features * 9;  // worth a bitmap probe per term from 10 % zeros
    }
    h->m.intercept = intercept;
    // LogisticRegressionModel.raw2prediction: rawThreshold = log(t / (1 - t)), -inf at 0, +inf at 1
    h->m.raw_threshold = threshold <= 0.0 ? -INFINITY : (threshold >= 1.0 ? INFINITY : std::log(threshold / (1.0 - threshold)));
    h->m.classifier = 1;
    h->m.n_classes = 2;
    h->host_nodes.clear();
    h->fused_dirty = true;
    return prepare_fused(h);
}

int fd_set_forest(fd_handle* h, int32_t kind, int32_t n_trees, const int32_t* tree_offsets, const int32_t* feature,
                  const double* threshold, const int32_t* left, const int32_t* right, const double* leaf_value,
                  const double* tree_weights, int32_t n_classes) {
    if (!h) return FD_ERR_INVALID;
    if (kind != FD_KIND_DT && kind != FD_KIND_RF && kind != FD_KIND_GBT) return set_err(h, FD_ERR_INVALID, "unknown forest kind %d", kind);
    if (n_trees <= 0 || !tree_offsets || !feature || !threshold || !left || !right || !leaf_value)
        return set_err(h, FD_ERR_INVALID, "null forest arrays");
    if (n_classes != 2) return set_err(h, FD_ERR_INVALID, "binary classification only on this path (numClasses = %d)", n_classes);
    if (kind == FD_KIND_DT && n_trees != 1) return set_err(h, FD_ERR_INVALID, "a decision tree model has one tree");
    if (kind == FD_KIND_GBT && !tree_weights) return set_err(h, FD_ERR_INVALID, "GBT needs tree weights");
    FD_CUDA(h, cudaSetDevice(h->device));
    const int total = tree_offsets[n_trees];
    std::vector<Node> nodes((size_t)total);
    std::vector<double2> leaf2((size_t)total, make_double2(0.0, 0.0));
    std::vector<double2> leafp((size_t)total, make_double2(0.0, 0.0));
    std::vector<int32_t> roots((size_t)n_trees);
    std::vector<std::pair<int, int>> queue;
    for (int t = 0; t < n_trees; ++t) {
        const int base = tree_offsets[t], cnt = tree_offsets[t + 1] - base;
        if (cnt <= 0) return set_err(h, FD_ERR_INVALID, "tree %d has no nodes", t);
        roots[t] = base;
        int next = base + 1;
        queue.clear();
        queue.push_back({0, base});
        for (size_t qi = 0; qi < queue.size(); ++qi) {
            const int o = queue[qi].first, n = queue[qi].second;
            const int go = base + o;
            if (feature[go] < 0) {
                nodes[n].feat = -1;
                nodes[n].child = n;
                nodes[n].thr = kind == FD_KIND_GBT ? leaf_value[go] : 0.0;
                if (kind != FD_KIND_GBT) {
                    const double c0 = leaf_value[2 * (size_t)go], c1 = leaf_value[2 * (size_t)go + 1];
                    if (kind == FD_KIND_DT) {
                        leaf2[n] = make_double2(c0, c1);
                        const double tot = c0 + c1;  // normalizeToProbabilitiesInPlace
                        leafp[n] = tot != 0.0 ? make_double2(c0 / tot, c1 / tot) : make_double2(c0, c1);
                    } else {  // votes(i) += classCounts(i) / total, skipped when total == 0
                        const double tot = c0 + c1;
                        leaf2[n] = tot != 0.0 ? make_double2(c0 / tot, c1 / tot) : make_double2(0.0, 0.0);
                    }
                }
            } else {
                if (feature[go] >= h->m.num_features && h->m.num_features > 0)
                    return set_err(h, FD_ERR_INVALID, "tree %d node %d splits on feature %d >= numFeatures", t, o, feature[go]);
                const int l = left[go], r = right[go];
                if (l <= 0 || r <= 0 || l >= cnt || r >= cnt || next + 2 > base + cnt)
                    return set_err(h, FD_ERR_INVALID, "tree %d node %d has bad children", t, o);
                nodes[n].feat = feature[go];
                nodes[n].thr = threshold[go];
                nodes[n].child = next;
                queue.push_back({l, next});
                queue.push_back({r, next + 1});
                next += 2;
            }
        }
        if (next != base + cnt) return set_err(h, FD_ERR_INVALID, "tree %d is not a connected binary tree", t);
    }
    Model& m = h->m;
    int rc = upload(h, &m.nodes, nodes.data(), nodes.size());
    if (!rc) rc = upload(h, &m.leaf2, leaf2.data(), leaf2.size());
    if (!rc) rc = upload(h, &m.leaf_prob, leafp.data(), kind == FD_KIND_DT ? leafp.size() : 0);
    if (!rc) rc = upload(h, &m.roots, roots.data(), roots.size());
    if (!rc && kind == FD_KIND_GBT) rc = upload(h, &m.tree_w, tree_weights, (size_t)n_trees);
    if (rc) return rc;
    {
        std::vector<uint32_t> used(((size_t)std::max(h->m.num_features, 1) + 31) / 32, 0u);
        size_t n_used = 0;
        for (const auto& nd : nodes)
            if (nd.feat >= 0 && (size_t)nd.feat / 32 < used.size() && !((used[(size_t)nd.feat >> 5] >> (nd.feat & 31)) & 1u)) {
                used[(size_t)nd.feat >> 5] |= 1u << (nd.feat & 31);
                ++n_used;
            }
        h->host_seen = used;
        h->filt_want = h->m.num_features > 0 && n_used * 2 <= (size_t)h->m.num_features;
    }
    h->host_leaf2 = leaf2;
    h->host_leafp = leafp;
    h->host_root0 = roots[0];
    h->host_nodes = nodes;
    h->fused_dirty = true;
    m.kind = kind;
    m.n_trees = n_trees;
    m.n_classes = 2;
    m.n_nodes = total;
    m.classifier = 2;
    return prepare_fused(h);
}

}  // extern "C"
