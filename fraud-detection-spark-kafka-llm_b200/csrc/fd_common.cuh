// fd_common.cuh -- handle layout, error plumbing and small device helpers shared by the kernels.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/fdb200.h"

namespace fd {

// ---- on-device model pieces -------------------------------------------------------------------
// One decision-tree node, 16 bytes so a visit is ONE 128-bit load.  Nodes of a forest are laid out
// breadth-first per tree with the two children of a node adjacent (right = left + 1).
//   internal: feat >= 0, thr = split threshold, child = absolute index of the left child
//   leaf    : feat <  0, thr = leaf prediction (GBT), child = absolute node index (row of leaf2[])
struct __align__(16) Node {
    double thr;
    int32_t feat;
    int32_t child;
};

constexpr uint32_t EMPTY_KEY = 0xFFFFFFFFu;

// control block at the start of the workspace (device memory, written by kernels)
constexpr int DENSE_MAX = 2048;          // forest-local feature ids of the dense walks (Model::dense_on): bytes of the vector
constexpr size_t CTRL_CLEAR_BYTES = 24;  // tmp_cursor, big_count, big2_count, tmp_overflow, next_item
struct Ctrl {
    unsigned long long tmp_cursor;  // entries allocated in the CSR staging buffer
    unsigned int big_count;         // dialogues deferred to the medium wide-table kernel (list grows from the front)
    unsigned int big2_count;        // dialogues deferred to the large wide-table kernel (list grows from the back)
    unsigned int tmp_overflow;      // staging buffer was too small
    unsigned int next_item;         // micro-batches: dialogues handed out after every warp's first (RowArgs::dyn_rows)
    int status;                     // sticky device-side status (FD_OK / FD_ERR_LIMIT); NOT cleared per call
    unsigned int pad[1];
};

struct Model {
    __host__ __device__ size_t sw_tab_bytes() const { return sw_perfect ? (size_t)sw_cap * 2 : (size_t)sw_cap * 4; }
    __host__ __device__ size_t sw_smem_bytes() const { return sw_tab_bytes() + (size_t)n_sw * 16; }
    // stop words: 2-choice cuckoo table of 4-byte slots keyed by murmur3(word, 42): a word sits at slot
    //   h & mask or (h >> 10) & mask; slot = (hash & 0xFFFF0000) | (word id + 1), 0 = empty
    uint32_t* sw_tab = nullptr;
    // perfect-hash mode (found on the host for lists of a few hundred words): ONE probe -- 16-bit slots holding
    // word id + 1 at index (h * sw_mul) >> sw_shift, no two words share a slot; otherwise the cuckoo layout above
    int sw_perfect = 0;
    uint32_t sw_mul = 0;
    int sw_shift = 0;
    int sw_cap = 0;             // slots, power of two, 0 = no stop words
    uint4* sw_words = nullptr;  // per word: x = len (|0x80000000 when > 12 bytes), y/z/w = first 12 bytes or y = pool offset
    int n_sw = 0;
    uint8_t* sw_pool = nullptr; // bytes of words longer than 12
    int empty_is_stop = 0;
    int empty_absent = 0;       // vocabulary mode: the empty token is not in the vocabulary
    int drop_empty = 0;         // empty_is_stop || empty_absent: empty tokens are not counted
    // CountVectorizerModel vocabulary (replaces the hashing trick when vocab_cap > 0): 2-choice cuckoo table in global
    // memory, slot = {murmur3 of the word, word id + 1}; a word sits at slot (h & mask) or (sw_key(...) & mask)
    uint2* vocab_tab = nullptr;
    int vocab_cap = 0;
    uint4* vocab_words = nullptr;  // per word: x = len (|0x80000000 when > 12 bytes), y/z/w = first 12 bytes or y = pool offset
    uint8_t* vocab_pool = nullptr;
    // hashing
    int32_t num_features = 0;
    int binary = 0;
    uint32_t empty_idx = 0;  // bucket of the empty token
    uint32_t mod_magic = 0;  // floor(2^32 / num_features)
    uint32_t mod_r32 = 0;    // 2^32 mod num_features
    int mod_pow2 = 0;
    // featurize through a token list ordered by a counting sort (fd_main.cu csr_sorted_emit): hashing trick with
    // numFeatures <= 2^22; bin_mul = floor(2^40 / numFeatures) maps a bucket to one of 256 bins (0: the bucket itself)
    int csr_list = 0;
    uint32_t bin_mul = 0;
    // idf
    double* idf = nullptr;
    // classifier
    int classifier = 0;  // 0 none, 1 linear, 2 forest
    double* coef = nullptr;
    // What the classifier can see.  A LogisticRegression term with an exact-zero coefficient adds +-0.0 to a sum that
    // starts at +0.0 (L1 / elastic-net fits; 59 % of the shipped model's coefficients), and a forest only ever looks at
    // the features its nodes split on: every other term is dropped before it reaches the count table.  The same bitmap
    // makes the stop-word lookup unnecessary for almost every token: a stop word only matters when its bucket is one
    // the classifier can see, so the fused kernels probe the stop-word table only for tokens whose bucket carries the
    // second bit.  Word i of filt covers features 32 i .. 32 i + 31: x = "visible to the classifier",
    // y = "visible AND some stop word maps to this feature" (rebuilt by prepare_fused from the host copies below).
    uint2* filt = nullptr;
    int filt_on = 0;         // consult filt (off when the classifier sees most features anyway)
    float lr_keep = 1.0f;    // fraction of the features a sparse LogisticRegression keeps
    double intercept = 0.0, raw_threshold = 0.0;
    int kind = 0, n_trees = 0, n_classes = 2;
    Node* nodes = nullptr;
    // same forest for the fused kernels: split thresholds pre-quantised to TERM COUNTS and two levels per 32-byte
    // record (see prepare_fused): record n = { X.feat, L.feat, R.feat, L.child, R.child, X.c*, L.c*, R.c* } for the
    // split node n = X and its children L, R.  A child that is a leaf has feat = 0x7FFFFFFF and its own node id in the
    // child word; bit 31 of L.feat / L.child says that L's left / right child is a leaf (same for R).
    uint4* nodes_s = nullptr;
    int int_walk = 0;          // 1: the fused kernels walk nodes_s with integer compares
    // forests over at most DENSE_MAX distinct split features whose count thresholds fit a byte: the end of a dialogue
    // spreads its counts over a dense byte vector indexed by forest-local feature id (f_local: feature -> local id,
    // 0xFFFF = not a split feature) and a node visit is one byte load; nodes_l = nodes_s with local ids
    uint4* nodes_l = nullptr;
    uint16_t* f_local = nullptr;
    int dense_on = 0;
    // A decision tree splits on a handful of features (at most 2^depth - 1), and its prediction depends on the counts
    // of those features only.  When there are at most 63 of them the fused DT kernel keeps just those counts: a perfect
    // hash (f * dt_mul) >> dt_shift over 2^dt_log2 slots (dt_tab[slot] = the feature, ~0 = none) indexes per-warp
    // counters, every other term is dropped after one compare, and nodes_dt is nodes_s with slots in place of features.
    int dt_small = 0;
    uint32_t dt_mul = 0;
    int dt_shift = 0, dt_log2 = 0;
    uint32_t* dt_tab = nullptr;  // bit 31 of an entry: some stop word maps to this feature (probe the stop-word table)
    uint4* nodes_dt = nullptr;
    // ... and when the tree is at most 5 levels deep the main kernel evaluates all of its <= 31 split nodes at once, one
    // per lane, in heap order (root 0, children 2 i + 1 and 2 i + 2; a position below a leaf repeats "go left"), and
    // walks the ballot: dt_heap[i] = {counter slot, count threshold}, dt_out[p] = the five result doubles
    // (prediction, probability 0/1, rawPrediction 0/1) of the leaf reached at depth-5 position p (0..31).
    int dt_heap_on = 0;
    int dt_stop = 0;           // some split feature is also the bucket of a stop word (bit 31 set somewhere in dt_tab)
    int2* dt_heap = nullptr;
    double* dt_out = nullptr;
    int32_t* roots = nullptr;  // absolute root index per tree
    double2* leaf2 = nullptr;  // DT: raw class stats, RF: stats / total (0,0 when total == 0)
    double2* leaf_prob = nullptr;  // DT: stats / total, divided once on the host (IEEE division: same bits as on the device)
    double* tree_w = nullptr;
    int n_nodes = 0;
};

}  // namespace fd

struct fd_handle {
    int device = 0;
    int sm_count = 148;
    std::string err;
    fd::Model m;
    // workspace
    char* ws = nullptr;
    size_t ws_bytes = 0;
    // featurize state between begin/finish
    int64_t feat_rows = -1;
    int64_t feat_nnz = 0;
    int64_t feat_unsorted = 0;
    size_t off_nnz = 0, off_tmpoff = 0, off_tmp = 0, off_bsum = 0;
    int64_t tmp_cap = 0;
    int64_t launches = 0;
    // host-transform staging
    cudaStream_t hs[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t hev[3] = {nullptr, nullptr, nullptr};
    char* hbuf[3] = {nullptr, nullptr, nullptr};
    size_t hbuf_bytes[3] = {0, 0, 0};
    fd::Ctrl* pinned_ctrl = nullptr;
    int dense_want = 1;   // forests may use the dense byte-count walk (FDB200_DENSE_WALK=0 keeps the lookup front)
    int direct_host = 1;  // fd_transform_host may read / write pinned caller buffers in place (FDB200_DIRECT_HOST=0 turns it off)
    // host copies needed to (re)build nodes_s when the idf vector or the forest changes
    std::vector<fd::Node> host_nodes;
    std::vector<double> host_idf;
    std::vector<double2> host_leaf2, host_leafp;     // per node: leaf statistics / probabilities (Model::leaf2, leaf_prob)
    int host_root0 = 0;                              // node index of the first tree's root
    bool filt_want = false;                          // the classifier sees few enough features for Model::filt to pay
    std::vector<uint32_t> host_seen;                 // bitmap of the features the classifier can see (non-zero coefficient / split feature)
    std::vector<std::vector<uint8_t>> host_stop;     // stop words as uploaded (lower-cased unless case sensitive), for the filt.y bits
    std::vector<std::vector<uint8_t>> host_vocab;    // CountVectorizerModel vocabulary (empty: hashing trick)
    bool fused_dirty = false;
    // launch configuration of the main kernel, cached per (instantiation, shared-memory size)
    struct LaunchCfg { const void* fn; size_t smem; int occ; };
    std::vector<LaunchCfg> launch_cfg;
};

namespace fd {

int set_err(fd_handle* h, int code, const char* fmt, ...);
int cuda_fail(fd_handle* h, cudaError_t e, const char* what);
int ensure_ws(fd_handle* h, size_t bytes);
int prepare_fused(fd_handle* h);

#define FD_CUDA(h, call)                                             \
    do {                                                             \
        cudaError_t e__ = (call);                                    \
        if (e__ != cudaSuccess) return fd::cuda_fail(h, e__, #call); \
    } while (0)

// ---- murmur3 x86_32 pieces (host + device) ------------------------------------------------------
__host__ __device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
__host__ __device__ __forceinline__ uint32_t mm_k(uint32_t k) {
    k *= 0xcc9e2d51u;
    k = rotl32(k, 15);
    return k * 0x1b873593u;
}
__host__ __device__ __forceinline__ uint32_t mm_h(uint32_t h, uint32_t k) {
    h ^= k;
    h = rotl32(h, 13);
    return h * 5u + 0xe6546b64u;
}
__host__ __device__ __forceinline__ uint32_t mm_fmix(uint32_t h, uint32_t len) {
    h ^= len;
    h ^= h >> 16;
    h *= 0x85ebca6bu;
    h ^= h >> 13;
    h *= 0xc2b2ae35u;
    h ^= h >> 16;
    return h;
}
inline uint32_t murmur3_host(const uint8_t* p, int n, uint32_t seed = 42u) {
    uint32_t h = seed;
    int nb = n / 4;
    for (int i = 0; i < nb; ++i) {
        uint32_t k = (uint32_t)p[4 * i] | ((uint32_t)p[4 * i + 1] << 8) | ((uint32_t)p[4 * i + 2] << 16) |
                     ((uint32_t)p[4 * i + 3] << 24);
        h = mm_h(h, mm_k(k));
    }
    uint32_t k = 0;
    for (int i = 0; i < (n & 3); ++i) k |= (uint32_t)p[4 * nb + i] << (8 * i);
    if (n & 3) h ^= mm_k(k);
    return mm_fmix(h, (uint32_t)n);
}
// second hash of a term for the vocabulary's cuckoo table: a cheap mix of its length and first 12 bytes (zero padded);
// host and device must agree
__host__ __device__ __forceinline__ uint32_t sw_key(uint32_t n, uint32_t b0, uint32_t b1, uint32_t b2) {
    uint32_t k = (b0 * 0x9E3779B1u) ^ (b1 * 0x85EBCA6Bu) ^ (b2 * 0xC2B2AE35u) ^ (n * 0x27D4EB2Fu);
    return k ^ (k >> 15);
}
// Utils.nonNegativeMod on the signed value of h
inline uint32_t nonneg_mod_host(uint32_t h, int32_t nf) {
    int32_t r = (int32_t)h % nf;
    return (uint32_t)(r < 0 ? r + nf : r);
}

}  // namespace fd
