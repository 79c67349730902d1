/*
 * fdb200.h -- C-ABI of libfdb200.so: the B200 (sm_100a) hot path behind a Spark ML PipelineModel.
 *
 * The reference (wangwang2111/fraud-detection-spark-kafka-llm) has NO C/FFI interface of its own for
 * this path: it calls pyspark.ml (third-party, JVM) -- PipelineModel.load (utils/agent_api.py:129),
 * model.transform (utils/agent_api.py:158, fraud_detection_spark.py:98), Pipeline.fit
 * (fraud_detection_spark.py:85-91).  Each entry point below therefore cites the reference call
 * site / Spark stage whose work it replaces.  Host code stays Python and binds these with ctypes
 * (see INTEGRATION.md); no torch types appear in any signature.
 *
 * Conventions
 *  - every function returns int: 0 = FD_OK, < 0 = error code; fd_last_error(h) gives the text
 *    (per handle, valid until the next call on that handle).
 *  - "dev" pointers are device memory owned by the caller (e.g. a torch tensor's data_ptr()),
 *    "host" pointers are host memory.  The library never frees or keeps caller buffers past the
 *    completion of the stream work it enqueued.
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  Compute calls are
 *    asynchronous on that stream unless stated otherwise.
 *  - a handle is bound to one device, owns a growable device workspace, and is not re-entrant;
 *    distinct handles are independent.  The compute calls of ONE handle share that workspace
 *    (fd_transform*, fd_featurize_*, fd_vocab_count, fd_doc_freq): issue them on one stream, or
 *    order the streams yourself, so that two of them never run at the same time.
 *  - packed text: `text` = all dialogues' UTF-8 bytes back to back, `offs` = int64[n_rows+1] byte
 *    offsets (offs[0] may be > 0).  `text` must be 16-byte aligned and readable up to the next
 *    multiple of 16 past offs[n_rows] (the kernels stage 16-byte-aligned windows with TMA bulk
 *    copies).
 */
#ifndef FDB200_H
#define FDB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FD_OK 0
#define FD_ERR_INVALID (-1)  /* bad argument */
#define FD_ERR_CAPACITY (-2) /* caller buffer too small; required size reported */
#define FD_ERR_CUDA (-3)     /* CUDA runtime error, text in fd_last_error */
#define FD_ERR_STATE (-4)    /* a model piece needed by the call has not been set */
#define FD_ERR_LIMIT (-5)    /* a dialogue exceeds a documented size limit */

/* flags for fd_featurize / fd_transform */
#define FD_FLAG_CLEAN 1u /* apply the caller-side cleaning lower + regexp_replace("[^a-zA-Z ]","")
                            (utils/agent_api.py:143-144) inside the kernel; without it the input is
                            the `clean_text` column and only the Tokenizer's ASCII lower-casing and
                            Java "\\s" split are applied */

/* classifier kinds for fd_set_forest */
#define FD_KIND_DT 1  /* DecisionTreeClassificationModel  (fraud_detection_spark.py:59-65) */
#define FD_KIND_RF 2  /* RandomForestClassificationModel  (fraud_detection_spark.py:67-74) */
#define FD_KIND_GBT 3 /* GBTClassificationModel           (BASELINE.json config 4) */

typedef struct fd_handle fd_handle;

/* ---- lifetime ------------------------------------------------------------------------------- */
/* Replaces SparkSession.builder...getOrCreate() + PipelineModel.load's JVM-side object
 * (utils/agent_api.py:126-129): creates the per-device context that holds the uploaded model. */
int fd_create(int device, fd_handle** out);
int fd_destroy(fd_handle* h);
const char* fd_last_error(fd_handle* h);
int fd_version(void);

/* ---- model upload (host pointers, synchronous) ---------------------------------------------- */
/* Order: featuriser first (stop words, then HashingTF or the vocabulary), IDF, classifier last -- as the stages of the
 * saved pipeline come.  The IDF vector, the coefficients and the forest are indexed by feature, and the uploads derive
 * per-feature tables from them (quantised thresholds, the bitmap of features the classifier can see and which of those a
 * stop word maps to): a call that changes the feature space (fd_set_hashing with another numFeatures, any
 * fd_set_vocabulary, or a switch between the two) DROPS the IDF vector and the classifier -- fd_transform / fd_score then
 * answer FD_ERR_STATE until they are set again.  Changing only the stop words or the binary flag keeps them. */
/* StopWordsRemover params (stages/1_StopWordsRemover_8c0b00b256b3/metadata/part-00000:1).
 * words = UTF-8 bytes back to back, offs = int32[n_words+1].  With case_sensitive == 0 ASCII
 * letters of the list are folded to lower case (the host folds non-ASCII ones before the call). */
int fd_set_stopwords(fd_handle* h, const uint8_t* words, const int32_t* offs, int32_t n_words,
                     int32_t case_sensitive);
/* HashingTF params (stages/2_HashingTF_e7eba1072633/metadata/part-00000:1): numFeatures, binary. */
int fd_set_hashing(fd_handle* h, int32_t num_features, int32_t binary);
/* CountVectorizerModel (fraud_detection_spark.py:52; data = `vocabulary: array<string>`): use a term dictionary instead
 * of the hashing trick.  words = the vocabulary's UTF-8 bytes back to back in index order, offs int32[n_words+1].
 * Terms outside the vocabulary are ignored, numFeatures becomes n_words; only minTF = 1.0 semantics are implemented.
 * A later fd_set_hashing switches back to HashingTF. */
int fd_set_vocabulary(fd_handle* h, const uint8_t* words, const int32_t* offs, int32_t n_words, int32_t binary);
/* CountVectorizer.fit, counting part (fraud_detection_spark.py:52 via Pipeline.fit at :85-91): tokenise the dialogues
 * exactly as fd_featurize does (cleaning flag, stop words) and accumulate every surviving term into a caller-owned
 * open-addressed table in device memory, so that several calls (batches, or ranks merged later) add up:
 *   keys  u64[1 << log2_slots]   initialised to all ones (empty); terms of <= 7 bytes are stored as their bytes
 *                                (length in the top byte), longer ones as 0xFF + a 56-bit hash
 *   info  u32[1 << log2_slots][4] initialised to zero: {term count, document frequency, pool offset, length}
 *   pool  u8[pool_bytes]          bytes of the terms longer than 7 bytes (offset 0xFFFFFFFF: not recorded)
 *   counters u64[3]               initialised to zero: {entries, pool bytes used, overflow flag}
 * The caller picks the vocabulary (top vocabSize by count, minDF / maxDF) from the table.  Asynchronous. */
int fd_vocab_count(fd_handle* h, const uint8_t* text_dev, const int64_t* offs_dev, int64_t n_rows, uint32_t flags,
                   uint64_t* keys_dev, uint32_t* info_dev, int32_t log2_slots, uint8_t* pool_dev, int64_t pool_bytes,
                   uint64_t* counters_dev, void* stream);
/* HashingTF.indexOf(term) for n terms on the HOST (no device work): out[i] = nonNegativeMod(
 * murmur3_x86_32(term_i bytes, seed 42), num_features).  terms = bytes back to back, offs int32[n+1]. */
int fd_hash_terms(const uint8_t* terms, const int32_t* offs, int32_t n, int32_t num_features, int32_t* out);
/* IDF.fit, counting part (fraud_detection_spark.py:53 via Pipeline.fit at :85-91; mllib IDF.DocumentFrequencyAggregator):
 * out_doc_freq int64[num_features] (zeroed by the call) receives, per feature, the number of rows of the CSR matrix that
 * hold it with a value > 0.  Ranks add their vectors (NCCL all-reduce by the caller) and the host applies
 * idf = log((numDocs + 1) / (docFreq + 1)).  Asynchronous. */
int fd_doc_freq(fd_handle* h, const int32_t* idx_dev, const double* val_dev, int64_t nnz, int32_t num_features,
                int64_t* out_doc_freq_dev, void* stream);
/* IDFModel.transform on a CSR matrix that is already on the device (the step of Pipeline.fit between IDF.fit and the
 * classifier's fit, fraud_detection_spark.py:85-91): val[k] *= idf_dev[idx[k]], one exact fp64 multiply, in place.
 * idf_dev: f64[num_features] on the device.  Asynchronous. */
int fd_idf_scale(fd_handle* h, const int32_t* idx_dev, double* val_dev, int64_t nnz, const double* idf_dev, int32_t num_features,
                 void* stream);
/* IDFModel.idf (stages/3_IDF_58bd96296a82/data): dense f64[num_features]; NULL removes the stage. */
int fd_set_idf(fd_handle* h, const double* idf, int32_t num_features);
/* Binomial LogisticRegressionModel (stages/4_LogisticRegression_65ef69d6dfd9): dense coefficient
 * row f64[num_features] (zeros where the saved sparse matrix has none), intercept, threshold. */
int fd_set_linear(fd_handle* h, const double* coef_dense, int32_t num_features, double intercept,
                  double threshold);
/* Tree ensemble as SoA node arrays (Spark NodeData order: per tree pre-order ids, root first).
 * tree_offsets int32[n_trees+1] indexes the concatenated node arrays; feature[n] < 0 marks a leaf;
 * left/right are node ids LOCAL to the tree.  leaf_value: f64[n_nodes, n_classes] class-weight stats
 * for FD_KIND_DT / FD_KIND_RF, f64[n_nodes] predictions for FD_KIND_GBT.  tree_weights f64[n_trees]
 * (used by GBT only).  Replaces {DecisionTree,RandomForest,GBT}ClassificationModel.transform
 * (fraud_detection_spark.py:98 on the models built at :59-74). */
int fd_set_forest(fd_handle* h, int32_t kind, int32_t n_trees, const int32_t* tree_offsets,
                  const int32_t* feature, const double* threshold, const int32_t* left,
                  const int32_t* right, const double* leaf_value, const double* tree_weights,
                  int32_t n_classes);

/* ---- featurise: Tokenizer -> StopWordsRemover -> HashingTF -> IDF  (fraud_detection_spark.py:50-53;
 *      saved stages 0..3) -------------------------------------------------------------------- */
/* Pass 1: tokenise/hash/count every dialogue into the handle's workspace and report the total number
 * of non-zeros through *nnz_host (this call synchronises `stream` to read it back). */
int fd_featurize_begin(fd_handle* h, const uint8_t* text_dev, const int64_t* offs_dev, int64_t n_rows,
                       uint32_t flags, int64_t* nnz_host, void* stream);
/* Pass 2: emit CSR for the rows of the last fd_featurize_begin: rowptr int64[n_rows+1],
 * idx int32[nnz] ascending per row, val f64[nnz] (term count x idf).  capacity = elements available in
 * out_idx/out_val; FD_ERR_CAPACITY if smaller than nnz. */
int fd_featurize_finish(fd_handle* h, int64_t* out_rowptr_dev, int32_t* out_idx_dev, double* out_val_dev,
                        int64_t capacity, void* stream);
/* Both passes in one call; *nnz_host receives the total (also on FD_ERR_CAPACITY). */
int fd_featurize(fd_handle* h, const uint8_t* text_dev, const int64_t* offs_dev, int64_t n_rows,
                 uint32_t flags, int64_t* out_rowptr_dev, int32_t* out_idx_dev, double* out_val_dev,
                 int64_t capacity, int64_t* nnz_host, void* stream);

/* ---- score: classifier stage on CSR features  (ProbabilisticClassificationModel.transform;
 *      result columns read at utils/agent_api.py:161-167, fraud_detection_spark.py:101-118) ---- */
/* out_pred f64[n_rows], out_prob f64[n_rows, n_classes], out_raw f64[n_rows, n_classes]. */
int fd_score(fd_handle* h, const int64_t* rowptr_dev, const int32_t* idx_dev, const double* val_dev,
             int64_t n_rows, double* out_pred_dev, double* out_prob_dev, double* out_raw_dev, void* stream);

/* ---- transform: the whole PipelineModel.transform, fused (utils/agent_api.py:158) ----------- */
/* text -> prediction / probability / rawPrediction without materialising the sparse vectors. */
int fd_transform(fd_handle* h, const uint8_t* text_dev, const int64_t* offs_dev, int64_t n_rows,
                 uint32_t flags, double* out_pred_dev, double* out_prob_dev, double* out_raw_dev,
                 void* stream);
/* fd_transform / fd_featurize_* are asynchronous; a dialogue over the size limit (more than ~49k
 * distinct hashed terms) is only detected on the device.  fd_status synchronises `stream` and returns
 * (and clears) that sticky status: FD_OK or FD_ERR_LIMIT. */
int fd_status(fd_handle* h, void* stream);
/* Same with HOST buffers (pinned memory recommended): the call chunks the batch, overlaps the
 * host->device copies, the kernels and the device->host copies on internal streams, and returns when
 * the three result columns are complete in host memory.  text_host needs no padding or alignment.
 * A micro-batch (at most 65,536 dialogues and 8 MiB of text) whose five buffers are all pinned (page-locked, mapped)
 * host memory, text_host 16-byte aligned, is not copied: the kernels read the text and offsets and write the columns in
 * place over PCIe -- same results, fewer driver calls on the latency path (utils/agent_api.py:158 one dialogue at a time,
 * app_ui.py:195-226 one poll at a time).  Environment FDB200_DIRECT_HOST=0 (read at fd_create) turns that off. */
int fd_transform_host(fd_handle* h, const uint8_t* text_host, const int64_t* offs_host, int64_t n_rows,
                      uint32_t flags, double* out_pred_host, double* out_prob_host, double* out_raw_host);
/* No counterpart in the reference (nor in Spark): the fused transform for callers that cannot allocate device
 * memory -- a host whose GPU memory is already owned by a pool allocator (an inference server, a Spark/RAPIDS
 * executor), or a latency-bound micro-batch consumer that must not take the driver's allocation lock or hit
 * an out-of-memory error mid-stream.  Same inputs, outputs and results as fd_transform, but every byte of
 * scratch lives in a fixed array inside the library image (about 4.5 MiB per GPU, held until the process
 * exits), so the call performs no cudaMalloc / cudaFree.  Limits and rules:
 *   - n_rows <= FD_NOALLOC_MAX_ROWS, else FD_ERR_LIMIT and nothing is launched;
 *   - all calls on one device share the array: calls whose GPU work overlaps race on it -- serialise them;
 *   - the first call on a device resolves the array's address (once per device): make it before starting any
 *     stream capture;
 *   - asynchronous on `stream`; out_status_dev (int32 in device memory, may be NULL) receives FD_OK or
 *     FD_ERR_LIMIT (a dialogue with more than ~49k distinct hashed terms) when the stream reaches it. */
#define FD_NOALLOC_MAX_ROWS 65536
int fd_transform_noalloc(fd_handle* h, const uint8_t* text_dev, const int64_t* offs_dev, int64_t n_rows,
                         uint32_t flags, double* out_pred_dev, double* out_prob_dev, double* out_raw_dev,
                         int32_t* out_status_dev, void* stream);
/* The handle owns one device workspace (control block, deferred-dialogue list, wide count tables, CSR staging) that
 * grows on demand: a compute call whose batch outgrows it SYNCHRONISES THE DEVICE, frees and reallocates it (cudaFree +
 * cudaMalloc) before it enqueues anything.  fd_workspace_bytes says how many bytes a batch of n_rows dialogues needs
 * (fd_transform and fd_featurize with its default staging estimate; the fixed part is about 204 MB), fd_reserve grows
 * the workspace to that size now, so that later calls of up to n_rows dialogues neither allocate nor synchronise
 * (fd_featurize on text with more than 96 distinct terms per dialogue on average still re-sizes its staging once).
 * No counterpart in the reference: Spark sizes its executors' memory through spark.executor.memory. */
int64_t fd_workspace_bytes(fd_handle* h, int64_t n_rows);
int fd_reserve(fd_handle* h, int64_t n_rows);
/* ---- evaluators on the transform output (fraud_detection_spark.py:101-119) -------------------------------- */
/* BinaryClassificationEvaluator(metricName = "areaUnderROC"): mllib BinaryClassificationMetrics(scoreAndLabels, numBins) on
 * one partition.  score_dev + i * score_stride (in doubles) is row i's score -- pass rawPrediction + 1 with stride 2 for the
 * reference's column; label_dev f64[n] (> 0.5 = positive).  out_dev f64[3] receives {area over the points up to the last
 * distinct score, positives, negatives}; the caller adds the closing segment to (1, 1), which is zero unless a class is
 * absent.  Allocates and frees its scratch, synchronises `stream`. */
int fd_eval_auc_roc(fd_handle* h, const double* score_dev, int64_t score_stride, const double* label_dev, int64_t n,
                    int32_t num_bins, double* out_dev, void* stream);
/* The counts behind MulticlassClassificationEvaluator (accuracy, weightedPrecision, weightedRecall, f1) and
 * crosstab("labels", "prediction") for binary labels: out_counts int64[5] = rows with (label, prediction) = (0,0), (0,1),
 * (1,0), (1,1), and rows holding any other value (the caller then falls back to a general tabulation).  Asynchronous. */
int fd_eval_confusion2(fd_handle* h, const double* label_dev, const double* pred_dev, int64_t n, int64_t* out_counts_dev,
                       void* stream);
/* Number of kernel launches the handle has issued since creation (bench.py's gpu_launches). */
int64_t fd_launch_count(fd_handle* h);

/* ---- RandomForest training: ml.tree.impl.RandomForest.run reached via Pipeline.fit
 *      (fraud_detection_spark.py:85-91) ----------------------------------------------------- */
/* TreePoint binning: bin[k] = lower_bound(thresholds of feature idx[k], val[k]) as u8.
 * thr_ptr int32[num_features+1] indexes thr f64[]. */
int fd_bin_csr(fd_handle* h, const int32_t* idx_dev, const double* val_dev, int64_t nnz,
               const int32_t* thr_ptr_dev, const double* thr_dev, uint8_t* out_bin_dev, void* stream);
/* The trainer's column view of this rank's binned rows (Spark keeps TreePoint rows; here a histogram pass walks COLUMNS):
 * out_rowidx u32[nnz] = row index with the row's label in bit 31, out_bin u8[nnz], both ordered by (feature, row) -- the
 * stable order of the CSR entries sorted by idx.  The column pointers are the prefix sums of the features' entry counts
 * (caller).  Radix sort over the bits of a feature id only.  scratch_dev: caller-owned device memory of at least
 * fd_csc_scratch_bytes(h, nnz, num_features) bytes -- about 32 per entry -- (FD_ERR_CAPACITY otherwise).  Asynchronous. */
int64_t fd_csc_scratch_bytes(fd_handle* h, int64_t nnz, int32_t num_features);
int fd_csc_build(fd_handle* h, const int64_t* rowptr_dev, const int32_t* idx_dev, const uint8_t* bin_dev, const uint8_t* label_dev,
                 int64_t n_rows, int64_t nnz, int32_t num_features, uint32_t* out_rowidx_dev, uint8_t* out_bin_dev, void* scratch_dev,
                 int64_t scratch_bytes, void* stream);
/* BaggedPoint.convertToBaggedRDD (bootstrap = true, reached from RandomForestClassifier, fraud_detection_spark.py:67-74):
 * out_bagw u8[n_trees, n_rows] = Poisson(rate) draws (clamped to 255) from a counter-based generator keyed by
 * (seed, tree, row) -- reproducible, independent of the launch shape; not Spark's XORShift stream (bag weights are an
 * INPUT of the parity contract).  out_root u64[n_trees, 2] (zeroed by the call) receives the weighted class counts
 * of this rank's rows: sum over rows of bagw[t, row] * [label[row] == c]. */
int fd_rf_bag_poisson(fd_handle* h, uint64_t seed, double rate, int32_t n_trees, int64_t n_rows, const uint8_t* label_dev,
                      uint8_t* out_bagw_dev, uint64_t* out_root_dev, void* stream);
/* binSeqOp over a node group.  An ITEM is (tree, node, feature): item_tree/item_node/item_feat int32[n_items],
 * items of one node contiguous and in the node's feature-subset order.  For every item the kernel walks the
 * feature's COLUMN (CSC over this rank's row shard: colptr int64[num_features+1], rowidx u32[nnz] = row index
 * with the row's LABEL in bit 31, bin u8[nnz]) and adds the bag weight of the entries whose row currently sits in
 * the node: hist[item, bin, label] += weight.  state u32[n_trees, n_rows] = node_id << 8 | bag_weight (heap
 * numbering, root 1; node ids below 2^24) -- one gather per column entry.  Long columns are cut into segments of
 * seg_len entries: UNIT u covers segment unit_seg[u] of item unit_item[u]'s column (one warp per unit).
 * item_chain int32[n_items] (or NULL = all zero): how many of the node's nearest ancestors reach it through DEFAULT links
 * only -- a row whose state still names one of them sits in the node (rows without a split feature are never rewritten, see
 * fd_rf_advance).
 * out_hist u32[n_items, 32, 2] is zeroed by the call.  Rows WITHOUT the feature are not added here:
 * fd_rf_best_split adds them after the cross-rank sum (NCCL all-reduce of out_hist by the caller). */
int fd_rf_histogram(fd_handle* h, const int64_t* colptr_dev, const uint32_t* rowidx_dev, const uint8_t* bin_dev,
                    const uint32_t* state_dev, int64_t n_rows, int32_t n_items, const int32_t* item_tree_dev,
                    const int32_t* item_node_dev, const int32_t* item_feat_dev, const int32_t* item_chain_dev, int32_t n_units,
                    const int32_t* unit_item_dev, const int32_t* unit_seg_dev, int32_t seg_len, uint32_t* out_hist_dev,
                    void* stream);
/* binsToBestSplit for a node group (histograms already summed over ranks).  node_item_ptr int32[n_nodes+1]
 * delimits each node's items; node_total f64[n_nodes, 2] = class totals of the node; n_splits
 * int32[num_features], bin_of_zero u8[num_features].  The implied zero-valued rows (node_total - explicit)
 * are placed in bin_of_zero.  Gini gain in Spark's operation order; first maximum over splits, then over the
 * node's items.  Outputs per node: best_item int32 (-1: no valid split), best_split int32 (bins <= it go
 * left), best_gain f64, left / right class counts f64[n_nodes, 2]. */
int fd_rf_best_split(fd_handle* h, const uint32_t* hist_dev, int32_t n_nodes, const int32_t* node_item_ptr_dev,
                     const int32_t* item_feat_dev, const double* node_total_dev, const int32_t* n_splits_dev,
                     const uint8_t* bin_of_zero_dev, int32_t min_instances, double min_info_gain,
                     int32_t* out_best_item_dev, int32_t* out_best_split_dev, double* out_best_gain_dev,
                     double* out_left_dev, double* out_right_dev, void* stream);
/* Move rows to their children after a level whose node ids are [level_base, 2*level_base): the node field of
 * state (see fd_rf_histogram) is rewritten, the bag weight kept.  For each split node s
 * (split_tree/node/feat/bin int32[n_split_nodes]) rows holding the feature go to 2*node when bin <= split_bin else
 * 2*node+1 (units as in fd_rf_histogram, unit_item indexes the split list); every other row of that level follows
 * action int32[n_trees, level_base]: the child id that bin_of_zero selects, or -1 when its node became a leaf (the
 * row keeps its id).  action_dev = NULL skips that pass over all of state: the rows keep naming the node they last reached
 * through an explicit entry, and the caller passes chain lengths (split_chain int32[n_split_nodes] here, item_chain to
 * fd_rf_histogram; NULL = zeros) so that the walks recognise them -- what training.ForestTrainer does. */
int fd_rf_advance(fd_handle* h, const int64_t* colptr_dev, const uint32_t* rowidx_dev, const uint8_t* bin_dev,
                  uint32_t* state_dev, int64_t n_rows, int32_t n_trees, int32_t level_base, int32_t n_split_nodes,
                  const int32_t* split_tree_dev, const int32_t* split_node_dev, const int32_t* split_feat_dev,
                  const int32_t* split_bin_dev, const int32_t* split_chain_dev, int32_t n_units, const int32_t* unit_item_dev,
                  const int32_t* unit_seg_dev, int32_t seg_len, const int32_t* action_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FDB200_H */
