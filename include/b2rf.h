/* b2rf — B200-native MLlib-compatible decision-tree / random-forest regression training.
 *
 * C-ABI of libb2rf.so.  This is the drop-in boundary for the hot path that the reference reaches through
 *   pyspark.ml.regression.DecisionTreeRegressor / RandomForestRegressor  .fit()
 *   (reference call sites: `ML 06 - Decision Trees.py:73-75,118,126`,
 *    `ML 07 - Random Forests and Hyperparameter Tuning.py:41,107-116,143-149`, `ML 08 - Hyperopt.py:71,97-98`)
 * i.e. JavaEstimator._fit_java -> py4j -> org.apache.spark.ml.tree.impl.RandomForest.run (Spark MLlib 3.1.2, not
 * vendored in the reference).  Each entry point names the MLlib function it replaces.
 *
 * Conventions
 *  - Every function returns 0 on success, <0 on error; b2rf_last_error() gives the thread-local message.
 *    Messages for user errors repeat MLlib's text (e.g. the maxBins requirement `ML 06:81,110-118` expects).
 *  - Pointers named d_* are DEVICE pointers owned by the caller (e.g. torch tensors); the library never frees them.
 *    The stage kernels (K1, K2, K4 stand-alone, K7) take caller-provided outputs / workspaces and allocate nothing.  The trainer
 *    object is different: it owns small descriptor buffers and, in segment mode, the per-tree row / label lists (12-24 bytes per
 *    in-bag row and tree), taken from a process-wide pool of device and pinned-host blocks that grows with cudaMalloc /
 *    cudaMallocHost on first use and is reused by later fits (blocks return to the pool when a trainer is destroyed).
 *    Pointers named h_* are host pointers.  `stream` is a cudaStream_t passed as void*; calls are asynchronous on it
 *    unless stated otherwise.
 *  - No CPU fallback exists: without a CUDA device every compute entry point fails with B2RF_ERR_CUDA.
 *  - Layouts: X is row-major f64 [rows][ldx]; the bin matrix is COLUMN-major bins[F][ldb] (ldb % 128 == 0) of uint8 codes,
 *    or uint16 codes when min(maxBins, numExamples) > 256;
 *    bag weights w[T][ldw] uint8; per-tree row->node ids node[T][ldn] int32 (-1 = row inactive for that tree);
 *    histograms hist[cells][4] f64 = (weight-sum, sum w*y, sum w*y*y, raw count) per (node-slot, feature, bin).
 */
#ifndef B2RF_H
#define B2RF_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2RF_OK 0
#define B2RF_ERR_ARG (-1)      /* bad argument */
#define B2RF_ERR_MAXBINS (-2)  /* MLlib: maxBins smaller than a categorical arity */
#define B2RF_ERR_DATA (-3)     /* MLlib: invalid categorical value */
#define B2RF_ERR_CUDA (-4)     /* CUDA runtime error / no device */
#define B2RF_ERR_STATE (-5)    /* call sequence error */

int b2rf_version(void);
const char* b2rf_last_error(void);

/* Parameters of one fit.  Mirrors the pyspark Params the labs set (`ML 06:118`, `ML 07:41,74-77`, `ML 08:71,97`). */
typedef struct b2rf_params {
  int32_t maxDepth;
  int32_t maxBins;
  int32_t minInstancesPerNode;
  int32_t numTrees;
  int32_t featuresPerNode;  /* resolved featureSubsetStrategy (b2rf_features_per_node) */
  int32_t maxMemoryInMB;
  int32_t bootstrap;
  int32_t histMode;  /* K4 implementation: 0 = segments (default), 1 = per-group counting sort, 2 = column scan + global RED */
  double minInfoGain;
  double minWeightFractionPerNode;
  double subsamplingRate;
  int64_t seed;
  int32_t impurity;  /* 0 = variance (regression); 1 = gini, 2 = entropy: binary classification (labels 0.0 / 1.0, numClasses == 2).
                      * Class counts ride on the regression statistics: with y in {0,1}, sum(w*y) is the class-1 count and
                      * sum(w) - sum(w*y) the class-0 count, both exact integers in fp64 (more than two classes: numClasses below). */
  int32_t numClasses;  /* 0: regression, or binary classification on the regression statistics (impurity != 0); 3..32: multiclass
                        * classification with per-class count histograms, impurity-ordered and unordered categorical splits
                        * (segment histogram mode only); 2 is accepted and means the same as 0 */
} b2rf_params;

/* DecisionTreeMetadata.buildMetadata featureSubsetStrategy table.  strategy: "auto","all","sqrt","log2","onethird",
 * "n" (integer > 0) or a fraction in (0,1].  Returns the count or B2RF_ERR_ARG. */
int b2rf_features_per_node(const char* strategy, int32_t F, int32_t numTrees);

/* DecisionTreeMetadata.buildMetadata checks: fails with B2RF_ERR_MAXBINS and MLlib's message when a
 * categorical arity exceeds maxPossibleBins = min(maxBins, numExamples) — the error `ML 06:81,91-92` expects.
 * On success returns metadata.maxBins (= max over features of the initial numBins), the value findSplits squares. */
int b2rf_check_metadata(const b2rf_params* params, int32_t F, const int32_t* h_arity, int64_t numExamples);

/* ---- K1: RandomForest.findSplits / findSplitsForContinuousFeature ------------------------ */
/* samplesFractionForFindSplits: max(maxBins_meta^2, 10000) / N if smaller than N, else 1.0 */
double b2rf_samples_fraction(int64_t numExamples, int32_t metaMaxBins);
/* Row sample used when fraction < 1 (replaces RDD.sample; per-global-row hash shared with the oracle).  Appends the
 * kept rows of X[row0_global + i] (i < n_rows) to d_sample_out [cap][F]; *d_count (device int64, caller-zeroed) counts them. */
int b2rf_sample_rows(const double* d_X, int64_t ldx, int64_t n_rows, int64_t row0_global, int32_t F, int64_t seed,
                     double fraction, double* d_sample_out, int64_t cap, int64_t* d_count, void* stream);
/* Host twin of b2rf_sample_rows: the local indices (ascending) of the rows it keeps, so a host-side caller can gather the
 * sample before any feature data reaches the device.  Returns the number kept (may exceed cap: then only cap are written). */
int64_t b2rf_sample_indices(int64_t seed, int64_t row0_global, int64_t n_rows, double fraction, int64_t* h_idx_out, int64_t cap);
/* Device form of the same decision without touching X (the rows may still be on the host): the local ids of the kept rows, in any
 * order, appended to d_ids_out [cap]; *d_count (device int64, caller-zeroed) counts them. */
int b2rf_sample_row_ids(int64_t n_rows, int64_t row0_global, int64_t seed, double fraction, int64_t* d_ids_out, int64_t cap, int64_t* d_count,
                        void* stream);
size_t b2rf_find_splits_workspace_bytes(int64_t n_sample, int32_t F);
/* d_sample [n_sample][F]; d_arity[F] (0 = continuous); out: d_thresholds [F][maxPossibleBins-1], d_nsplits[F]. */
int b2rf_find_splits(const double* d_sample, int64_t n_sample, int32_t F, const int32_t* d_arity, int32_t maxPossibleBins,
                     double weightedNumSamples, double* d_thresholds, int32_t* d_nsplits, void* d_workspace,
                     size_t workspace_bytes, void* stream);

/* Split finding for callers that must not allocate.  (No counterpart in the reference: MLlib allocates freely on the JVM
 * heap; this exists for hosts of this library whose GPU memory is entirely owned by their own pool — e.g. a Spark
 * executor plugin — or that cannot tolerate the device synchronisation of cudaMalloc/cudaFree.)
 * Same inputs, outputs and results as b2rf_find_splits, but the sort workspace is a fixed 8 MiB array built into the
 * library (resident on every GPU that loads it), so neither the caller nor the library allocates anything.
 * Limit: next_pow2(n_sample) * F * 8 <= B2RF_NOALLOC_WORKSPACE_BYTES, otherwise B2RF_ERR_ARG with an explanatory message.
 * The first call on a device resolves the array's address (do it before any stream capture); calls on one device share
 * the array, so calls that could overlap on the GPU must be serialised by the caller.  The array needs no clearing. */
#define B2RF_NOALLOC_WORKSPACE_BYTES (8u << 20)
size_t b2rf_find_splits_noalloc_max_bytes(void);
int b2rf_find_splits_noalloc(const double* d_sample, int64_t n_sample, int32_t F, const int32_t* d_arity, int32_t maxPossibleBins,
                             double weightedNumSamples, double* d_thresholds, int32_t* d_nsplits, void* stream);

/* ---- K2: TreePoint.convertToTreeRDD / findBin (row a3) ------------------------------------------------------ */
/* Bins rows [0,n_rows) of d_X into d_bins[f][row0 + i].  *d_err (int32, caller-zeroed) becomes feature+1 on an invalid
 * categorical value. */
int b2rf_bin(const double* d_X, int64_t ldx, int64_t n_rows, int32_t F, const double* d_thresholds, const int32_t* d_nsplits,
             const int32_t* d_arity, int32_t maxSplits, void* d_bins, int64_t ldb, int64_t row0, int32_t* d_err,
             void* d_bins_rm /* optional row-major twin [row][ldr], may be NULL */, int64_t ldr,
             int32_t code_bytes /* 1: uint8 codes (maxBins <= 256); 2: uint16.  ldb / ldr count codes */, void* stream);

/* ---- K3: BaggedPoint.convertToBaggedRDD (row a4) ------------------------------------------------------------- */
/* One emulated Spark partition per WARP (Bernoulli sampling and tiny rates: per thread): partition p (global index part0 + p) covers local rows
 * [h_part_offsets[p], h_part_offsets[p+1]).  Writes d_w[t][row] and, when d_node is not NULL, the initial d_node[t][row]
 * (0 if w > 0 else -1; only the sort / scan histogram modes need it). */
int b2rf_bag(int64_t seed, const int64_t* h_part_offsets, int32_t P, int32_t part0, int32_t T, double subsamplingRate,
             int32_t bootstrap, uint8_t* d_w, int64_t ldw, int32_t* d_node, int64_t ldn, void* stream);

/* ---- trainer: RandomForest.run driver loop (rows a1, a5, a9, a11) + K4/K5/K6 launches ------------------------ */
typedef struct b2rf_trainer b2rf_trainer;

/* h_arity[F], h_nsplits[F], h_thresholds[F][maxPossibleBins-1] (after K1).  numExamples = GLOBAL row count (all ranks).
 * Fails with B2RF_ERR_MAXBINS and MLlib's message when a categorical arity exceeds min(maxBins, numExamples). */
int b2rf_trainer_create(const b2rf_params* params, int32_t F, const int32_t* h_arity, const int32_t* h_nsplits,
                        const double* h_thresholds, int64_t numExamples, b2rf_trainer** out);
void b2rf_trainer_destroy(b2rf_trainer* t);
/* Largest histogram (in cells of 4 doubles) any node group can need; the caller allocates d_hist with that many cells. */
int64_t b2rf_trainer_max_group_cells(const b2rf_trainer* t);
/* d_bins_rm (optional, may be NULL): row-major twin of the bin matrix written by b2rf_bin.  With it the trainer buckets each
 * group's rows by node (counting sort) and builds the histograms in shared memory; without it, it scans the column-major
 * matrix and accumulates with global reductions (also selectable with B2RF_HIST_MODE=scan). */
int b2rf_trainer_code_bytes(const b2rf_trainer* t);   /* 1 or 2: the bin-code width this fit needs (min(maxBins, N) > 256 => 2) */
int b2rf_trainer_bind(b2rf_trainer* t, const void* d_bins, int64_t ldb, const void* d_bins_rm, int64_t ldr, const double* d_y,
                      const uint8_t* d_w, int64_t ldw, int32_t* d_node, int64_t ldn, int64_t n_rows_local, double* d_hist,
                      int64_t hist_cells);
/* selectNodesToSplit + K4 (findBestSplits.binSeqOp, row a6) with the fused K6 row routing (row a10).
 * Returns 1 and sets *n_cells (histogram cells written, to be summed across ranks by the caller — row a7) when a group
 * was launched, 0 when training is finished. */
int b2rf_trainer_begin_group(b2rf_trainer* t, int64_t* n_cells, void* stream);
/* K5 binsToBestSplit (row a8) on the (reduced) histogram, then the driver tail (row a9).  Synchronises the stream. */
int b2rf_trainer_end_group(b2rf_trainer* t, void* stream);
/* Rows sharded over `world` processes (reference: the reduceByKey + collectAsMap between the partition aggregates and
 * the driver in RandomForest.findBestSplits — MLlib also reduces BY NODE and finds each node's best split on the executor that owns
 * the node, sending only the split back): call b2rf_trainer_set_shard after create and before max_group_cells / bind.  A group's
 * node slots are then cut into `world` ranges of equal slot count and the histogram buffer is laid out as `world` chunks of equal
 * size (begin_group's *n_cells = world * chunk; padding cells stay zero), rank r's slots in chunk r.  Per group the caller does
 *   begin_group; reduce-scatter d_hist (sum; chunk r -> rank r, in place: recv = d_hist + r * chunk * 4 doubles);
 *   n = b2rf_trainer_shard_block_doubles(t); b2rf_trainer_end_group_shard_launch(t, d_block [n doubles], stream)   -- K5 on the own slots
 *   all-gather d_block -> d_blocks (world * n doubles); b2rf_trainer_end_group_shard_finish(t, d_blocks, stream)     -- driver tail on every rank
 * instead of all-reduce + b2rf_trainer_end_group (which still works on the padded layout: every rank then runs K5 on all slots).
 * Segment mode only; in the other modes set_shard changes nothing and shard_block_doubles returns 0. */
int b2rf_trainer_set_shard(b2rf_trainer* t, int32_t rank, int32_t world);
int64_t b2rf_trainer_shard_block_doubles(const b2rf_trainer* t);
int b2rf_trainer_end_group_shard_launch(b2rf_trainer* t, void* d_block, void* stream);
int b2rf_trainer_end_group_shard_finish(b2rf_trainer* t, const void* d_blocks, void* stream);
/* fitMultiple shared passes (reference: CrossValidator around the tree estimator, `ML 07 - Random Forests and
 * Hyperparameter Tuning.py:74-77,143-149`): the candidates of one fold are separate trainers bound to the SAME bin matrices.
 * b2rf_trainers_begin_group_shared = begin_group of every listed trainer (started[i] = 1 / 0, n_cells[i] as there) with ONE K4
 * launch whose task table spans the node slots of all of them (trainers whose launch shape differs get their own launch);
 * b2rf_trainers_end_group_shared = end_group of every listed trainer that started a group, with one stream synchronisation.
 * Every trainer ends with exactly the forest it builds alone. */
int b2rf_trainers_begin_group_shared(b2rf_trainer** ts, int32_t n, int64_t* n_cells, int32_t* started, void* stream);
int b2rf_trainers_end_group_shared(b2rf_trainer** ts, int32_t n, void* stream);
/* Introspection of the group in flight (valid between begin_group and end_group; debug / parity tests: the reduced histogram the
 * caller holds can then be compared cell by cell with DTStatsAggregator's contents).  Slot s is node h_node_id[s] (heap numbering,
 * root = 1) of tree h_tree[s]; its j-th drawn feature and that feature's first cell inside the slot are h_slot_fo[s][j][0..1]; the
 * slot's first cell in d_hist is h_slot_base[s].  Arrays may be NULL (pass only n_slots first to size them). */
int b2rf_trainer_group_info(const b2rf_trainer* t, int64_t* n_slots, int32_t* h_tree, int32_t* h_node_id, uint32_t* h_slot_fo,
                            int64_t* h_slot_base);
/* Process-wide switches, for A/B runs and tests: "k4_fp64" (1: segment mode uses the fp64 compare-and-swap histogram kernels
 * instead of the limb kernel; env B2RF_K4=fp64), "k4_tma" (1: the limb kernel stages code rows with one bulk async copy per row on
 * the TMA engine instead of register-staged 128-bit loads; env B2RF_K4_STAGE=tma), "k4_binary" (0: fits whose labels are all 0 / 1 use
 * the general integer-label form instead of the two-plane form; env B2RF_K4_BINARY=0), "carry_labels" (0: row lists do not carry labels, K4 gathers them; env
 * B2RF_CARRY_Y=0).  Takes effect at the next b2rf_trainer_bind / begin_group. */
int b2rf_set_option(const char* name, int64_t value);
/* h_seconds5: init(0), findSplits(0), findBestSplits (K4 device time from CUDA events), chooseSplits (K5 + driver tail),
 * total — MLlib TimeTracker phase names.  h_counts9: node groups, K4 launches, rows*trees scanned, learning nodes,
 * kernels launched by the trainer, algorithmic bytes summed over K4 launches (SURVEY.md §8d formula), rows visited by the
 * segment-mode K4 launches (times featuresPerNode = physical (row, feature) bin updates), featuresPerNode, weighted rows summed over
 * all evaluated nodes (times featuresPerNode = the updates MLlib's per-instance loop executes). */
int b2rf_trainer_stats(const b2rf_trainer* t, double* h_seconds5, int64_t* h_counts9);
/* finished forest, LearningNode.toNode(prune=true) in pre-order (row a11) */
int b2rf_trainer_forest_sizes(b2rf_trainer* t, int64_t* n_nodes, int64_t* n_cats);
int b2rf_trainer_export(b2rf_trainer* t, int64_t* h_tree_off, int32_t* h_feature, double* h_threshold, int32_t* h_left,
                        int32_t* h_right, double* h_prediction, double* h_impurity, double* h_gain, double* h_stats4,
                        int32_t* h_ncat, int32_t* h_cat_off, int32_t* h_cat_len, double* h_cats);
/* per-node decision trace in processing order: ints [n][5] = group, tree, heap node id, feature, split index;
 * doubles [n][5] = gain, impurity, count, sum, sumsq */
/* Classification: numClasses of the fit (2 for the binary form) and the class counts of every exported node,
 * h_counts[n_nodes][numClasses] (the model's impurityStats).  0 / B2RF_ERR_STATE for a regression trainer. */
int b2rf_trainer_num_classes(const b2rf_trainer* t);
int b2rf_trainer_export_class_counts(b2rf_trainer* t, double* h_counts);
/* The trace is recorded only after b2rf_trainer_set_trace(t, 1) (call before the first begin_group); off by default: one 72-byte row per
 * evaluated node is measurable host work on the critical path of large forests. */
int b2rf_trainer_set_trace(b2rf_trainer* t, int32_t on);
int64_t b2rf_trainer_trace_size(const b2rf_trainer* t);
/* Evidence / test hook: how many times this process launched each K4 implementation, h_counts[i] for i < n:
 * 0 limb kernel (native 32-bit shared-memory adds on fixed-point limbs; default of the segment mode), 1 fp64 kernel for
 * <= 32 features per node, 2 general fp64 kernel, 3 multiclass class-count launches (either form), 4 column-scan kernels,
 * 5 multiclass launches that took the staged form (code rows staged through shared memory, labels from the row lists). */
int b2rf_k4_launch_counts(int64_t* h_counts, int32_t n);
int b2rf_trainer_trace(const b2rf_trainer* t, int32_t* h_ints, double* h_dbls);

/* ---- stand-alone kernels (same code the trainer launches), for testing and for callers with their own driver -- */
/* K4: histogram of one node group.  d_slot_of_row [Tg][ldn] gives the group slot of each row per listed tree (-1 = skip);
 * d_tree_ids[Tg]; d_slot_fo [slots][Fsub] pairs (feature, cell offset inside the slot); d_slot_base[slots] cells. */
int b2rf_hist_build(const void* d_bins, int64_t ldb, int32_t code_bytes, int32_t F, const double* d_y, const uint8_t* d_w, int64_t ldw,
                    const int32_t* d_slot_of_row, int64_t ldn, const int32_t* d_tree_ids, int32_t Tg, int64_t n_rows,
                    const uint32_t* d_slot_fo, const int64_t* d_slot_base, int32_t Fsub, double* d_hist, void* stream);
/* K7: TreeEnsembleModel.predict on raw features (rows a11): mean over trees, tree-order summation. */
int b2rf_predict(const double* d_X, int64_t ldx, int64_t n_rows, int32_t T, const int64_t* d_tree_off, const int32_t* d_feature,
                 const double* d_threshold, const int32_t* d_left, const int32_t* d_right, const double* d_prediction,
                 const int32_t* d_ncat, const int32_t* d_cat_off, const int32_t* d_cat_len, const double* d_cats,
                 double* d_out, void* stream);
/* K7, packed form (the `_transform` fast path): the forest arrays are first packed into one 16-byte record per
 * node (b2rf_forest_pack, once per model and device; d_packed holds b2rf_forest_pack_bytes(n_nodes) bytes and is caller-owned),
 * after which a prediction visits one record per node.  Requires the pre-order numbering MLlib's NodeData uses (left child =
 * next node); the caller checks that and otherwise stays on b2rf_predict.  d_out[n_rows] (mean over trees, tree-order sum) and
 * d_leaf[n_rows][T] (tree-local index of the leaf each row reaches: leafCol / varianceCol) are each optional, not both NULL. */
size_t b2rf_forest_pack_bytes(int64_t n_nodes);
/* d_mask_idx (optional): for categorical nodes the row of d_masks that holds the node's left categories as a bit mask; with it
 * b2rf_predict_packed tests membership with one bit test (pass the same d_masks / maskWords there), without it it scans d_cats. */
int b2rf_forest_pack(int64_t n_nodes, const int32_t* d_feature, const double* d_threshold, const int32_t* d_right, const double* d_prediction,
                     const int32_t* d_ncat, const int32_t* d_cat_off, const int32_t* d_cat_len, const int32_t* d_mask_idx, void* d_packed,
                     void* stream);
int b2rf_predict_packed(const double* d_X, int64_t ldx, int64_t n_rows, int32_t T, const int64_t* d_tree_off, const void* d_packed,
                        const double* d_cats, const uint32_t* d_masks, int32_t maskWords, double* d_out, int32_t* d_leaf, void* stream);
/* K7 on bin codes (all-continuous forests): d_codes[n_rows][ldr] are the rows binned by b2rf_bin against the forest's OWN per-feature
 * threshold tables (sorted distinct split thresholds; d_bins may be NULL in that b2rf_bin call, d_bins_rm receives the codes);
 * d_nodes8[node] = { feature | split code << 16 (0xffff = leaf), right child }, pre-order numbering; leaf values from d_prediction.
 * x <= threshold <=> code(x) <= split code, so the leaves and the sums equal b2rf_predict's. */
int b2rf_predict_coded(const void* d_codes, int64_t ldr, int32_t code_bytes, int64_t n_rows, int32_t T, const int64_t* d_tree_off,
                       const void* d_nodes8, const double* d_prediction, double* d_out, int32_t* d_leaf, void* stream);
/* K7 for any numClasses >= 2: d_class_counts[node][numClasses] are the exported class counts; d_out[n_rows][numClasses]. */
int b2rf_predict_raw_mc(const double* d_X, int64_t ldx, int64_t n_rows, int32_t T, const int64_t* d_tree_off, const int32_t* d_feature,
                        const double* d_threshold, const int32_t* d_left, const int32_t* d_right, const double* d_class_counts,
                        int32_t numClasses, const int32_t* d_ncat, const int32_t* d_cat_off, const int32_t* d_cat_len,
                        const double* d_cats, int32_t mode, double* d_out, void* stream);
/* K7, classification (reference call sites `Solutions/Labs/ML 07L:80-82,105-108`): rawPrediction of a binary
 * tree-ensemble classifier, d_out2[n_rows][2].  d_stats4[node][4] are the exported node statistics (sum w, sum w*y, ., .), i.e.
 * class counts (stats[0] - stats[1], stats[1]).  mode 0 = RandomForestClassificationModel.predictRaw (per-tree normalised class
 * counts summed in tree order), mode 1 = DecisionTreeClassificationModel.predictRaw (class counts of tree 0's leaf). */
int b2rf_predict_raw(const double* d_X, int64_t ldx, int64_t n_rows, int32_t T, const int64_t* d_tree_off, const int32_t* d_feature,
                     const double* d_threshold, const int32_t* d_left, const int32_t* d_right, const double* d_stats4,
                     const int32_t* d_ncat, const int32_t* d_cat_off, const int32_t* d_cat_len, const double* d_cats,
                     int32_t mode, double* d_out2, void* stream);

#ifdef __cplusplus
}
#endif
#endif
