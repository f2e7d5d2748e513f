// Internal declarations shared by the kernel files (k1_find_splits.cu ... k7_predict.cu: device code + launchers) and b2rf_trainer.cu (host driver).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>

namespace b2rf {

// Device-side mirror of one LearningNode (16 bytes, one 128-bit load).
//  feature : split feature, or -1 while the node has no split (pending on the stack, or a finished leaf)
//  left    : tree-local index of the left child (right child = left + 1)
//  slot    : >= 0 node is in the current group (histogram slot); -1 pending; -2 finished leaf (rows retire)
//  aux     : continuous split: highest bin that goes left; categorical: 0x80000000 | index into the cat-mask pool
struct __align__(16) DNode {
  int32_t feature;
  int32_t left;
  int32_t slot;
  uint32_t aux;
};
static constexpr int32_t SLOT_PENDING = -1;
static constexpr int32_t SLOT_DEAD = -2;

struct DNodeUpdate {  // scatter record: dnodes[pos] = node
  int64_t pos;
  DNode node;
};

// Result of binsToBestSplit for one node slot.
struct __align__(8) SplitRec {
  double gain, impurity, leftImp, rightImp;
  double parent[4], left[4], right[4];
  int32_t feature, splitIdx, valid, pad;
};   // the left-category set of a categorical best split goes to SplitLaunch::catmask_out[slot][maskWords]

struct HistLaunch {
  const uint8_t* bins; int64_t ldb; int32_t F;   // ldb counts codes (elements), not bytes
  const double* y;
  const uint8_t* w; int64_t ldw;
  int32_t* node; int64_t ldn;       // traverse mode: tree-local node index per (tree,row); direct mode: slot per (g,row)
  const int32_t* tree_ids; int32_t Tg;
  int64_t n_rows;
  const DNode* dnodes; int64_t nodeStride; const uint32_t* catmasks;
  const uint2* slot_fo; const int64_t* slot_base; int32_t Fsub;
  double* hist;
  uint32_t* slot_count;             // MODE 2 only: rows per slot
  int32_t codeBytes;                // 1: uint8 bin codes (maxBins <= 256), 2: uint16
  int32_t maskWords;                // 32-bit words per categorical left-set mask
};

// Row segment of one node in a tree's row list (segment mode): rows [start, start+count) of list `buf` (0/1 ping-pong).
struct __align__(16) DSeg {
  uint64_t start;
  uint32_t count;
  uint32_t buf;
};

// Everything a segment-mode K4 task needs to know about its node slot, resolved on the host: lets one launch serve slots of
// several trainers (fitMultiple's shared passes: every candidate forest of a fold rides in the same K4 launch).
struct __align__(16) SlotSrc {
  const DSeg* seg;                              // the node's row segment
  const uint32_t* idx0; const uint32_t* idx1;   // the owning trainer's packed row lists (both ping-pong buffers)
  const double* y0; const double* y1;           // ... and label lists
  const uint2* fo;                              // [Fsub] (feature, first cell of that feature inside the slot)
  double* hb;                                   // first cell of the slot in its trainer's histogram
  int32_t cells; int32_t pad;
};

struct BucketLaunch {
  const uint8_t* bins_rm; int64_t ldr;   // row-major bin codes [row][ldr]
  const double* y;
  const uint32_t* rowlist;               // packed row | weight << 27, grouped by slot
  const uint64_t* slot_start;            // [nSlots + 1]
  const uint32_t* slot_count;            // [nSlots]
  const uint32_t* chunk_prefix;          // [nSlots + 1] exclusive prefix of ceil(count / CH)
  int32_t nSlots, CH;
  const uint2* slot_fo; const int64_t* slot_base; const int32_t* slot_cells; int32_t Fsub;
  int32_t maxCells;                      // shared-memory histogram capacity in cells
  int32_t padB;                          // > 0: padded [bin][feature-in-group] shared layout with padB bins; 0: compact cells
  int32_t G;                             // padded layout: features per group / task (power of two <= 32)
  int32_t codeBytes;                     // 1 / 2 bytes per bin code; ldr counts codes
  int32_t zeroReg;                       // generic kernel: accumulate bin 0 of every feature in registers, one add per task
  double* hist;
  // segment mode (dseg != nullptr): slot s owns rows idx[seg.buf][seg.start ...] with seg = dseg[slot_pos[s]]
  const DSeg* dseg; const int64_t* slot_pos; const uint32_t* idx0; const uint32_t* idx1;
  // labels carried with the row lists (same positions as idx0 / idx1); nullptr: gather a.y[row]
  const double* y0; const double* y1;
  unsigned long long* rows_acc;          // optional device counter: += rows visited by this launch (segment mode; physical update count)
  const SlotSrc* slots;                  // limb kernel only: per-slot sources (overrides dseg / slot_pos / idx / y / slot_fo / slot_base / hist)
};

// Which K4 implementation a launch used (b2rf_k4_launch_counts; tests assert on it).
enum { B2RF_K4_LIMB = 0, B2RF_K4_FP64_32 = 1, B2RF_K4_FP64_GENERAL = 2, B2RF_K4_CLASS_COUNTS = 3, B2RF_K4_SCAN = 4, B2RF_K4_CLASS_STAGED = 5, B2RF_K4_VARIANTS = 6 };
void count_k4_variant(int v);
long long k4_variant_count(int v);

// Limb kernel (k4_limb.cu): launch shape chosen by limb_plan from the histogram size and the code-row width.
// NP = planes per cell: 9 (wide windows, exact for any realistic label range), 7 (narrow: fits whose labels are all modest integers)
// or 2 (labels all 0 / 1).
struct LimbPlan { int G, BR, threads, padB, NP, tma; };
bool limb_plan(const BucketLaunch& b, int maxBinsAll, int form, bool wantTma, LimbPlan* p);   // form: 0 wide, 1 narrow, 2 binary
cudaError_t launch_limb_hist(const BucketLaunch& b, const LimbPlan& p, cudaStream_t s);
cudaError_t launch_seg_tasks_slots(const SlotSrc* slots, int n, int CH, uint32_t* chunkPrefix, cudaStream_t s);   // task table + ticket from SlotSrc segments
cudaError_t launch_labels_integer(const double* y, int64_t n, int32_t* d_flag, cudaStream_t s);   // d_flag[0] / [1] (preset to 1) cleared unless every label is an integer, |y| <= 2^24 / is 0 or 1

constexpr int SEG_CH = 4096;            // rows per (node, chunk) task
constexpr int COMPACT_ROWS = 2048;      // rows per compaction block

struct PartLaunch {                      // stable partition of the row segments of the nodes split in one group
  const int64_t* split_pos; int32_t nSplits;          // flat node-table positions (tree * stride + node) of the split nodes
  const DNode* dnodes; DSeg* dseg; const uint32_t* catmasks;
  const uint8_t* bins; int64_t ldb;                    // column-major bin matrix
  uint32_t* idx0; uint32_t* idx1;
  double* y0; double* y1;                              // labels carried with the lists (nullptr: not carried)
  uint32_t* chunk_prefix;                              // [nSplits + 1]
  uint32_t* left_cnt; uint32_t* left_off;              // [tasks]
  uint32_t* predbits;                                  // [tasks][SEG_CH / 32]
  int32_t codeBytes, maskWords;
};
cudaError_t launch_partition(const PartLaunch& p, cudaStream_t s);
cudaError_t launch_segment_hist(const BucketLaunch& b, cudaStream_t s);
// chunkPrefix[i] = exclusive prefix of ceil(count_i / CH) over the row segments dseg[pos[i]], i < n; chunkPrefix[n] = total
// ticket: also clears chunkPrefix[n + 1], the work counter of kernels that hand out tasks dynamically
cudaError_t launch_seg_tasks(const DSeg* dseg, const int64_t* pos, int n, int CH, uint32_t* chunkPrefix, cudaStream_t s, bool ticket = false,
                             unsigned long long* rows_acc = nullptr);
cudaError_t launch_compact_count(const uint8_t* w, int64_t ldw, int T, int64_t n_rows, uint32_t* blockCnt, unsigned long long* totals, cudaStream_t s);
cudaError_t launch_compact_write(const uint8_t* w, int64_t ldw, int T, int64_t n_rows, const uint32_t* blockOff, const uint64_t* treeBase, uint32_t* idx,
                                 const double* y, double* ylist, cudaStream_t s);
cudaError_t launch_regrow_segs(const DSeg* src, int64_t oldStride, DSeg* dst, int64_t newStride, int T, cudaStream_t s);

cudaError_t launch_hist(const HistLaunch& a, int mode, cudaStream_t s, std::string* err);
cudaError_t launch_bucket_pipeline(const HistLaunch& h, const BucketLaunch& b, uint32_t* cursor, cudaStream_t s, std::string* err);
cudaError_t launch_count_active(const int32_t* node, int64_t ldn, int T, int64_t n_rows, unsigned long long* out, cudaStream_t s);
cudaError_t launch_scatter_dnodes(DNode* dnodes, const DNodeUpdate* upd, int64_t n, cudaStream_t s);

struct SplitLaunch {
  const double* hist;
  const uint2* slot_fo; const int64_t* slot_base; int32_t Fsub; int32_t nSlots;
  const double* parent4;      // [slots][4]
  const int32_t* hasParent;   // [slots]
  const int32_t* numBins;     // [F]
  const int32_t* isCat;       // [F]
  double minInstances, minWeight, minInfoGain;
  SplitRec* out;
  uint32_t* catmask_out;      // [slots][maskWords]
  int32_t maxB;               // largest numBins of any feature (sizes the per-warp shared arrays)
  int32_t maskWords;
  int32_t impurity;           // 0 variance, 1 gini, 2 entropy (binary classification on the regression statistics, see b2rf.h)
  int32_t allHaveParent;      // every slot has parent statistics (no root node in the group): several warps may share a slot
};
cudaError_t launch_best_split(const SplitLaunch& a, cudaStream_t s);

// ---- multiclass classification (numClasses > 2): per-class counts, k8_multiclass.cu ----
struct ClsHistLaunch {                   // segment mode only
  const void* bins_rm; int64_t ldr; int32_t codeBytes;
  const double* y;                       // labels 0 .. C-1
  const uint32_t* chunk_prefix; int32_t nSlots, CH;
  const uint2* slot_fo; const int64_t* slot_base; int32_t Fsub;
  int32_t padB;                          // bins per feature plane in shared memory (largest cell count of any feature)
  int32_t C, M;                          // classes; a global histogram cell is 4*M doubles, the first C of them the class counts
  int32_t direct;                        // set by the launcher when C * padB * 32 counters do not fit shared memory
  unsigned long long* rows_acc;          // optional device counter: += rows visited by this launch
  double* hist;
  const DSeg* dseg; const int64_t* slot_pos; const uint32_t* idx0; const uint32_t* idx1;
  const double* y0; const double* y1;    // labels carried with the row lists (nullptr: gather y[row]); the staged kernel needs them
  const int32_t* slot_cells;             // [nSlots] histogram cells of the slot (over its Fsub features)
  int32_t staged;                        // 0: always the gather kernel (B2RF_K4C=gather)
};
cudaError_t launch_cls_hist(const ClsHistLaunch& a, cudaStream_t s);

struct ClsSplitLaunch {
  const double* hist;
  const uint2* slot_fo; const int64_t* slot_base; int32_t Fsub; int32_t nSlots;
  const double* parentCls;    // [slots][Cp] class counts of the node (from its parent's split); ignored for roots
  const int32_t* hasParent;   // [slots]
  const int32_t* cellBins;    // [F] histogram cells of the feature: splits + 1 (continuous) or arity (categorical)
  const int32_t* kind;        // [F] 0 continuous, 1 ordered categorical, 2 unordered categorical
  double minInstances, minWeight, minInfoGain;
  SplitRec* out;              // gain, impurities, feature, split index; parent[0] / left[0] / right[0] = weighted counts
  double* clsOut;             // [slots][3][Cp]: class counts of the node, of the best split's left and right child
  uint32_t* catmask_out;      // [slots][maskWords] left categories of a categorical best split
  int32_t maxB, maskWords, impurity, C, Cp, M;
};
cudaError_t launch_best_split_cls(const ClsSplitLaunch& a, cudaStream_t s);

int device_sm_count();
void set_error(const std::string& m);
const char* last_error_cstr();
bool cuda_ok(cudaError_t e, const char* what);

}  // namespace b2rf
