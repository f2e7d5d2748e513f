// Internal declarations shared by b2rf_kernels.cu (device code + launchers) and b2rf_trainer.cu (host driver).
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
  uint32_t catmask[8];
};

struct HistLaunch {
  const uint8_t* bins; int64_t ldb; int32_t F;
  const double* y;
  const uint8_t* w; int64_t ldw;
  int32_t* node; int64_t ldn;       // traverse mode: tree-local node index per (tree,row); direct mode: slot per (g,row)
  const int32_t* tree_ids; int32_t Tg;
  int64_t n_rows;
  const DNode* dnodes; int64_t nodeStride; const uint32_t* catmasks;
  const uint2* slot_fo; const int64_t* slot_base; int32_t Fsub;
  double* hist;
  uint32_t* slot_count;             // MODE 2 only: rows per slot
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
  double* hist;
};

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
};
cudaError_t launch_best_split(const SplitLaunch& a, cudaStream_t s);

int device_sm_count();
void set_error(const std::string& m);
bool cuda_ok(cudaError_t e, const char* what);

}  // namespace b2rf
