// b2rf host driver: the RandomForest.run loop of Spark MLlib 3.1.2 (selectNodesToSplit -> findBestSplits -> driver tail),
// re-designed around device-resident data: the host keeps only the tiny LearningNode trees, the LIFO node stack and the
// RNGs that define MLlib's order-dependent semantics; every pass over rows is a kernel.
#include "b2rf_internal.h"
#include "../../include/b2rf.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <functional>
#include <mutex>
#include <thread>
#include <string>
#include <unordered_map>
#include <vector>

using namespace b2rf;

namespace {

// ---- java.util.Random (feature-subset master stream; RandomForest.run `rng`) ----
struct JRandom {
  uint64_t s;
  explicit JRandom(int64_t seed) : s(((uint64_t)seed ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1)) {}
  int32_t bits32() { s = (s * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1); return (int32_t)((int64_t)s >> 16); }
  int64_t nextLong() { const int64_t hi = bits32(); const int64_t lo = bits32(); return (int64_t)((uint64_t)hi << 32) + lo; }
};

// ---- org.apache.spark.util.random.XORShiftRandom (reservoir sampler stream) ----
inline uint32_t rotl(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
uint32_t murmur_long(uint64_t be_value, uint32_t seed) {
  // MurmurHash3_x86_32 over the 8 big-endian bytes of a long
  uint8_t d[8];
  for (int i = 0; i < 8; i++) d[i] = (uint8_t)(be_value >> (56 - 8 * i));
  uint32_t h = seed;
  for (int i = 0; i < 8; i += 4) {
    uint32_t k = (uint32_t)d[i] | ((uint32_t)d[i + 1] << 8) | ((uint32_t)d[i + 2] << 16) | ((uint32_t)d[i + 3] << 24);
    k *= 0xcc9e2d51u; k = rotl(k, 15); k *= 0x1b873593u;
    h ^= k; h = rotl(h, 13); h = h * 5 + 0xe6546b64u;
  }
  h ^= 8u; h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
struct XorShift {
  uint64_t s;
  explicit XorShift(int64_t init) {
    const uint32_t lo = murmur_long((uint64_t)init, 0x3c074a61u);
    const uint32_t hi = murmur_long((uint64_t)init, lo);
    s = ((uint64_t)hi << 32) | lo;
  }
  int32_t next(int bits) { uint64_t n = s ^ (s << 21); n ^= n >> 35; n ^= n << 4; s = n; return (int32_t)(n & ((1ULL << bits) - 1)); }
  double nextDouble() { return (double)(((int64_t)next(26) << 27) + next(27)) * 0x1.0p-53; }
};

// SamplingUtils.reservoirSampleAndCount(0 until F, k, seed)
void reservoir_features(int F, int k, int64_t seed, uint32_t* out) {
  for (int i = 0; i < k; i++) out[i] = (uint32_t)i;
  if (k >= F) return;
  XorShift r(seed);
  long l = k;
  for (int item = k; item < F; item++) {
    l++;
    const long rep = (long)(r.nextDouble() * (double)l);
    if (rep < k) out[rep] = (uint32_t)item;
  }
}

// Iteration order of `nodesForGroup` (immutable.Map built by .toMap from a mutable.HashMap) — Appendix A.6.
uint32_t bswap_hash(uint32_t v) {
  uint32_t hc = v * 0x9e3775cdu;
  hc = __builtin_bswap32(hc);
  return hc * 0x9e3775cdu;
}
uint32_t trie_improve(uint32_t hcode) {
  uint32_t h = hcode + ~(hcode << 9);
  h ^= h >> 14;
  h += h << 4;
  return h ^ (h >> 10);
}
void group_tree_order(const std::vector<int>& inserted, std::vector<int>& out) {
  int tbl = 16;
  while ((int)inserted.size() > tbl * 3 / 4) tbl *= 2;
  const uint32_t ones = (uint32_t)tbl - 1;
  const int rot = __builtin_popcount(ones) % 32;
  const int shift = __builtin_clz(ones);
  std::vector<std::vector<int>> bucket(tbl);
  for (int key : inserted) {
    const uint32_t i = bswap_hash((uint32_t)key);
    const uint32_t imp = rot ? ((i >> rot) | (i << (32 - rot))) : i;
    auto& b = bucket[(imp >> shift) & ones];
    b.insert(b.begin(), key);
  }
  std::vector<int> mut;
  for (int b = tbl - 1; b >= 0; b--) for (int key : bucket[b]) mut.push_back(key);
  if (mut.size() <= 4) { out = mut; return; }
  std::vector<std::pair<uint64_t, int>> ks;
  for (int key : mut) {
    const uint32_t h = trie_improve((uint32_t)key);
    uint64_t k = 0;
    for (int lvl = 0; lvl < 7; lvl++) k = (k << 5) | ((h >> (5 * lvl)) & 31u);
    ks.push_back({k, key});
  }
  std::sort(ks.begin(), ks.end());
  out.clear();
  for (auto& p : ks) out.push_back(p.second);
}

struct NodeStats {  // ImpurityStats (regression)
  double gain = 0, impurity = 0;
  double calc[4] = {0, 0, 0, 0};
  bool valid = true;
};
struct HNode {  // LearningNode
  int32_t id = 1;
  int32_t left = -1;  // tree-local index; right = left + 1
  bool isLeaf = false, hasSplit = false;
  int32_t feature = -1;
  int32_t splitBin = -1;
  double threshold = 0;
  int32_t catMaskIdx = -1;
  int64_t cls = -1;   // multiclass: offset of this node's class counts in b2rf_trainer::clsPool
  NodeStats stats;
};
struct GNode { int32_t tree, nidx; };

inline int level_of(int id) { return 31 - __builtin_clz((unsigned)id); }
// ImpurityCalculator.predict: regression -> mean label; binary classification -> indexOfLargestArrayElement(class counts), the
// first maximum, i.e. class 1 only when its count (c[1]) strictly exceeds class 0's (c[0] - c[1])
inline double vpred(const double* c, int impurity) {
  if (impurity == 0) return c[0] == 0 ? 0.0 : c[1] / c[0];
  return c[1] > c[0] - c[1] ? 1.0 : 0.0;
}

// Library-owned scratch comes from two small process-wide pools (device memory / pinned host memory): a fit allocates dozens
// of descriptor buffers, and cudaMalloc / cudaMallocHost / cudaFree each cost 0.1-10 ms and synchronise the device, which
// would otherwise be paid again by every fit (CrossValidator runs many).  Blocks return to the pool when a trainer dies.
struct BlockPool {
  struct Block { void* p; size_t bytes; int dev; };   // device blocks are only reused on the device that owns them
  std::mutex mu;
  std::vector<Block> free_;
  bool pinned;
  explicit BlockPool(bool pin) : pinned(pin) {}
  static int current_device() { int d = 0; cudaGetDevice(&d); return d; }
  // Requests are rounded up to a coarse size class (1/8 steps between powers of two, <= 12.5 % slack) so that fits whose row counts
  // differ slightly — the folds of a cross-validation — reuse each other's blocks instead of paying cudaMalloc for gigabyte lists.
  static size_t size_class(size_t bytes) {
    if (bytes <= 4096) return 4096;
    size_t p = 1;
    while (p < bytes) p <<= 1;
    const size_t step = p >> 4;               // 1/8 of the lower power of two
    return (bytes + step - 1) / step * step;
  }
  void* take(size_t bytes, size_t* got) {
    bytes = size_class(bytes);
    const int dev = pinned ? -1 : current_device();
    {
      std::lock_guard<std::mutex> lk(mu);
      int best = -1;
      for (int i = 0; i < (int)free_.size(); i++)
        if (free_[i].dev == dev && free_[i].bytes >= bytes && free_[i].bytes <= bytes * 4 + 4096 && (best < 0 || free_[i].bytes < free_[best].bytes)) best = i;
      if (best >= 0) { Block b = free_[best]; free_.erase(free_.begin() + best); *got = b.bytes; return b.p; }
    }
    void* p = nullptr;
    cudaError_t e = pinned ? cudaMallocHost(&p, bytes) : cudaMalloc(&p, bytes);
    if (e == cudaErrorMemoryAllocation && trim() > 0) {   // the cached blocks of earlier fits are in the way: give them back and try again
      cudaGetLastError();
      e = pinned ? cudaMallocHost(&p, bytes) : cudaMalloc(&p, bytes);
    }
    if (!cuda_ok(e, pinned ? "trainer pinned alloc" : "trainer device alloc")) return nullptr;
    *got = bytes;
    return p;
  }
  // bytes of cached (unused) blocks of the current device / of pinned memory
  size_t cached() {
    std::lock_guard<std::mutex> lk(mu);
    const int dev = pinned ? -1 : current_device();
    size_t n = 0;
    for (const Block& b : free_) if (b.dev == dev) n += b.bytes;
    return n;
  }
  // frees every cached block of the current device; returns the bytes released
  size_t trim() {
    std::vector<Block> drop;
    {
      std::lock_guard<std::mutex> lk(mu);
      const int dev = pinned ? -1 : current_device();
      for (size_t i = 0; i < free_.size();) {
        if (free_[i].dev == dev) { drop.push_back(free_[i]); free_[i] = free_.back(); free_.pop_back(); } else i++;
      }
    }
    size_t n = 0;
    for (const Block& b : drop) { if (pinned) cudaFreeHost(b.p); else cudaFree(b.p); n += b.bytes; }
    return n;
  }
  void give(void* p, size_t bytes) {
    if (!p) return;
    std::lock_guard<std::mutex> lk(mu);
    const int dev = pinned ? -1 : current_device();
    if (free_.size() >= 768) { if (pinned) cudaFreeHost(p); else cudaFree(p); return; }   // a 9-candidate fitMultiple round returns ~400 blocks
    free_.push_back({p, bytes, dev});
  }
};
BlockPool& device_pool() { static BlockPool* p = new BlockPool(false); return *p; }
BlockPool& pinned_pool() { static BlockPool* p = new BlockPool(true); return *p; }

template <class T> struct DevBuf {
  T* p = nullptr; size_t cap = 0; size_t bytes = 0;
  bool ensure(size_t n) {
    if (n <= cap) return true;
    const size_t nc = std::max(n, cap * 2);
    device_pool().give(p, bytes);
    p = nullptr; cap = 0; bytes = 0;
    size_t got = 0;
    void* q = device_pool().take(nc * sizeof(T), &got);
    if (!q) return false;
    p = static_cast<T*>(q); bytes = got; cap = got / sizeof(T); return true;
  }
  void release() { device_pool().give(p, bytes); p = nullptr; cap = 0; bytes = 0; }
};
template <class T> struct PinBuf {
  T* p = nullptr; size_t cap = 0; size_t bytes = 0;
  bool ensure(size_t n) {
    if (n <= cap) return true;
    const size_t nc = std::max(n, cap * 2);
    pinned_pool().give(p, bytes);
    p = nullptr; cap = 0; bytes = 0;
    size_t got = 0;
    void* q = pinned_pool().take(nc * sizeof(T), &got);
    if (!q) return false;
    p = static_cast<T*>(q); bytes = got; cap = got / sizeof(T); return true;
  }
  void release() { pinned_pool().give(p, bytes); p = nullptr; cap = 0; bytes = 0; }
};

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// Process-wide switches (b2rf_set_option; initial values from the environment: B2RF_K4=fp64, B2RF_CARRY_Y=0).
std::atomic<int>& opt_k4_fp64() { static std::atomic<int> v{[] { const char* m = std::getenv("B2RF_K4"); return (m && std::string(m) == "fp64") ? 1 : 0; }()}; return v; }
std::atomic<int>& opt_k4_tma() { static std::atomic<int> v{[] { const char* m = std::getenv("B2RF_K4_STAGE"); return (m && std::string(m) == "tma") ? 1 : 0; }()}; return v; }
std::atomic<int>& opt_k4_bin() { static std::atomic<int> v{[] { const char* m = std::getenv("B2RF_K4_BINARY"); return (m && *m == '0') ? 0 : 1; }()}; return v; }   // two-plane form for 0 / 1 labels
std::atomic<int>& opt_carry_labels() { static std::atomic<int> v{[] { const char* m = std::getenv("B2RF_CARRY_Y"); return (m && *m == '0') ? 0 : 1; }()}; return v; }

}  // namespace

struct b2rf_trainer {
  b2rf_params p;
  int32_t F = 0, T = 0;
  int64_t numExamples = 0;
  int32_t maxPossibleBins = 0, maxSplits = 0, Fsub = 0;
  int32_t codeBytes = 1;    // bytes per bin code: 1 (maxBins <= 256) or 2
  int32_t maskWords = 8;    // 32-bit words per categorical left-set mask = ceil(max arity / 32)
  bool subsampling = false;
  double minWeightPerNode = 0;
  std::vector<int32_t> arity, numBins;
  // Multiclass classification (C >= 3, k8_multiclass.cu); C == 0 otherwise.  numBins stays MLlib's metadata.numBins (an unordered
  // categorical feature counts 2^(arity-1) - 1 of them: node-group memory accounting), cellBins is what the histogram really holds
  // (splits + 1, or one cell per category), kind = 0 continuous / 1 ordered categorical / 2 unordered categorical.
  int32_t C = 0, M = 1, Cp = 4;
  std::vector<int32_t> cellBins, kind;
  std::vector<double> clsPool;
  std::vector<int64_t> drawCells;    // histogram cells of a draw's features
  std::vector<double> thresholds;  // [F][maxSplits]
  int64_t allBins = 0;

  // bound device data (caller-owned)
  const uint8_t* d_bins = nullptr; int64_t ldb = 0;
  const double* d_y = nullptr;
  const uint8_t* d_w = nullptr; int64_t ldw = 0;
  int32_t* d_node = nullptr; int64_t ldn = 0;
  int64_t nLocal = 0;
  double* d_hist = nullptr; int64_t histCells = 0;
  const uint8_t* d_bins_rm = nullptr; int64_t ldr = 0;   // optional row-major twin of the bin matrix => bucket mode
  bool bucketMode = false;
  bool segMode = false;   // bucket mode with per-tree row lists kept partitioned by node (default); else per-group counting sort
  // rows sharded over `world` processes (b2rf_trainer_set_shard): a group's node slots are cut into `world` contiguous ranges of
  // shardSpr slots, rank r's range starts at cell r * shardChunk of the histogram buffer (reduce-scatter by node), K5 runs on the
  // own range only and the split records travel through an all-gather (b2rf_trainer_end_group_shard_launch / _finish)
  int32_t shardRank = 0, shardWorld = 1;
  int64_t shardSpr = 0, shardChunk = 0;        // block capacity in slots (largest range), cells per chunk
  std::vector<int64_t> shardLo;                // [world + 1] first slot of every rank's range
  int32_t chunkRows = 4096;

  // host trees + scheduler
  std::vector<std::vector<HNode>> trees;
  std::vector<std::pair<int32_t, int32_t>> stack;  // back() == head of MLlib's nodeStack
  JRandom rng{0};
  // Feature subsets are a pure function of the position in the master stream (k-th popped node <- k-th nextLong), not of the tree
  // state, so they are drawn ahead of need while the host waits for the GPU (end_group) and consumed in order by begin_group.
  std::vector<uint32_t> drawFeats;   // [draw][Fsub]
  std::vector<int64_t> drawBins;     // total bins of the draw's features
  size_t drawHead = 0;               // next unconsumed draw
  std::vector<uint32_t> updStamp, updSlot; uint32_t updGen = 0;   // flush_node_updates: last-writer-wins without hashing
  void ensure_draws(size_t n) {      // at least n unconsumed draws
    if (drawHead > 0 && drawHead == drawBins.size()) { drawFeats.clear(); drawBins.clear(); drawCells.clear(); drawHead = 0; }
    while (drawBins.size() - drawHead < n) {
      const size_t k = drawBins.size();
      drawFeats.resize((k + 1) * (size_t)Fsub);
      uint32_t* out = drawFeats.data() + k * (size_t)Fsub;
      reservoir_features(F, Fsub, rng.nextLong(), out);
      int64_t tb = 0, tc = 0;
      for (int j = 0; j < Fsub; j++) { tb += numBins[out[j]]; tc += cellBins[out[j]]; }
      drawBins.push_back(tb); drawCells.push_back(tc);
    }
  }
  std::vector<uint32_t> catPool;  // 8 words per categorical split
  size_t catUploaded = 0;

  // device-side descriptors (library-owned, small)
  int64_t nodeStride = 0;
  DevBuf<DNode> dn;
  DevBuf<DNodeUpdate> dnUpd;
  DevBuf<uint32_t> dCat;
  DevBuf<uint2> dSlotFo;
  DevBuf<int64_t> dSlotBase;
  DevBuf<double> dParent;
  DevBuf<int32_t> dHasParent, dNumBins, dIsCat, dTreeIds, dCellBins, dKind;
  DevBuf<double> dParentCls, dClsOut;
  PinBuf<double> hParentCls, hClsOut;
  DevBuf<SplitRec> dRec;
  DevBuf<uint32_t> dRecMask;
  PinBuf<uint32_t> hRecMask;
  DevBuf<uint32_t> dSlotCount, dCursor, dChunkPrefix, dRowList;
  DevBuf<uint64_t> dSlotStart;
  DevBuf<int32_t> dSlotCells;
  PinBuf<int32_t> hSlotCells;
  std::vector<int64_t> inBag;   // rows with weight > 0 per tree (bound of the per-group row list)
  // segment mode
  DevBuf<DSeg> dSeg;
  DevBuf<uint32_t> dIdx0, dIdx1, dPartPrefix, dLeftCnt, dLeftOff, dPredBits;
  DevBuf<double> dY0, dY1;      // labels carried with the row lists (same positions), when memory allows: K4 then reads them as a stream
  bool carryY = false;
  // fitMultiple shared passes: begin_group can leave its (limb) K4 launch to the caller, who merges the slots of several trainers
  bool deferK4 = false, deferred = false;
  BucketLaunch deferB; LimbPlan deferPlan;
  DevBuf<SlotSrc> dShared; PinBuf<SlotSrc> hShared; DevBuf<uint32_t> dSharedPrefix;
  bool labelsInt = false;       // every local label is an integer of magnitude <= 2^24: the narrow (7-plane) limb kernel is exact
  bool labels01 = false;        // every local label is 0 or 1: the two-plane form (count, sum of w*y)
  DevBuf<unsigned long long> dBindTmp;
  DevBuf<unsigned long long> dRowsAcc;   // rows visited by the K4 launches of this fit (segment mode)
  int64_t physRows = 0;
  double weightedRows = 0;      // sum over evaluated nodes of their weighted row count (x featuresPerNode = MLlib's per-instance updates)
  DevBuf<int64_t> dSlotPos, dSplitPos;
  PinBuf<int64_t> hSlotPos, hSplitPos;
  std::vector<uint64_t> treeBase;
  PinBuf<DNodeUpdate> hUpd, hUpdTail;   // staging of node-table updates: one buffer per call site (begin_group / driver tail), see flush_node_updates
  PinBuf<uint2> hSlotFo;
  PinBuf<int64_t> hSlotBase;
  PinBuf<double> hParent;
  PinBuf<int32_t> hHasParent, hTreeIds;
  PinBuf<SplitRec> hRec;
  std::vector<DNodeUpdate> pending;  // node-table changes produced by the last driver tail

  // current group
  std::vector<GNode> group;
  std::vector<int> groupTreeInsertOrder;
  int64_t groupCells = 0;
  bool inGroup = false;
  int groupNo = 0;

  // trace + stats
  struct Trace { int32_t g, t, id, f, s; double gain, imp, cnt, sum, sq; };
  std::vector<Trace> trace;
  bool keepTrace = false;   // b2rf_trainer_set_trace: record one Trace row per evaluated node (tests / parity tooling)
  double tHist = 0, tChoose = 0, tStart = 0, tTotal = 0;
  double tSelect = 0, tTail = 0, tWait = 0;   // host shares of tChoose (begin_group / end_group) and the end_group wait; printed with B2RF_TIMING=1
  int64_t nHistLaunches = 0, nScanned = 0, nKernels = 0;
  double histAlgBytes = 0;  // SURVEY §8(d) algorithmic bytes summed over K4 launches
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool evPending = false;
  int64_t l2CellBudget = 1500000;  // cells per K4 launch kept L2-resident (32 B each => 48 MB)

  // flat export cache
  bool exported = false;
  struct TreePlan { std::vector<uint8_t> leaf; std::vector<double> pred; std::vector<int32_t> sz, csz; };   // export pass 1 (see export_plan)
  std::vector<TreePlan> x_plan;
  std::vector<int64_t> x_tree_off, x_cat_base;

  ~b2rf_trainer() {
    dn.release(); dnUpd.release(); dCat.release(); dSlotFo.release(); dSlotBase.release(); dParent.release();
    dHasParent.release(); dNumBins.release(); dIsCat.release(); dTreeIds.release(); dRec.release();
    dCellBins.release(); dKind.release(); dParentCls.release(); dClsOut.release(); hParentCls.release(); hClsOut.release();
    dSeg.release(); dIdx0.release(); dIdx1.release(); dY0.release(); dY1.release(); dBindTmp.release(); dRowsAcc.release(); dShared.release(); hShared.release(); dSharedPrefix.release(); dPartPrefix.release(); dLeftCnt.release(); dLeftOff.release(); dPredBits.release();
    dSlotPos.release(); dSplitPos.release(); hSlotPos.release(); hSplitPos.release();
    dRecMask.release(); hRecMask.release();
    dSlotCount.release(); dCursor.release(); dChunkPrefix.release(); dRowList.release(); dSlotStart.release(); dSlotCells.release(); hSlotCells.release();
    hUpd.release(); hUpdTail.release(); hSlotFo.release(); hSlotBase.release(); hParent.release(); hHasParent.release(); hTreeIds.release(); hRec.release();
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
  }

  DNode mirror(int t, int ni, int slot) const {
    const HNode& n = trees[t][ni];
    DNode d;
    d.feature = n.hasSplit ? n.feature : -1;
    d.left = n.left;
    d.slot = slot;
    d.aux = n.hasSplit ? (n.catMaskIdx >= 0 ? (0x80000000u | (uint32_t)n.catMaskIdx) : (uint32_t)n.splitBin) : 0u;
    return d;
  }
  int state_slot(const HNode& n) const { return (n.isLeaf && !n.hasSplit) ? SLOT_DEAD : SLOT_PENDING; }
  double node_pred(const HNode& n) const {   // ImpurityCalculator.predict
    if (!C) return vpred(n.stats.calc, p.impurity);
    if (n.cls < 0) return 0.0;
    int best = 0;
    for (int k = 1; k < C; k++) if (clsPool[(size_t)n.cls + k] > clsPool[(size_t)n.cls + best]) best = k;   // indexOfLargestArrayElement: first maximum
    return (double)best;
  }
};

extern "C" {

int b2rf_features_per_node(const char* strategy, int32_t F, int32_t numTrees) {
  std::string s = strategy ? strategy : "auto";
  if (s == "auto") s = numTrees == 1 ? "all" : "onethird";  // regression
  if (s == "all") return F;
  if (s == "sqrt") return (int)std::ceil(std::sqrt((double)F));
  if (s == "log2") return std::max(1, (int)std::ceil(std::log((double)F) / std::log(2.0)));
  if (s == "onethird") return (int)std::ceil((double)F / 3.0);
  char* end = nullptr;
  const long n = std::strtol(s.c_str(), &end, 10);
  if (end && *end == 0 && !s.empty()) { if (n > 0) return (int)std::min<long>(n, F); }
  else {
    const double x = std::strtod(s.c_str(), &end);
    if (end && *end == 0 && !s.empty() && x > 0.0 && x <= 1.0) return (int)std::ceil(x * F);
  }
  set_error("Supported values: auto, all, onethird, sqrt, log2, (0.0-1.0], [1-n]. Got \"" + s + "\"");
  return B2RF_ERR_ARG;
}

static int check_arity(int32_t maxBins, int32_t F, const int32_t* h_arity, int64_t numExamples, int* maxPossibleBinsOut) {
  const int mpb = (int)std::min<int64_t>(maxBins, numExamples);
  int maxCat = 0, maxCatF = -1;
  for (int f = 0; f < F; f++) if (h_arity[f] > maxCat) { maxCat = h_arity[f]; maxCatF = f; }
  if (maxCatF >= 0 && maxCat > mpb) {
    char buf[512];
    snprintf(buf, sizeof buf,
             "requirement failed: DecisionTree requires maxBins (= %d) to be at least as large as the number of values in each "
             "categorical feature, but categorical feature %d has %d values. Consider removing this and other categorical "
             "features with a large number of values, or add more training examples.",
             mpb, maxCatF, maxCat);
    set_error(buf);
    return B2RF_ERR_MAXBINS;
  }
  *maxPossibleBinsOut = mpb;
  return B2RF_OK;
}

int b2rf_check_metadata(const b2rf_params* params, int32_t F, const int32_t* h_arity, int64_t numExamples) {
  if (!params || F <= 0 || !h_arity || numExamples <= 0) { set_error("b2rf_check_metadata: bad argument"); return B2RF_ERR_ARG; }
  int mpb = 0;
  const int rc = check_arity(params->maxBins, F, h_arity, numExamples, &mpb);
  if (rc) return rc;
  int mx = 0;
  for (int f = 0; f < F; f++) mx = std::max(mx, h_arity[f] > 1 ? h_arity[f] : mpb);
  return mx;
}

int b2rf_trainer_create(const b2rf_params* params, int32_t F, const int32_t* h_arity, const int32_t* h_nsplits, const double* h_thresholds,
                        int64_t numExamples, b2rf_trainer** out) {
  if (!params || !out || F <= 0 || numExamples <= 0 || params->numTrees <= 0 || params->maxDepth < 0 || params->maxDepth > 30) {
    set_error("b2rf_trainer_create: bad argument"); return B2RF_ERR_ARG;
  }
  if (params->impurity < 0 || params->impurity > 2) { set_error("b2rf_trainer_create: impurity must be 0 (variance), 1 (gini) or 2 (entropy)"); return B2RF_ERR_ARG; }
  if (params->numClasses != 0 && (params->impurity == 0 || params->numClasses < 2 || params->numClasses > 32)) {
    set_error("b2rf_trainer_create: numClasses must be 0 (regression / binary default) or 2..32 with impurity gini / entropy"); return B2RF_ERR_ARG;
  }
  b2rf_trainer* t = new b2rf_trainer();
  t->p = *params; t->F = F; t->T = params->numTrees; t->numExamples = numExamples;
  { int mpb = 0; const int rc = check_arity(params->maxBins, F, h_arity, numExamples, &mpb); if (rc) { delete t; return rc; } t->maxPossibleBins = mpb; }
  t->maxSplits = std::max(t->maxPossibleBins - 1, 1);
  t->arity.assign(h_arity, h_arity + F);
  if (t->maxPossibleBins > 2048) { set_error("b2rf: maxBins > 2048 is not supported (split-scan shared-memory arrays)"); delete t; return B2RF_ERR_ARG; }
  t->codeBytes = t->maxPossibleBins > 256 ? 2 : 1;
  { int maxAr = 1; for (int f = 0; f < F; f++) maxAr = std::max(maxAr, h_arity[f]); t->maskWords = std::max(1, (std::max(maxAr, t->maxPossibleBins) + 31) / 32); }
  t->numBins.assign(F, t->maxPossibleBins);
  for (int f = 0; f < F; f++) {
    if (h_arity[f] > 1) t->numBins[f] = h_arity[f];
    else if (h_arity[f] == 0) t->numBins[f] = h_nsplits[f] + 1;
  }
  t->cellBins = t->numBins;
  t->kind.assign(F, 0);
  for (int f = 0; f < F; f++) t->kind[f] = h_arity[f] > 1 ? 1 : 0;
  if (params->impurity != 0 && params->numClasses >= 3) {
    // DecisionTreeMetadata.buildMetadata, numClasses > 2: categorical features with few categories are UNORDERED (every subset of
    // categories is a candidate split): arity <= floor(log2(maxPossibleBins / 2 + 1) + 1), with numBins = 2^(arity-1) - 1
    t->C = params->numClasses; t->M = (t->C + 3) / 4; t->Cp = 4 * t->M;
    const int maxCatUnordered = (int)std::floor(std::log((double)(t->maxPossibleBins / 2 + 1)) / std::log(2.0) + 1);
    for (int f = 0; f < F; f++)
      if (h_arity[f] > 1 && h_arity[f] <= maxCatUnordered) { t->kind[f] = 2; t->numBins[f] = (1 << (h_arity[f] - 1)) - 1; }
  }
  t->thresholds.assign(h_thresholds, h_thresholds + (size_t)F * t->maxSplits);
  t->allBins = 0;
  for (int f = 0; f < F; f++) t->allBins += t->numBins[f];
  t->Fsub = params->featuresPerNode > 0 ? std::min(params->featuresPerNode, F) : F;
  t->subsampling = t->Fsub != F;
  t->minWeightPerNode = params->minWeightFractionPerNode * (double)numExamples;
  t->trees.assign(t->T, {});
  for (int i = 0; i < t->T; i++) { t->trees[i].reserve(64); t->trees[i].push_back(HNode()); t->stack.push_back({i, 0}); }
  t->rng = JRandom(params->seed);
  *out = t;
  return B2RF_OK;
}

void b2rf_trainer_destroy(b2rf_trainer* t) { delete t; }

int64_t b2rf_trainer_max_group_cells(const b2rf_trainer* t) {
  // selectNodesToSplit admits nodes while statsSize*totalBins*8 bytes fit maxMemoryInMB; always at least one node
  const int64_t maxMem = (int64_t)t->p.maxMemoryInMB * 1024 * 1024;
  const int64_t statsSize = t->p.impurity == 0 ? 3 : (t->C ? t->C : 2);   // aggregateSizeForNode: 3 for regression, numClasses for classification
  std::vector<int32_t> nb(t->cellBins);
  std::sort(nb.begin(), nb.end(), std::greater<int32_t>());
  int64_t worstNode = 0;
  for (int j = 0; j < t->Fsub; j++) worstNode += nb[j];
  // admitted MLlib bins <= maxMem / (statsSize * 8) + one node; an unordered feature can hold up to 2x more cells than bins (arity 2)
  const int64_t byMem = (maxMem / (statsSize * 8)) * (t->C ? 2 : 1) + worstNode;
  const int64_t maxNodesTotal = (int64_t)t->T << std::min(t->p.maxDepth, 24);
  // sharded layout: every rank's range is padded to the largest one (< one node more per rank)
  const int64_t pad = t->shardWorld > 1 ? (int64_t)t->shardWorld * worstNode : 0;
  return t->M * (std::max<int64_t>(worstNode, std::min<int64_t>(byMem, maxNodesTotal * worstNode)) + pad);   // in units of 4 doubles
}

int b2rf_trainer_code_bytes(const b2rf_trainer* t) { return t ? t->codeBytes : 0; }

int b2rf_trainer_bind(b2rf_trainer* t, const void* d_bins_v, int64_t ldb, const void* d_bins_rm_v, int64_t ldr, const double* d_y,
                      const uint8_t* d_w, int64_t ldw, int32_t* d_node, int64_t ldn, int64_t n_rows_local, double* d_hist, int64_t hist_cells) {
  const uint8_t* d_bins = static_cast<const uint8_t*>(d_bins_v);
  const uint8_t* d_bins_rm = static_cast<const uint8_t*>(d_bins_rm_v);
  if (!t || !d_bins || !d_y || !d_w || !d_hist || n_rows_local < 0 || ((ldb * (t ? t->codeBytes : 1)) & 15) || ldb < n_rows_local || ldw < n_rows_local || (d_node && ldn < n_rows_local)) {
    set_error("b2rf_trainer_bind: bad argument (ldb must be a multiple of 16 and every leading dimension >= n_rows)"); return B2RF_ERR_ARG;
  }
  if ((ldw & 7) != 0 || (reinterpret_cast<uintptr_t>(d_w) & 7) != 0) {
    set_error("b2rf_trainer_bind: d_w must be 8-byte aligned with ldw a multiple of 8 (the row lists are built from 8-byte loads of bag weights)"); return B2RF_ERR_ARG;
  }
  t->d_bins = d_bins; t->ldb = ldb; t->d_y = d_y; t->d_w = d_w; t->ldw = ldw; t->d_node = d_node; t->ldn = ldn; t->nLocal = n_rows_local;
  t->d_hist = d_hist; t->histCells = hist_cells;
  t->d_bins_rm = d_bins_rm; t->ldr = ldr;
  t->bucketMode = d_bins_rm != nullptr && ldr >= t->F && n_rows_local < ((int64_t)1 << 27);
  t->segMode = t->bucketMode;
  if (t->p.histMode == 2) { t->bucketMode = false; t->segMode = false; }
  else if (t->p.histMode == 1) t->segMode = false;
  if (const char* m = std::getenv("B2RF_HIST_MODE")) {
    const std::string mode(m);
    if (mode == "scan") { t->bucketMode = false; t->segMode = false; }
    else if (mode == "sort") t->segMode = false;
  }
  // device node table: one stride per tree
  const int64_t full = (int64_t)1 << std::min(t->p.maxDepth + 1, 17);
  t->nodeStride = full;
  if (!t->segMode && d_node == nullptr) { set_error(n_rows_local >= ((int64_t)1 << 27)
                ? "b2rf_trainer_bind: a shard of 2^27 (134,217,728) rows or more cannot use the packed row lists of the segment / sort modes (27-bit row ids); "
                  "pass d_node so the scan mode can run, or shard the rows over more GPUs"
                : "b2rf_trainer_bind: d_node is required unless the segment histogram mode is used");
    return B2RF_ERR_ARG; }
  if (t->bucketMode) {
    // in-bag rows per tree; in segment mode also the initial row lists (ascending rows, packed row | weight << 27)
    const int nb = (int)((n_rows_local + COMPACT_ROWS - 1) / COMPACT_ROWS);
    DevBuf<uint32_t> blockCnt;
    std::vector<unsigned long long> h_tot(t->T, 0);
    // scratch from the trainer's pool (no cudaMalloc / cudaFree per fit): [0, T) in-bag totals, [T, 2T + 1) list bases
    if (!blockCnt.ensure((size_t)std::max(nb, 1) * t->T) || !t->dBindTmp.ensure((size_t)2 * t->T + 1)) { blockCnt.release(); return B2RF_ERR_CUDA; }
    unsigned long long* d_tot = t->dBindTmp.p;
    bool ok = cuda_ok(cudaMemsetAsync(d_tot, 0, (size_t)t->T * 8, 0), "bind memset");
    ok = ok && cuda_ok(launch_compact_count(d_w, ldw, t->T, n_rows_local, blockCnt.p, d_tot, 0), "count in-bag rows");
    ok = ok && cuda_ok(cudaMemcpy(h_tot.data(), d_tot, (size_t)t->T * 8, cudaMemcpyDeviceToHost), "bind copy");
    t->inBag.assign(h_tot.begin(), h_tot.end());
    if (ok && t->segMode) {
      t->treeBase.assign(t->T + 1, 0);
      for (int i = 0; i < t->T; i++) t->treeBase[i + 1] = t->treeBase[i] + h_tot[i];
      const size_t total = std::max<uint64_t>(t->treeBase[t->T], 1);
      uint64_t* d_base = reinterpret_cast<uint64_t*>(t->dBindTmp.p + t->T);
      ok = ok && t->dIdx0.ensure(total) && t->dIdx1.ensure(total) && t->dSeg.ensure((size_t)t->nodeStride * t->T);
      // Labels ride with the row lists (two more lists of 8 bytes per in-bag row) unless that would not leave room on the device:
      // K4 then streams them instead of gathering a 32-byte sector per (row, tree).  B2RF_CARRY_Y=0 keeps the gathers.
      t->carryY = false;
      if (ok) {
        const bool want = opt_carry_labels().load() != 0;
        size_t freeB = 0, totB = 0;
        const size_t have = t->dY0.bytes + t->dY1.bytes;
        const size_t needY = 2 * total * 8 + ((size_t)4 << 30);
        bool fits = want && cudaMemGetInfo(&freeB, &totB) == cudaSuccess && (have >= 2 * total * 8 || freeB + have > needY);
        if (want && !fits && freeB + have + device_pool().cached() > needY) {   // blocks cached from earlier (smaller) fits would make the difference
          device_pool().trim();
          fits = cudaMemGetInfo(&freeB, &totB) == cudaSuccess && freeB + have > needY;
        }
        if (fits) t->carryY = t->dY0.ensure(total) && t->dY1.ensure(total);
        if (!t->carryY) { t->dY0.release(); t->dY1.release(); }
      }
      t->labelsInt = false; t->labels01 = false;
      if (ok && t->carryY) {   // one pass over the labels decides which limb form is exact for this fit
        int32_t ones[2] = {1, 1}, flag[2] = {0, 0};
        int32_t* d_flag = reinterpret_cast<int32_t*>(t->dBindTmp.p);    // the in-bag totals have been read back already
        ok = ok && cuda_ok(cudaMemcpy(d_flag, ones, 8, cudaMemcpyHostToDevice), "bind copy");
        ok = ok && cuda_ok(launch_labels_integer(d_y, n_rows_local, d_flag, 0), "label scan");
        ok = ok && cuda_ok(cudaMemcpy(flag, d_flag, 8, cudaMemcpyDeviceToHost), "bind copy");
        t->labelsInt = ok && flag[0] != 0;
        t->labels01 = ok && flag[1] != 0;
      }
      ok = ok && cuda_ok(cudaMemcpy(d_base, t->treeBase.data(), (size_t)(t->T + 1) * 8, cudaMemcpyHostToDevice), "bind copy");
      ok = ok && cuda_ok(launch_compact_write(d_w, ldw, t->T, n_rows_local, blockCnt.p, d_base, t->dIdx0.p, d_y, t->carryY ? t->dY0.p : nullptr, 0),
                         "build row lists");
      std::vector<DSeg> roots(t->T);
      for (int i = 0; i < t->T; i++) { roots[i].start = t->treeBase[i]; roots[i].count = (uint32_t)h_tot[i]; roots[i].buf = 0; }
      ok = ok && cuda_ok(cudaMemcpy2D(t->dSeg.p, (size_t)t->nodeStride * sizeof(DSeg), roots.data(), sizeof(DSeg), sizeof(DSeg), t->T,
                                      cudaMemcpyHostToDevice), "bind copy root segments");
      ok = ok && cuda_ok(cudaDeviceSynchronize(), "bind sync");
    }
    blockCnt.release();
    if (!ok) return B2RF_ERR_CUDA;
  }
  if (t->C && !t->segMode) { set_error("b2rf: multiclass classification needs the segment histogram mode (row-major bin matrix, histMode 0)"); return B2RF_ERR_ARG; }
  if (!t->dn.ensure((size_t)t->nodeStride * t->T) || !t->dRowsAcc.ensure(1)) return B2RF_ERR_CUDA;
  if (!cuda_ok(cudaMemset(t->dRowsAcc.p, 0, 8), "bind memset")) return B2RF_ERR_CUDA;
  if (!t->dNumBins.ensure(t->F) || !t->dIsCat.ensure(t->F) || !t->dCellBins.ensure(t->F) || !t->dKind.ensure(t->F)) return B2RF_ERR_CUDA;
  if (!cuda_ok(cudaMemcpy(t->dCellBins.p, t->cellBins.data(), t->F * 4, cudaMemcpyHostToDevice), "bind copy")) return B2RF_ERR_CUDA;
  if (!cuda_ok(cudaMemcpy(t->dKind.p, t->kind.data(), t->F * 4, cudaMemcpyHostToDevice), "bind copy")) return B2RF_ERR_CUDA;
  std::vector<int32_t> isCat(t->F);
  for (int f = 0; f < t->F; f++) isCat[f] = t->arity[f] != 0;
  if (!cuda_ok(cudaMemcpy(t->dNumBins.p, t->numBins.data(), t->F * 4, cudaMemcpyHostToDevice), "bind copy")) return B2RF_ERR_CUDA;
  if (!cuda_ok(cudaMemcpy(t->dIsCat.p, isCat.data(), t->F * 4, cudaMemcpyHostToDevice), "bind copy")) return B2RF_ERR_CUDA;
  if (!cuda_ok(cudaEventCreate(&t->ev0), "event") || !cuda_ok(cudaEventCreate(&t->ev1), "event")) return B2RF_ERR_CUDA;
  t->pending.clear();
  for (int i = 0; i < t->T; i++) t->pending.push_back({(int64_t)i * t->nodeStride, t->mirror(i, 0, SLOT_PENDING)});
  t->tStart = now_s();
  return B2RF_OK;
}

// Uploads the queued node-table changes (and new categorical masks) and applies them with one scatter launch.  The upload is an
// asynchronous copy from a pinned staging buffer; begin_group and the driver tail each have their own, because the tail's copy can
// still be queued behind other trainers' kernels on a shared stream when the next begin_group fills its buffer (a stream
// synchronisation always separates two uses of the same buffer).
static int flush_node_updates(b2rf_trainer* t, cudaStream_t s, bool tail = false) {
  {  // one scatter launch must not carry two records for the same node: keep the latest
    const size_t tableSize = (size_t)t->nodeStride * (size_t)t->T;
    if (t->updStamp.size() < tableSize) t->updStamp.resize(tableSize, 0);
    if (++t->updGen == 0) { std::fill(t->updStamp.begin(), t->updStamp.end(), 0); t->updGen = 1; }
    if (t->updSlot.size() < tableSize) t->updSlot.resize(tableSize);
    size_t w = 0;
    for (size_t i = 0; i < t->pending.size(); i++) {
      const size_t pos = (size_t)t->pending[i].pos;
      if (t->updStamp[pos] != t->updGen) { t->updStamp[pos] = t->updGen; t->updSlot[pos] = (uint32_t)w; t->pending[w++] = t->pending[i]; }
      else t->pending[t->updSlot[pos]] = t->pending[i];
    }
    t->pending.resize(w);
  }
  const size_t nUpd = t->pending.size();
  bool ok = true;
  if (nUpd) {
    PinBuf<DNodeUpdate>& hb = tail ? t->hUpdTail : t->hUpd;
    if (!hb.ensure(nUpd) || !t->dnUpd.ensure(nUpd)) return B2RF_ERR_CUDA;
    memcpy(hb.p, t->pending.data(), nUpd * sizeof(DNodeUpdate));
    t->pending.clear();
    ok &= cuda_ok(cudaMemcpyAsync(t->dnUpd.p, hb.p, nUpd * sizeof(DNodeUpdate), cudaMemcpyHostToDevice, s), "upload node updates");
    ok &= cuda_ok(launch_scatter_dnodes(t->dn.p, t->dnUpd.p, (int64_t)nUpd, s), "scatter node updates");
    t->nKernels++;
  }
  if (t->catPool.size() > t->catUploaded) {
    if (t->catPool.size() > t->dCat.cap) {  // grow, keeping old contents
      DevBuf<uint32_t> nb;
      if (!nb.ensure(std::max<size_t>(t->catPool.size() * 2, 1024))) return B2RF_ERR_CUDA;
      cudaStreamSynchronize(s);
      t->dCat.release(); t->dCat = nb; t->catUploaded = 0;
    }
    ok &= cuda_ok(cudaMemcpyAsync(t->dCat.p + t->catUploaded, t->catPool.data() + t->catUploaded, (t->catPool.size() - t->catUploaded) * 4,
                                  cudaMemcpyHostToDevice, s), "upload cat masks");
    cudaStreamSynchronize(s);  // catPool is pageable host memory
    t->catUploaded = t->catPool.size();
  }
  return ok ? B2RF_OK : B2RF_ERR_CUDA;
}

static int grow_node_table(b2rf_trainer* t, int64_t need, cudaStream_t s) {
  int64_t ns = t->nodeStride;
  while (ns < need) ns *= 2;
  cudaStreamSynchronize(s);
  const int64_t oldStride = t->nodeStride;
  t->nodeStride = ns;
  if (!t->dn.ensure((size_t)ns * t->T)) return B2RF_ERR_CUDA;
  if (t->segMode) {   // the row segments exist only on the device: copy them into the wider table
    DevBuf<DSeg> wide;
    if (!wide.ensure((size_t)ns * t->T)) return B2RF_ERR_CUDA;
    if (!cuda_ok(launch_regrow_segs(t->dSeg.p, oldStride, wide.p, ns, t->T, s), "regrow segments")) return B2RF_ERR_CUDA;
    cudaStreamSynchronize(s);
    t->dSeg.release();
    t->dSeg = wide;
  }
  // rebuild every entry from the host trees
  t->pending.clear();
  for (int i = 0; i < t->T; i++)
    for (size_t k = 0; k < t->trees[i].size(); k++)
      t->pending.push_back({(int64_t)i * ns + (int64_t)k, t->mirror(i, (int)k, t->state_slot(t->trees[i][k]))});
  return B2RF_OK;
}

int b2rf_trainer_begin_group(b2rf_trainer* t, int64_t* n_cells, void* stream) {
  if (!t || !t->d_bins) { set_error("b2rf_trainer_begin_group: trainer not bound"); return B2RF_ERR_STATE; }
  if (t->inGroup) { set_error("b2rf_trainer_begin_group: previous group not finished"); return B2RF_ERR_STATE; }
  cudaStream_t s = (cudaStream_t)stream;
  if (t->stack.empty()) {
    if (n_cells) *n_cells = 0;
    if (t->tTotal == 0) {
      if (t->dRowsAcc.p) { unsigned long long r = 0; if (cudaMemcpy(&r, t->dRowsAcc.p, 8, cudaMemcpyDeviceToHost) == cudaSuccess) t->physRows = (int64_t)r; }
      t->tTotal = now_s() - t->tStart;
      static const bool timing = [] { const char* m = std::getenv("B2RF_TIMING"); return m && *m == '1'; }();
      if (timing) fprintf(stderr, "[b2rf timing] groups %d host select %.4f tail %.4f wait-in-end_group %.4f total %.4f\n", t->groupNo, t->tSelect, t->tTail, t->tWait, t->tTotal);
    }
    return 0;
  }
  const double tc0 = now_s();

  // ---- RandomForest.selectNodesToSplit ----
  const int64_t maxMem = (int64_t)t->p.maxMemoryInMB * 1024 * 1024;
  t->group.clear(); t->groupTreeInsertOrder.clear();
  std::vector<uint32_t> feats;  // [slots][Fsub]
  std::vector<uint32_t> tmp(t->Fsub);
  std::vector<char> treeSeen(t->T, 0);
  int64_t mem = 0, cells = 0;
  std::vector<int64_t> slotBase;
  while (!t->stack.empty()) {
    const auto [tr, ni] = t->stack.back();
    int64_t totalBins, nodeCells;
    if (t->subsampling) {
      t->ensure_draws(1);                                                  // drawn before the memory check, consumed either way
      memcpy(tmp.data(), t->drawFeats.data() + t->drawHead * (size_t)t->Fsub, (size_t)t->Fsub * 4);
      totalBins = t->drawBins[t->drawHead];
      nodeCells = t->drawCells[t->drawHead];
      t->drawHead++;
    } else {
      for (int j = 0; j < t->Fsub; j++) tmp[j] = (uint32_t)j;
      totalBins = t->allBins;
      nodeCells = 0;
      for (int f = 0; f < t->F; f++) nodeCells += t->cellBins[f];
    }
    const int64_t nodeMem = (t->p.impurity == 0 ? 3 : (t->C ? t->C : 2)) * totalBins * 8;   // aggregateSizeForNode * 8: 3 stats (regression) or numClasses
    if (!(mem + nodeMem <= maxMem || mem == 0)) break;
    t->stack.pop_back();
    if (!treeSeen[tr]) { treeSeen[tr] = 1; t->groupTreeInsertOrder.push_back(tr); }
    t->group.push_back({tr, ni});
    feats.insert(feats.end(), tmp.begin(), tmp.end());
    slotBase.push_back(cells);
    cells += nodeCells;
    mem += nodeMem;
  }
  const size_t nSlots = t->group.size();
  t->shardSpr = 0; t->shardChunk = 0;
  if (t->shardWorld > 1 && t->segMode) {
    // reduce-scatter by node: contiguous slot ranges of balanced CELL counts (a range ends with the first slot that carries its prefix
    // past r/W of the group's cells, so no range exceeds cells / W by more than one node), every range padded to the largest one
    const int64_t W = t->shardWorld;
    t->shardLo.assign((size_t)W + 1, (int64_t)nSlots);
    t->shardLo[0] = 0;
    {
      int64_t r = 1;
      for (size_t sl = 0; sl < nSlots && r < W; sl++) {
        const int64_t end = sl + 1 < nSlots ? slotBase[sl + 1] : cells;     // cells up to and including slot sl
        while (r < W && end * W >= r * cells) t->shardLo[(size_t)r++] = (int64_t)sl + 1;
      }
    }
    int64_t chunk = 0, spr = 0;
    for (int64_t r = 0; r < W; r++) {
      const int64_t lo = t->shardLo[(size_t)r], hi = t->shardLo[(size_t)r + 1];
      if (hi > lo) chunk = std::max(chunk, (hi < (int64_t)nSlots ? slotBase[hi] : cells) - slotBase[lo]);
      spr = std::max(spr, hi - lo);
    }
    for (int64_t r = W - 1; r >= 0; r--) {
      const int64_t lo = t->shardLo[(size_t)r], hi = t->shardLo[(size_t)r + 1];
      const int64_t first = hi > lo ? slotBase[lo] : 0;
      for (int64_t sl = lo; sl < hi; sl++) slotBase[sl] = r * chunk + (slotBase[sl] - first);
    }
    cells = W * chunk;
    t->shardSpr = spr; t->shardChunk = chunk;
  }
  if (cells * t->M > t->histCells) { set_error("b2rf_trainer_begin_group: histogram buffer too small for this node group"); return B2RF_ERR_ARG; }
  t->groupCells = cells;

  // ---- descriptors ----
  if (!t->hSlotFo.ensure(nSlots * t->Fsub) || !t->hSlotBase.ensure(nSlots) || !t->hParent.ensure(nSlots * 4) || !t->hHasParent.ensure(nSlots) ||
      !t->dSlotFo.ensure(nSlots * t->Fsub) || !t->dSlotBase.ensure(nSlots) || !t->dParent.ensure(nSlots * 4) || !t->dHasParent.ensure(nSlots) ||
      (t->C && (!t->hParentCls.ensure(nSlots * t->Cp) || !t->dParentCls.ensure(nSlots * t->Cp) || !t->hClsOut.ensure(nSlots * 3 * t->Cp) ||
                !t->dClsOut.ensure(nSlots * 3 * t->Cp))) ||
      !t->hSlotCells.ensure(nSlots) || !t->dRec.ensure(nSlots) || !t->hRec.ensure(nSlots) || !t->dRecMask.ensure(nSlots * t->maskWords) || !t->hRecMask.ensure(nSlots * t->maskWords) || !t->hTreeIds.ensure(t->T) || !t->dTreeIds.ensure(t->T))
    return B2RF_ERR_CUDA;
  int64_t maxNodes = 0;
  for (size_t sl = 0; sl < nSlots; sl++) {
    const GNode& g = t->group[sl];
    const HNode& n = t->trees[g.tree][g.nidx];
    uint32_t off = 0;
    for (int j = 0; j < t->Fsub; j++) {
      const uint32_t f = feats[sl * t->Fsub + j];
      t->hSlotFo.p[sl * t->Fsub + j] = make_uint2(f, off);
      off += (uint32_t)t->cellBins[f];
    }
    t->hSlotCells.p[sl] = (int32_t)off;
    t->hSlotBase.p[sl] = slotBase[sl];
    const bool hasParent = level_of(n.id) > 0;
    t->hHasParent.p[sl] = hasParent;
    for (int q = 0; q < 4; q++) t->hParent.p[sl * 4 + q] = hasParent ? n.stats.calc[q] : 0.0;
    if (t->C) for (int k = 0; k < t->Cp; k++) t->hParentCls.p[sl * t->Cp + k] = (hasParent && k < t->C && n.cls >= 0) ? t->clsPool[(size_t)n.cls + k] : 0.0;
    t->pending.push_back({(int64_t)g.tree * t->nodeStride + g.nidx, t->mirror(g.tree, g.nidx, (int)sl)});
    maxNodes = std::max<int64_t>(maxNodes, (int64_t)t->trees[g.tree].size());
  }
  (void)maxNodes;
  // trees of this group, ascending; scan mode cuts them into launches whose histogram cells stay L2-resident
  std::vector<int64_t> cellsPerTree;
  if (!t->bucketMode) {
    cellsPerTree.assign(t->T, 0);
    for (size_t sl = 0; sl < nSlots; sl++) cellsPerTree[t->group[sl].tree] += (sl + 1 < nSlots ? slotBase[sl + 1] : cells) - slotBase[sl];
  }
  int nTg = 0;
  for (int i = 0; i < t->T; i++) if (treeSeen[i]) t->hTreeIds.p[nTg++] = i;

  // ---- uploads ----
  bool ok = true;
  { const int rc = flush_node_updates(t, s); if (rc) return rc; }
  ok &= cuda_ok(cudaMemcpyAsync(t->dSlotFo.p, t->hSlotFo.p, nSlots * t->Fsub * sizeof(uint2), cudaMemcpyHostToDevice, s), "upload slot_fo");
  ok &= cuda_ok(cudaMemcpyAsync(t->dSlotBase.p, t->hSlotBase.p, nSlots * 8, cudaMemcpyHostToDevice, s), "upload slot_base");
  ok &= cuda_ok(cudaMemcpyAsync(t->dParent.p, t->hParent.p, nSlots * 32, cudaMemcpyHostToDevice, s), "upload parent stats");
  ok &= cuda_ok(cudaMemcpyAsync(t->dHasParent.p, t->hHasParent.p, nSlots * 4, cudaMemcpyHostToDevice, s), "upload hasParent");
  ok &= cuda_ok(cudaMemcpyAsync(t->dTreeIds.p, t->hTreeIds.p, (size_t)nTg * 4, cudaMemcpyHostToDevice, s), "upload tree ids");
  if (t->C) ok &= cuda_ok(cudaMemcpyAsync(t->dParentCls.p, t->hParentCls.p, nSlots * t->Cp * 8, cudaMemcpyHostToDevice, s), "upload parent class counts");
  ok &= cuda_ok(cudaMemsetAsync(t->d_hist, 0, (size_t)cells * 32 * t->M, s), "zero histogram");
  if (!ok) return B2RF_ERR_CUDA;
  t->tChoose += now_s() - tc0; t->tSelect += now_s() - tc0;

  // ---- K4 launches (findBestSplits.binSeqOp over all local rows) ----
  cudaEventRecord(t->ev0, s);
  HistLaunch a;
  a.bins = t->d_bins; a.ldb = t->ldb; a.F = t->F; a.y = t->d_y; a.w = t->d_w; a.ldw = t->ldw; a.node = t->d_node; a.ldn = t->ldn;
  a.n_rows = t->nLocal; a.dnodes = t->dn.p; a.nodeStride = t->nodeStride; a.catmasks = t->dCat.p;
  a.slot_fo = t->dSlotFo.p; a.slot_base = t->dSlotBase.p; a.Fsub = t->Fsub; a.hist = t->d_hist; a.slot_count = nullptr;
  a.codeBytes = t->codeBytes; a.maskWords = t->maskWords;
  if (t->bucketMode) {
    // shared-memory histograms per (slot, chunk).  Segment mode: every pending node already owns a contiguous row segment
    // (kept by the stable partitions launched from end_group).  Sort mode: counting sort of the group's rows by slot.
    int64_t listCap = 0;
    for (int i = 0; i < nTg; i++) listCap += t->inBag[t->hTreeIds.p[i]];
    int32_t maxCells = 0;
    if (!t->hSlotCells.ensure(nSlots) || !t->dSlotCells.ensure(nSlots) || !t->dChunkPrefix.ensure(nSlots + 2)) return B2RF_ERR_CUDA;
    if (t->segMode) {
      if (!t->hSlotPos.ensure(nSlots) || !t->dSlotPos.ensure(nSlots)) return B2RF_ERR_CUDA;
      for (size_t sl = 0; sl < nSlots; sl++) t->hSlotPos.p[sl] = (int64_t)t->group[sl].tree * t->nodeStride + t->group[sl].nidx;
      ok &= cuda_ok(cudaMemcpyAsync(t->dSlotPos.p, t->hSlotPos.p, nSlots * 8, cudaMemcpyHostToDevice, s), "upload slot positions");
    } else if (!t->dSlotCount.ensure(nSlots) || !t->dCursor.ensure(nSlots) || !t->dSlotStart.ensure(nSlots + 1) ||
               !t->dRowList.ensure((size_t)std::max<int64_t>(listCap, 1)))
      return B2RF_ERR_CUDA;
    for (size_t sl = 0; sl < nSlots; sl++) maxCells = std::max(maxCells, t->hSlotCells.p[sl]);
    // shared-memory layout: padded [bin][feature-in-group] with the widest power-of-two feature group G that keeps a task's
    // histogram (maxBins * G cells) within ~48 KB, i.e. four or more CTAs per SM; compact cells only if even G = 4 is too big
    const int32_t maxBinsAll = *std::max_element(t->cellBins.begin(), t->cellBins.end());
    int32_t padB = 0, G = 32;
    static const bool noPad = [] { const char* m = std::getenv("B2RF_NO_PAD"); return m && *m == '1'; }();
    // budget: 48 KB (four 256-thread CTAs per SM); histograms with more than 128 bins get 100 KB (two 512-thread CTAs) — measured
    // on 12.5M x 128, 256 bins: G = 8 / 16 / 32 -> K4 1.34 / 1.03 / 1.69 s
    const int64_t histKB = maxBinsAll > 128 ? 96 : 48;   // + staging tables + 1 KB reserve: 2 resp. 4 CTAs per SM
    while (G > 4 && ((int64_t)maxBinsAll * G * 20 + 16 > histKB * 1024 || G / 2 >= t->Fsub)) G /= 2;
    if (!noPad && (int64_t)maxBinsAll * G * 20 + 16 <= (int64_t)200 * 1024) { padB = maxBinsAll; maxCells = maxBinsAll * G; }
    if ((size_t)maxCells * 20 + 16 > (size_t)208 * 1024) maxCells = (int32_t)(((size_t)208 * 1024 - 16) / 20);   // larger slots take the direct path
    ok &= cuda_ok(cudaMemcpyAsync(t->dSlotCells.p, t->hSlotCells.p, nSlots * 4, cudaMemcpyHostToDevice, s), "upload slot cells");
    if (!t->segMode) ok &= cuda_ok(cudaMemsetAsync(t->dSlotCount.p, 0, nSlots * 4, s), "zero slot counts");
    if (!ok) return B2RF_ERR_CUDA;
    a.tree_ids = t->dTreeIds.p; a.Tg = nTg; a.slot_count = t->dSlotCount.p;
    BucketLaunch b;
    b.bins_rm = t->d_bins_rm; b.ldr = t->ldr; b.y = t->d_y; b.rowlist = t->dRowList.p; b.slot_start = t->dSlotStart.p; b.slot_count = t->dSlotCount.p;
    b.chunk_prefix = t->dChunkPrefix.p; b.nSlots = (int32_t)nSlots; b.CH = t->segMode ? SEG_CH : t->chunkRows; b.slot_fo = t->dSlotFo.p;
    b.slot_base = t->dSlotBase.p; b.slot_cells = t->dSlotCells.p; b.Fsub = t->Fsub; b.maxCells = maxCells; b.padB = padB; b.G = G; b.codeBytes = t->codeBytes; b.hist = t->d_hist;
    { static const bool zr = [] { const char* m = std::getenv("B2RF_ZERO_REG"); return !(m && *m == '0'); }(); b.zeroReg = zr ? 1 : 0; }
    b.dseg = nullptr; b.slot_pos = nullptr; b.idx0 = nullptr; b.idx1 = nullptr; b.y0 = nullptr; b.y1 = nullptr; b.rows_acc = t->dRowsAcc.p; b.slots = nullptr;
    std::string err;
    cudaError_t e;
    if (t->C) {   // multiclass: class-count histograms (segment mode only, checked in bind)
      ClsHistLaunch c;
      c.bins_rm = t->d_bins_rm; c.ldr = t->ldr; c.codeBytes = t->codeBytes; c.y = t->d_y; c.chunk_prefix = t->dChunkPrefix.p;
      c.nSlots = (int32_t)nSlots; c.CH = SEG_CH; c.slot_fo = t->dSlotFo.p; c.slot_base = t->dSlotBase.p; c.Fsub = t->Fsub;
      c.padB = maxBinsAll; c.C = t->C; c.M = t->M; c.direct = 0; c.hist = t->d_hist; c.rows_acc = t->dRowsAcc.p;
      c.dseg = t->dSeg.p; c.slot_pos = t->dSlotPos.p; c.idx0 = t->dIdx0.p; c.idx1 = t->dIdx1.p;
      c.y0 = t->carryY ? t->dY0.p : nullptr; c.y1 = t->carryY ? t->dY1.p : nullptr; c.slot_cells = t->dSlotCells.p;
      { static const bool gather = [] { const char* m = std::getenv("B2RF_K4C"); return m && !std::strcmp(m, "gather"); }(); c.staged = gather ? 0 : 1; }
      e = launch_cls_hist(c, s);
      t->nKernels += 2;
    } else if (t->segMode) {
      b.dseg = t->dSeg.p; b.slot_pos = t->dSlotPos.p; b.idx0 = t->dIdx0.p; b.idx1 = t->dIdx1.p;
      if (t->carryY) { b.y0 = t->dY0.p; b.y1 = t->dY1.p; }
      // default: limb kernel (native 32-bit shared-memory adds on fixed-point limbs, k4_limb.cu); B2RF_K4=fp64 keeps the
      // compare-and-swap kernels, which also serve what the limb kernel does not (uint16 codes, labels not carried, huge histograms)
      const bool fp64Only = opt_k4_fp64().load() != 0;
      LimbPlan plan;
      t->deferred = false;
      if (!fp64Only && limb_plan(b, maxBinsAll, t->labels01 && opt_k4_bin().load() ? 2 : (t->labelsInt ? 1 : 0), opt_k4_tma().load() != 0, &plan)) {
        if (t->deferK4) { t->deferB = b; t->deferPlan = plan; t->deferred = true; e = cudaSuccess; }   // launched by b2rf_trainers_begin_group_shared
        else e = launch_limb_hist(b, plan, s);
      } else e = launch_segment_hist(b, s);
      t->nKernels += 2;
    } else {
      e = launch_bucket_pipeline(a, b, t->dCursor.p, s, &err);
      t->nKernels += 4;
    }
    if (e != cudaSuccess) { if (!err.empty()) set_error(err); else cuda_ok(e, "histogram launch"); return B2RF_ERR_CUDA; }
    t->nHistLaunches++;
    t->nScanned += (int64_t)nTg * t->nLocal;
    t->histAlgBytes += (double)t->nLocal * ((double)t->F + 8.0 + 3.0 * nTg) + 20.0 * (double)cells;
  } else {
    int start = 0;
    while (start < nTg) {
      int end = start; int64_t c = 0;
      while (end < nTg && (end == start || c + cellsPerTree[t->hTreeIds.p[end]] <= t->l2CellBudget)) { c += cellsPerTree[t->hTreeIds.p[end]]; end++; }
      a.tree_ids = t->dTreeIds.p + start; a.Tg = end - start;
      std::string err;
      cudaError_t e = launch_hist(a, 1, s, &err);
      if (e != cudaSuccess) { if (!err.empty()) set_error(err); else cuda_ok(e, "hist launch"); return B2RF_ERR_CUDA; }
      t->nHistLaunches++; t->nKernels++;
      t->nScanned += (int64_t)(end - start) * t->nLocal;
      // algorithmic bytes of this launch: bin matrix + labels + (bag weight u8 + node slot u16) per active tree + histogram cells*20
      t->histAlgBytes += (double)t->nLocal * ((double)t->F + 8.0 + 3.0 * (end - start)) + 20.0 * (double)c;
      start = end;
    }
  }
  cudaEventRecord(t->ev1, s);
  t->evPending = true;
  t->inGroup = true;
  if (n_cells) *n_cells = cells * t->M;   // in units of 4 doubles: what the caller sums across ranks
  return 1;
}

}  // extern "C"

// end_group, first half: K5 + the copies of its results, all asynchronous on the stream.
// K5 over the node slots [lo, lo + cnt) of the group in flight (the whole group unless the fit is sharded): every per-slot array is
// passed from its lo-th entry, so the kernels see an ordinary group of cnt slots.
static int end_group_k5(b2rf_trainer* t, cudaStream_t s, size_t lo, size_t cnt) {
  if (cnt == 0) return B2RF_OK;
  SplitLaunch a;
  a.hist = t->d_hist; a.slot_fo = t->dSlotFo.p + lo * t->Fsub; a.slot_base = t->dSlotBase.p + lo; a.Fsub = t->Fsub; a.nSlots = (int32_t)cnt;
  a.parent4 = t->dParent.p + lo * 4; a.hasParent = t->dHasParent.p + lo; a.numBins = t->dNumBins.p; a.isCat = t->dIsCat.p;
  a.minInstances = (double)t->p.minInstancesPerNode; a.minWeight = t->minWeightPerNode; a.minInfoGain = t->p.minInfoGain;
  a.out = t->dRec.p + lo; a.catmask_out = t->dRecMask.p + lo * t->maskWords; a.maskWords = t->maskWords; a.impurity = t->p.impurity;
  a.maxB = *std::max_element(t->cellBins.begin(), t->cellBins.end());
  a.allHaveParent = 1;
  for (size_t sl = lo; sl < lo + cnt; sl++) if (!t->hHasParent.p[sl]) { a.allHaveParent = 0; break; }
  if (t->C) {
    ClsSplitLaunch c;
    c.hist = t->d_hist; c.slot_fo = a.slot_fo; c.slot_base = a.slot_base; c.Fsub = t->Fsub; c.nSlots = (int32_t)cnt;
    c.parentCls = t->dParentCls.p + lo * t->Cp; c.hasParent = a.hasParent; c.cellBins = t->dCellBins.p; c.kind = t->dKind.p;
    c.minInstances = a.minInstances; c.minWeight = a.minWeight; c.minInfoGain = a.minInfoGain;
    c.out = a.out; c.clsOut = t->dClsOut.p + lo * 3 * t->Cp; c.catmask_out = a.catmask_out;
    c.maxB = a.maxB; c.maskWords = t->maskWords; c.impurity = t->p.impurity; c.C = t->C; c.Cp = t->Cp; c.M = t->M;
    if (!cuda_ok(launch_best_split_cls(c, s), "best_split_cls launch")) return B2RF_ERR_CUDA;
  } else if (!cuda_ok(launch_best_split(a, s), "best_split launch")) return B2RF_ERR_CUDA;
  t->nKernels++;
  return B2RF_OK;
}

// The copies of K5's results for all slots of the group (asynchronous on the stream), plus host work that overlaps them.
static int end_group_download(b2rf_trainer* t, cudaStream_t s) {
  const size_t nSlots = t->group.size();
  if (t->C && !cuda_ok(cudaMemcpyAsync(t->hClsOut.p, t->dClsOut.p, nSlots * 3 * t->Cp * 8, cudaMemcpyDeviceToHost, s), "download class counts")) return B2RF_ERR_CUDA;
  if (!cuda_ok(cudaMemcpyAsync(t->hRec.p, t->dRec.p, nSlots * sizeof(SplitRec), cudaMemcpyDeviceToHost, s), "download split records")) return B2RF_ERR_CUDA;
  if (!cuda_ok(cudaMemcpyAsync(t->hRecMask.p, t->dRecMask.p, nSlots * t->maskWords * 4, cudaMemcpyDeviceToHost, s), "download split masks")) return B2RF_ERR_CUDA;
  const double tw0 = now_s();
  if (t->subsampling) {   // the next groups pop at most two children per node of this one (+1 rejected draw); the GPU is busy with K4 anyway
    size_t kids = 1;
    for (size_t sl = 0; sl < nSlots; sl++)
      if (level_of(t->trees[t->group[sl].tree][t->group[sl].nidx].id) + 1 < t->p.maxDepth) kids += 2;
    t->ensure_draws(kids + t->stack.size());
  }
  t->tWait -= tw0;   // completed by end_group_tail
  return B2RF_OK;
}

// end_group, first half: K5 + the copies of its results, all asynchronous on the stream.
static int end_group_launch(b2rf_trainer* t, cudaStream_t s) {
  { const int rc = end_group_k5(t, s, 0, t->group.size()); if (rc) return rc; }
  return end_group_download(t, s);
}

// ---- sharded K5: the records of the own slot range are packed into one block, all-gathered by the caller, and unpacked ----
static size_t shard_block_bytes(const b2rf_trainer* t) {
  const size_t perSlot = sizeof(SplitRec) + (size_t)t->maskWords * 4 + (t->C ? (size_t)3 * t->Cp * 8 : 0);
  return ((size_t)t->shardSpr * perSlot + 7) / 8 * 8;
}

constexpr int SHARD_MAX_WORLD = 64;
struct ShardRanges { int32_t world; int32_t lo[SHARD_MAX_WORLD + 1]; };

__global__ void unpack_records_kernel(const uint32_t* x, int64_t blockWords, int64_t spr, int64_t nSlots, ShardRanges rg, int recW, int maskW, int clsW,
                                      uint32_t* rec, uint32_t* mask, uint32_t* cls) {
  const int perSlot = recW + maskW + clsW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nSlots * perSlot; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t sl = i / perSlot;
    const int w = (int)(i - sl * perSlot);
    int r = 0;
    while (r + 1 < rg.world && sl >= rg.lo[r + 1]) r++;
    const int64_t k = sl - rg.lo[r];
    const uint32_t* b = x + (int64_t)r * blockWords;           // [spr x rec][spr x mask][spr x cls]
    if (w < recW) rec[sl * recW + w] = b[k * recW + w];
    else if (w < recW + maskW) mask[sl * maskW + (w - recW)] = b[spr * recW + k * maskW + (w - recW)];
    else cls[sl * clsW + (w - recW - maskW)] = b[spr * (recW + maskW) + k * clsW + (w - recW - maskW)];
  }
}

// end_group, second half (after the stream has been synchronised): driver tail + K6 partitions.
static int end_group_tail(b2rf_trainer* t, cudaStream_t s) {
  const size_t nSlots = t->group.size();
  const double tc0 = now_s();
  t->tWait += tc0;
  if (t->evPending) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, t->ev0, t->ev1) == cudaSuccess) t->tHist += ms * 1e-3;
    t->evPending = false;
    static const bool timing = [] { const char* m = std::getenv("B2RF_TIMING"); return m && *m == '2'; }();
    if (timing) fprintf(stderr, "[b2rf timing] group %d slots %zu cells %lld K4 %.3f ms\n", t->groupNo, nSlots, (long long)t->groupCells, ms);
  }

  // ---- driver tail of findBestSplits: trees in Scala map order, nodes in pop order ----
  std::vector<int> order;
  group_tree_order(t->groupTreeInsertOrder, order);
  std::vector<std::vector<int>> slotsOfTree(t->T);
  for (size_t sl = 0; sl < nSlots; sl++) slotsOfTree[t->group[sl].tree].push_back((int)sl);
  int64_t needStride = t->nodeStride;
  std::vector<std::pair<int32_t, std::pair<int32_t, int32_t>>> splits;   // (tree, (node, left child)) of the nodes split in this group
  int64_t splitRowsBound = 0;
  for (int tr : order) {
    std::vector<HNode>& nodes = t->trees[tr];
    for (int sl : slotsOfTree[tr]) {
      const SplitRec& r = t->hRec.p[sl];
      const int ni = t->group[sl].nidx;
      const int id = nodes[ni].id;
      const int lvl = level_of(id);
      const bool isLeaf = (r.gain <= 0) || (lvl == t->p.maxDepth);
      HNode& n = nodes[ni];
      n.isLeaf = isLeaf;
      n.stats.gain = r.gain; n.stats.impurity = r.impurity; n.stats.valid = r.valid != 0;
      for (int q = 0; q < 4; q++) n.stats.calc[q] = r.parent[q];
      const double* cls3 = t->C ? t->hClsOut.p + (size_t)sl * 3 * t->Cp : nullptr;
      if (t->C) { n.cls = (int64_t)t->clsPool.size(); t->clsPool.insert(t->clsPool.end(), cls3, cls3 + t->C); }
      if (t->keepTrace) t->trace.push_back({t->groupNo, tr, id, r.feature, r.splitIdx, r.gain, r.impurity, r.parent[0], r.parent[1], r.parent[2]});
      t->weightedRows += r.parent[0];
      if (!isLeaf) {
        n.hasSplit = true; n.feature = r.feature;
        if (t->arity[r.feature] == 0) { n.splitBin = r.splitIdx; n.threshold = t->thresholds[(size_t)r.feature * t->maxSplits + r.splitIdx]; }
        else {
          const uint32_t* m = t->hRecMask.p + (size_t)sl * t->maskWords;
          n.catMaskIdx = (int32_t)(t->catPool.size() / t->maskWords);
          t->catPool.insert(t->catPool.end(), m, m + t->maskWords);
          n.splitBin = r.splitIdx;
        }
        const bool childLeaf = (lvl + 1) == t->p.maxDepth;
        const bool lLeaf = childLeaf || std::fabs(r.leftImp) < 0x1.0p-52;
        const bool rLeaf = childLeaf || std::fabs(r.rightImp) < 0x1.0p-52;
        HNode L, R;
        L.id = 2 * id; L.isLeaf = lLeaf; L.stats.gain = NAN; L.stats.impurity = r.leftImp; for (int q = 0; q < 4; q++) L.stats.calc[q] = r.left[q];
        R.id = 2 * id + 1; R.isLeaf = rLeaf; R.stats.gain = NAN; R.stats.impurity = r.rightImp; for (int q = 0; q < 4; q++) R.stats.calc[q] = r.right[q];
        if (t->C) {
          L.cls = (int64_t)t->clsPool.size(); t->clsPool.insert(t->clsPool.end(), cls3 + t->Cp, cls3 + t->Cp + t->C);
          R.cls = (int64_t)t->clsPool.size(); t->clsPool.insert(t->clsPool.end(), cls3 + 2 * t->Cp, cls3 + 2 * t->Cp + t->C);
        }
        const int li = (int)nodes.size();
        nodes.push_back(L); nodes.push_back(R);
        nodes[ni].left = li;
        needStride = std::max<int64_t>(needStride, (int64_t)nodes.size());
        t->pending.push_back({(int64_t)tr * t->nodeStride + ni, t->mirror(tr, ni, SLOT_PENDING)});
        t->pending.push_back({(int64_t)tr * t->nodeStride + li, t->mirror(tr, li, lLeaf ? SLOT_DEAD : SLOT_PENDING)});
        t->pending.push_back({(int64_t)tr * t->nodeStride + li + 1, t->mirror(tr, li + 1, rLeaf ? SLOT_DEAD : SLOT_PENDING)});
        if (!lLeaf) t->stack.push_back({tr, li});
        if (!rLeaf) t->stack.push_back({tr, li + 1});
        if (!lLeaf || !rLeaf) splits.push_back({tr, {ni, li}});   // children that are both leaves never need their rows again
      } else {
        t->pending.push_back({(int64_t)tr * t->nodeStride + ni, t->mirror(tr, ni, SLOT_DEAD)});
      }
    }
  }
  if (needStride > t->nodeStride) { const int rc = grow_node_table(t, needStride, s); if (rc) return rc; }
  if (t->segMode && !splits.empty()) {
    // K6 in segment mode: stable partition of every split node's row segment into its children's segments
    { const int rc = flush_node_updates(t, s, /*tail=*/true); if (rc) return rc; }   // the partition kernels read the parents' split descriptors
    const size_t k = splits.size();
    std::vector<char> seen(t->T, 0);
    for (auto& sp : splits) if (!seen[sp.first]) { seen[sp.first] = 1; splitRowsBound += t->inBag[sp.first]; }
    const size_t maxTasks = (size_t)(splitRowsBound / SEG_CH) + k + 1;
    if (!t->hSplitPos.ensure(2 * k) || !t->dSplitPos.ensure(2 * k) || !t->dPartPrefix.ensure(k + 1) || !t->dLeftCnt.ensure(maxTasks) ||
        !t->dLeftOff.ensure(maxTasks) || !t->dPredBits.ensure(maxTasks * (SEG_CH / 32)))
      return B2RF_ERR_CUDA;
    for (size_t i = 0; i < k; i++) {
      t->hSplitPos.p[i] = (int64_t)splits[i].first * t->nodeStride + splits[i].second.first;
      t->hSplitPos.p[k + i] = (int64_t)splits[i].first * t->nodeStride + splits[i].second.second;
    }
    if (!cuda_ok(cudaMemcpyAsync(t->dSplitPos.p, t->hSplitPos.p, 2 * k * 8, cudaMemcpyHostToDevice, s), "upload split list")) return B2RF_ERR_CUDA;
    PartLaunch pl;
    pl.split_pos = t->dSplitPos.p; pl.nSplits = (int32_t)k; pl.dnodes = t->dn.p; pl.dseg = t->dSeg.p; pl.catmasks = t->dCat.p;
    pl.bins = t->d_bins; pl.ldb = t->ldb; pl.idx0 = t->dIdx0.p; pl.idx1 = t->dIdx1.p; pl.chunk_prefix = t->dPartPrefix.p;
    pl.y0 = t->carryY ? t->dY0.p : nullptr; pl.y1 = t->carryY ? t->dY1.p : nullptr;
    pl.left_cnt = t->dLeftCnt.p; pl.left_off = t->dLeftOff.p; pl.predbits = t->dPredBits.p; pl.codeBytes = t->codeBytes; pl.maskWords = t->maskWords;
    if (!cuda_ok(launch_partition(pl, s), "partition launch")) return B2RF_ERR_CUDA;
    t->nKernels += 4;
  }
  t->inGroup = false;
  t->groupNo++;
  t->exported = false;
  t->tChoose += now_s() - tc0; t->tTail += now_s() - tc0;
  return B2RF_OK;
}

extern "C" {

int b2rf_trainer_end_group(b2rf_trainer* t, void* stream) {
  if (!t || !t->inGroup) { set_error("b2rf_trainer_end_group: no group in flight"); return B2RF_ERR_STATE; }
  cudaStream_t s = (cudaStream_t)stream;
  { const int rc = end_group_launch(t, s); if (rc) return rc; }
  if (!cuda_ok(cudaStreamSynchronize(s), "group synchronize")) return B2RF_ERR_CUDA;
  return end_group_tail(t, s);
}

// ---- rows sharded over several processes: reduce-scatter by node, K5 on the own slots, all-gather of the split records ----
int b2rf_trainer_set_shard(b2rf_trainer* t, int32_t rank, int32_t world) {
  if (!t || world < 1 || world > SHARD_MAX_WORLD || rank < 0 || rank >= world) { set_error("b2rf_trainer_set_shard: bad argument (1 <= world <= 64)"); return B2RF_ERR_ARG; }
  if (t->d_bins) { set_error("b2rf_trainer_set_shard: call before b2rf_trainer_bind (the histogram buffer is sized for the padded layout)"); return B2RF_ERR_STATE; }
  t->shardRank = rank; t->shardWorld = world;
  return B2RF_OK;
}

int64_t b2rf_trainer_shard_block_doubles(const b2rf_trainer* t) {
  if (!t || !t->inGroup || t->shardSpr <= 0) return 0;
  return (int64_t)(shard_block_bytes(t) / 8);
}

int b2rf_trainer_end_group_shard_launch(b2rf_trainer* t, void* d_block, void* stream) {
  if (!t || !t->inGroup) { set_error("b2rf_trainer_end_group_shard_launch: no group in flight"); return B2RF_ERR_STATE; }
  if (t->shardSpr <= 0 || !d_block) { set_error("b2rf_trainer_end_group_shard_launch: the group was not laid out for sharding (b2rf_trainer_set_shard, segment mode)"); return B2RF_ERR_STATE; }
  cudaStream_t s = (cudaStream_t)stream;
  const size_t spr = (size_t)t->shardSpr;
  const size_t lo = (size_t)t->shardLo[(size_t)t->shardRank], cnt = (size_t)t->shardLo[(size_t)t->shardRank + 1] - lo;
  { const int rc = end_group_k5(t, s, lo, cnt); if (rc) return rc; }
  unsigned char* b = static_cast<unsigned char*>(d_block);
  const size_t recB = sizeof(SplitRec), maskB = (size_t)t->maskWords * 4, clsB = t->C ? (size_t)3 * t->Cp * 8 : 0;
  bool ok = true;
  if (cnt) {
    ok &= cuda_ok(cudaMemcpyAsync(b, t->dRec.p + lo, cnt * recB, cudaMemcpyDeviceToDevice, s), "pack split records");
    ok &= cuda_ok(cudaMemcpyAsync(b + spr * recB, t->dRecMask.p + lo * t->maskWords, cnt * maskB, cudaMemcpyDeviceToDevice, s), "pack split masks");
    if (clsB) ok &= cuda_ok(cudaMemcpyAsync(b + spr * (recB + maskB), t->dClsOut.p + lo * 3 * t->Cp, cnt * clsB, cudaMemcpyDeviceToDevice, s), "pack class counts");
  }
  return ok ? B2RF_OK : B2RF_ERR_CUDA;
}

int b2rf_trainer_end_group_shard_finish(b2rf_trainer* t, const void* d_blocks, void* stream) {
  if (!t || !t->inGroup || t->shardSpr <= 0 || !d_blocks) { set_error("b2rf_trainer_end_group_shard_finish: no sharded group in flight"); return B2RF_ERR_STATE; }
  cudaStream_t s = (cudaStream_t)stream;
  const size_t nSlots = t->group.size();
  const int recW = (int)(sizeof(SplitRec) / 4), maskW = t->maskWords, clsW = t->C ? 3 * t->Cp * 2 : 0;
  const int64_t total = (int64_t)nSlots * (recW + maskW + clsW);
  const int grid = (int)std::min<int64_t>((total + 255) / 256, 4096);
  ShardRanges rg;
  rg.world = t->shardWorld;
  for (int r = 0; r <= t->shardWorld; r++) rg.lo[r] = (int32_t)t->shardLo[(size_t)r];
  unpack_records_kernel<<<grid, 256, 0, s>>>(static_cast<const uint32_t*>(d_blocks), (int64_t)(shard_block_bytes(t) / 4), t->shardSpr, (int64_t)nSlots, rg, recW, maskW,
                                               clsW, reinterpret_cast<uint32_t*>(t->dRec.p), t->dRecMask.p, reinterpret_cast<uint32_t*>(t->dClsOut.p));
  if (!cuda_ok(cudaGetLastError(), "unpack split records")) return B2RF_ERR_CUDA;
  t->nKernels++;
  { const int rc = end_group_download(t, s); if (rc) return rc; }
  if (!cuda_ok(cudaStreamSynchronize(s), "group synchronize")) return B2RF_ERR_CUDA;
  return end_group_tail(t, s);
}

// ---- fitMultiple shared passes (`ML 07 - Random Forests and Hyperparameter Tuning.py:74-77,143-149`) ----
// begin_group of every listed trainer (host scheduling + descriptor uploads), then ONE K4 launch whose task table spans the node
// slots of all of them: a shallow candidate's few nodes ride with the deep one's.  Trainers whose launch shape differs (other bin
// matrix, features per node, bin count, label form) or that do not run the limb kernel get their own launch.
int b2rf_trainers_begin_group_shared(b2rf_trainer** ts, int32_t n, int64_t* n_cells, int32_t* started, void* stream) {
  if (!ts || n <= 0 || !n_cells || !started) { set_error("b2rf_trainers_begin_group_shared: bad argument"); return B2RF_ERR_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  std::vector<b2rf_trainer*> def;
  for (int i = 0; i < n; i++) {
    b2rf_trainer* t = ts[i];
    if (!t) { set_error("b2rf_trainers_begin_group_shared: null trainer"); return B2RF_ERR_ARG; }
    t->deferK4 = true; t->deferred = false;
    const int rc = b2rf_trainer_begin_group(t, &n_cells[i], stream);
    t->deferK4 = false;
    if (rc < 0) return rc;
    started[i] = rc;
    if (rc == 1 && t->deferred) def.push_back(t);
  }
  auto same = [](const b2rf_trainer* a, const b2rf_trainer* b) {
    const BucketLaunch &x = a->deferB, &y = b->deferB; const LimbPlan &p = a->deferPlan, &q = b->deferPlan;
    return x.bins_rm == y.bins_rm && x.ldr == y.ldr && x.Fsub == y.Fsub && x.CH == y.CH && p.G == q.G && p.BR == q.BR && p.threads == q.threads &&
           p.padB == q.padB && p.NP == q.NP && p.tma == q.tma;
  };
  std::vector<char> done(def.size(), 0);
  for (size_t i = 0; i < def.size(); i++) {
    if (done[i]) continue;
    std::vector<b2rf_trainer*> cls{def[i]};
    done[i] = 1;
    for (size_t j = i + 1; j < def.size(); j++) if (!done[j] && same(def[i], def[j])) { cls.push_back(def[j]); done[j] = 1; }
    b2rf_trainer* o = cls[0];                       // owner of the merged slot table
    cudaError_t e;
    if (cls.size() == 1) e = launch_limb_hist(o->deferB, o->deferPlan, s);
    else {
      size_t total = 0;
      for (auto* t : cls) total += t->group.size();
      if (!o->hShared.ensure(total) || !o->dShared.ensure(total) || !o->dSharedPrefix.ensure(total + 2)) return B2RF_ERR_CUDA;
      size_t k = 0;
      for (auto* t : cls)
        for (size_t sl = 0; sl < t->group.size(); sl++) {
          SlotSrc& ss = o->hShared.p[k++];
          ss.seg = t->dSeg.p + t->hSlotPos.p[sl];
          ss.idx0 = t->dIdx0.p; ss.idx1 = t->dIdx1.p; ss.y0 = t->dY0.p; ss.y1 = t->dY1.p;
          ss.fo = t->dSlotFo.p + sl * (size_t)t->Fsub;
          ss.hb = t->d_hist + (size_t)t->hSlotBase.p[sl] * 4;
          ss.cells = t->hSlotCells.p[sl]; ss.pad = 0;
        }
      if (!cuda_ok(cudaMemcpyAsync(o->dShared.p, o->hShared.p, total * sizeof(SlotSrc), cudaMemcpyHostToDevice, s), "upload merged slot table")) return B2RF_ERR_CUDA;
      BucketLaunch b = o->deferB;
      b.slots = o->dShared.p; b.nSlots = (int32_t)total; b.chunk_prefix = o->dSharedPrefix.p; b.rows_acc = nullptr;
      e = launch_limb_hist(b, o->deferPlan, s);
    }
    if (!cuda_ok(e, "shared histogram launch")) return B2RF_ERR_CUDA;
    for (auto* t : cls) { cudaEventRecord(t->ev1, s); t->deferred = false; }   // each member's K4 time = the merged launch (shared, not additive)
  }
  return B2RF_OK;
}

// end_group of every listed trainer that has a group in flight: all K5 launches and result copies first, ONE synchronisation, then
// the driver tails (each followed by its K6 partitions, which run while the next tail is computed on the host).
int b2rf_trainers_end_group_shared(b2rf_trainer** ts, int32_t n, void* stream) {
  if (!ts || n <= 0) { set_error("b2rf_trainers_end_group_shared: bad argument"); return B2RF_ERR_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  for (int i = 0; i < n; i++) if (ts[i] && ts[i]->inGroup) { const int rc = end_group_launch(ts[i], s); if (rc) return rc; }
  if (!cuda_ok(cudaStreamSynchronize(s), "group synchronize")) return B2RF_ERR_CUDA;
  for (int i = 0; i < n; i++) if (ts[i] && ts[i]->inGroup) { const int rc = end_group_tail(ts[i], s); if (rc) return rc; }
  return B2RF_OK;
}

int b2rf_set_option(const char* name, int64_t value) {
  const std::string n = name ? name : "";
  if (n == "k4_fp64") { opt_k4_fp64().store(value ? 1 : 0); return B2RF_OK; }
  if (n == "carry_labels") { opt_carry_labels().store(value ? 1 : 0); return B2RF_OK; }
  if (n == "k4_tma") { opt_k4_tma().store(value ? 1 : 0); return B2RF_OK; }
  if (n == "k4_binary") { opt_k4_bin().store(value ? 1 : 0); return B2RF_OK; }
  set_error("b2rf_set_option: unknown option \"" + n + "\" (k4_fp64, k4_tma, carry_labels)");
  return B2RF_ERR_ARG;
}

int b2rf_trainer_group_info(const b2rf_trainer* t, int64_t* n_slots, int32_t* h_tree, int32_t* h_node_id, uint32_t* h_slot_fo, int64_t* h_slot_base) {
  if (!t || !t->inGroup) { set_error("b2rf_trainer_group_info: no group in flight"); return B2RF_ERR_STATE; }
  const size_t nSlots = t->group.size();
  if (n_slots) *n_slots = (int64_t)nSlots;
  for (size_t sl = 0; sl < nSlots; sl++) {
    if (h_tree) h_tree[sl] = t->group[sl].tree;
    if (h_node_id) h_node_id[sl] = t->trees[t->group[sl].tree][t->group[sl].nidx].id;
    if (h_slot_base) h_slot_base[sl] = t->hSlotBase.p[sl];
  }
  if (h_slot_fo) memcpy(h_slot_fo, t->hSlotFo.p, nSlots * (size_t)t->Fsub * sizeof(uint2));
  return B2RF_OK;
}

int b2rf_trainer_stats(const b2rf_trainer* t, double* h_seconds5, int64_t* h_counts4) {
  // h_counts4 has room for 9 entries (the last: weighted rows of all evaluated nodes): groups, hist launches, rows*trees scanned, learning nodes, kernel launches, hist algorithmic
  // bytes, rows visited by the segment-mode K4 launches (x featuresPerNode = physical bin updates), featuresPerNode
  if (!t) return B2RF_ERR_ARG;
  if (h_seconds5) { h_seconds5[0] = 0; h_seconds5[1] = 0; h_seconds5[2] = t->tHist; h_seconds5[3] = t->tChoose; h_seconds5[4] = t->tTotal > 0 ? t->tTotal : now_s() - t->tStart; }
  if (h_counts4) {
    int64_t nodes = 0; for (auto& tr : t->trees) nodes += (int64_t)tr.size();
    h_counts4[0] = t->groupNo; h_counts4[1] = t->nHistLaunches; h_counts4[2] = t->nScanned; h_counts4[3] = nodes;
    h_counts4[4] = t->nKernels; h_counts4[5] = (int64_t)t->histAlgBytes; h_counts4[6] = t->physRows; h_counts4[7] = t->Fsub; h_counts4[8] = (int64_t)t->weightedRows;
  }
  return B2RF_OK;
}

// LearningNode.toNode(prune = true): pre-order flattening; an internal node whose children are leaves with equal predictions
// collapses into a leaf.  Two passes per tree, trees in parallel on a few host threads:
//   1. (export_plan, cached) bottom-up over the node vector (children always sit behind their parent): does the node end up a leaf,
//      its prediction, the size of its flattened subtree and of its category lists; tree and category offsets of the forest;
//   2. (export_write) pre-order walk with an explicit stack that writes every field straight into the CALLER's arrays (node i of a
//      subtree at position p has its left child at p + 1 and its right child at p + 1 + size(left)); any array may be nullptr.
struct ExportOut {
  int32_t* feature; double* threshold; int32_t* left; int32_t* right; double* prediction; double* impurity; double* gain; double* stats4;
  int32_t* ncat; int32_t* cat_off; int32_t* cat_len; double* cats; double* cls;   // cls: [node][C] class counts (multiclass) or [node][2] (binary)
};

static void export_parallel(const b2rf_trainer* t, const std::function<void(int)>& fn) {
  const int T = t->T;
  size_t total = 0;
  for (int tr = 0; tr < T; tr++) total += t->trees[tr].size();
  const int workers = (int)std::max<size_t>(1, std::min<size_t>({(size_t)T, (size_t)8, total / 4096 + 1}));
  std::atomic<int> next{0};
  auto work = [&]() { for (int tr = next.fetch_add(1); tr < T; tr = next.fetch_add(1)) fn(tr); };
  std::vector<std::thread> pool;
  for (int w = 1; w < workers; w++) pool.emplace_back(work);
  work();
  for (auto& th : pool) th.join();
}

static void export_plan(b2rf_trainer* t) {
  if (t->exported) return;
  const int T = t->T;
  t->x_plan.assign(T, {});
  export_parallel(t, [&](int tr) {
    const std::vector<HNode>& nodes = t->trees[tr];
    const int n = (int)nodes.size();
    b2rf_trainer::TreePlan& p = t->x_plan[tr];
    p.leaf.resize(n); p.pred.resize(n); p.sz.resize(n); p.csz.resize(n);
    for (int i = n - 1; i >= 0; i--) {
      const HNode& nd = nodes[i];
      if (nd.left < 0) { p.leaf[i] = 1; p.pred[i] = t->node_pred(nd); p.sz[i] = 1; p.csz[i] = 0; continue; }
      const int l = nd.left, r = l + 1;
      if (p.leaf[l] && p.leaf[r] && p.pred[l] == p.pred[r]) { p.leaf[i] = 1; p.pred[i] = p.pred[l]; p.sz[i] = 1; p.csz[i] = 0; continue; }
      int own = 0;
      if (nd.catMaskIdx >= 0) { const uint32_t* m = &t->catPool[(size_t)nd.catMaskIdx * t->maskWords]; for (int q = 0; q < t->maskWords; q++) own += __builtin_popcount(m[q]); }
      p.leaf[i] = 0; p.pred[i] = t->node_pred(nd); p.sz[i] = 1 + p.sz[l] + p.sz[r]; p.csz[i] = own + p.csz[l] + p.csz[r];
    }
  });
  t->x_tree_off.assign(T + 1, 0);
  t->x_cat_base.assign(T + 1, 0);
  for (int tr = 0; tr < T; tr++) {
    t->x_tree_off[tr + 1] = t->x_tree_off[tr] + t->x_plan[tr].sz[0];
    t->x_cat_base[tr + 1] = t->x_cat_base[tr] + t->x_plan[tr].csz[0];
  }
  t->exported = true;
}

static void export_write(b2rf_trainer* t, const ExportOut& o) {
  export_plan(t);
  const int C = t->C;
  export_parallel(t, [&](int tr) {
    const std::vector<HNode>& nodes = t->trees[tr];
    const b2rf_trainer::TreePlan& p = t->x_plan[tr];
    const size_t base = (size_t)t->x_tree_off[tr];
    int64_t catAt = t->x_cat_base[tr];
    std::vector<std::pair<int32_t, int32_t>> stack;   // (node index, tree-local flat position)
    stack.reserve(64);
    stack.push_back({0, 0});
    while (!stack.empty()) {
      const auto [ni, pos] = stack.back();
      stack.pop_back();
      const HNode& nd = nodes[ni];
      const size_t at = base + (size_t)pos;
      if (o.stats4) for (int q = 0; q < 4; q++) o.stats4[at * 4 + q] = nd.stats.calc[q];
      if (o.cls) {
        if (C) for (int k = 0; k < C; k++) o.cls[at * (size_t)C + k] = nd.cls >= 0 ? t->clsPool[(size_t)nd.cls + k] : 0.0;
        else { o.cls[at * 2] = nd.stats.calc[0] - nd.stats.calc[1]; o.cls[at * 2 + 1] = nd.stats.calc[1]; }
      }
      const bool learnedLeaf = nd.left < 0, internal = !learnedLeaf && !p.leaf[ni];
      if (o.cat_off) o.cat_off[at] = (int32_t)catAt;
      if (o.prediction) o.prediction[at] = p.pred[ni];
      if (o.impurity) o.impurity[at] = learnedLeaf ? (nd.stats.valid ? nd.stats.impurity : -1.0) : nd.stats.impurity;
      const int l = nd.left, r = l + 1;
      if (o.feature) o.feature[at] = internal ? nd.feature : -1;
      if (o.left) o.left[at] = internal ? pos + 1 : -1;
      if (o.right) o.right[at] = internal ? pos + 1 + p.sz[l] : -1;
      if (o.threshold) o.threshold[at] = internal ? nd.threshold : 0.0;
      if (o.gain) o.gain[at] = internal ? nd.stats.gain : 0.0;
      int cnt = 0, ncat = 0;
      if (internal && nd.catMaskIdx >= 0) {
        ncat = t->arity[nd.feature];
        const uint32_t* m = &t->catPool[(size_t)nd.catMaskIdx * t->maskWords];
        for (int q = 0; q < t->maskWords; q++)
          for (uint32_t w = m[q]; w; w &= w - 1u) { if (o.cats) o.cats[(size_t)catAt] = (double)(q * 32 + __builtin_ctz(w)); catAt++; cnt++; }
      }
      if (o.ncat) o.ncat[at] = ncat;
      if (o.cat_len) o.cat_len[at] = cnt;
      if (!internal) continue;
      stack.push_back({r, pos + 1 + p.sz[l]});   // right first: the left subtree is popped (and its category lists placed) before it
      stack.push_back({l, pos + 1});
    }
  });
}

int b2rf_trainer_forest_sizes(b2rf_trainer* t, int64_t* n_nodes, int64_t* n_cats) {
  if (!t) return B2RF_ERR_ARG;
  export_plan(t);
  if (n_nodes) *n_nodes = t->x_tree_off[t->T];
  if (n_cats) *n_cats = t->x_cat_base[t->T];
  return B2RF_OK;
}

int b2rf_trainer_export(b2rf_trainer* t, int64_t* h_tree_off, int32_t* h_feature, double* h_threshold, int32_t* h_left, int32_t* h_right,
                        double* h_prediction, double* h_impurity, double* h_gain, double* h_stats4, int32_t* h_ncat, int32_t* h_cat_off,
                        int32_t* h_cat_len, double* h_cats) {
  if (!t) return B2RF_ERR_ARG;
  export_plan(t);
  if (h_tree_off) memcpy(h_tree_off, t->x_tree_off.data(), t->x_tree_off.size() * 8);
  ExportOut o{h_feature, h_threshold, h_left, h_right, h_prediction, h_impurity, h_gain, h_stats4, h_ncat, h_cat_off, h_cat_len, h_cats, nullptr};
  export_write(t, o);
  return B2RF_OK;
}

int b2rf_trainer_num_classes(const b2rf_trainer* t) { return t ? (t->C ? t->C : (t->p.impurity != 0 ? 2 : 0)) : 0; }

int b2rf_trainer_export_class_counts(b2rf_trainer* t, double* h_counts) {
  if (!t || !h_counts) { set_error("b2rf_trainer_export_class_counts: bad argument"); return B2RF_ERR_ARG; }
  if (t->p.impurity == 0) { set_error("b2rf_trainer_export_class_counts: this trainer fits a regression forest"); return B2RF_ERR_STATE; }
  ExportOut o{};
  o.cls = h_counts;
  export_write(t, o);
  return B2RF_OK;
}

int b2rf_trainer_set_trace(b2rf_trainer* t, int32_t on) {
  if (!t) return B2RF_ERR_ARG;
  t->keepTrace = on != 0;
  return B2RF_OK;
}

int64_t b2rf_trainer_trace_size(const b2rf_trainer* t) { return t ? (int64_t)t->trace.size() : 0; }
int b2rf_trainer_trace(const b2rf_trainer* t, int32_t* h_ints, double* h_dbls) {
  if (!t) return B2RF_ERR_ARG;
  for (size_t i = 0; i < t->trace.size(); i++) {
    const auto& r = t->trace[i];
    h_ints[i * 5] = r.g; h_ints[i * 5 + 1] = r.t; h_ints[i * 5 + 2] = r.id; h_ints[i * 5 + 3] = r.f; h_ints[i * 5 + 4] = r.s;
    h_dbls[i * 5] = r.gain; h_dbls[i * 5 + 1] = r.imp; h_dbls[i * 5 + 2] = r.cnt; h_dbls[i * 5 + 3] = r.sum; h_dbls[i * 5 + 4] = r.sq;
  }
  return B2RF_OK;
}

}  // extern "C"
