// b2rf device code (sm_100a) + kernel launchers.
//
// Kernels:  K1 find_splits (RandomForest.findSplitsForContinuousFeature), K2 bin (TreePoint.findBin),
// K3 bag (BaggedPoint.convertToBaggedRDD), K4 hist (DTStatsAggregator.update via binSeqOp) with the fused K6 row
// routing (LearningNode.predictImpl), K5 best_split (RandomForest.binsToBestSplit), K7 predict.
// Compiled with -fmad=false: MLlib's arithmetic is unfused JVM double arithmetic.
#include "b2rf_internal.h"
#include "../../include/b2rf.h"

#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>

namespace b2rf {

static thread_local std::string g_err;
void set_error(const std::string& m) { g_err = m; }
const char* last_error_cstr() { return g_err.c_str(); }
bool cuda_ok(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return true;
  g_err = std::string(what) + ": " + cudaGetErrorString(e);
  return false;
}
int device_sm_count() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) return 148;
    cached = p.multiProcessorCount; cached_dev = dev;
  }
  return cached;
}

// =================================================================================================
// small PTX helpers: mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP)
// =================================================================================================
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// Bin codes are uint8 when maxBins <= 256 and uint16 otherwise (codeBytes = 1 / 2, uniform per launch).
__device__ __forceinline__ uint32_t load_code(const void* base, size_t idx, int codeBytes) {
  return codeBytes == 2 ? (uint32_t) reinterpret_cast<const uint16_t*>(base)[idx] : (uint32_t) reinterpret_cast<const uint8_t*>(base)[idx];
}
__device__ __forceinline__ bool in_cat_mask(const uint32_t* masks, uint32_t aux, int maskWords, uint32_t bin) {
  return (masks[(size_t)(aux & 0x7fffffffu) * maskWords + (bin >> 5)] >> (bin & 31)) & 1u;
}

// =================================================================================================
// K4 (+K6): histogram build over TMA-staged column tiles of the bin matrix
// =================================================================================================
constexpr int HIST_THREADS = 512;
constexpr int HIST_RPT = 2;                       // rows per thread per tile
constexpr int HIST_R = HIST_THREADS * HIST_RPT;   // rows per tile (1024)
constexpr int HIST_PAD = 16;                      // smem pitch padding (keeps 16-byte alignment for bulk copies)

struct HistKArgs {
  HistLaunch a;
  int32_t R;         // rows per tile actually used (multiple of 128, <= HIST_R)
  int32_t pitch;     // smem bytes per feature column
  int32_t nbuf;      // 1 or 2 staging buffers
  int64_t numTiles;
};

// MODE 0: slot ids given per (listed tree, row), histogram by global RED   (stand-alone b2rf_hist_build)
// MODE 1: K6 routing from the cached node + histogram by global RED        (trainer, scan mode)
// MODE 2: K6 routing from the cached node + per-slot row COUNT only        (trainer, bucket mode, pass 1)
template <int MODE>
__global__ void __launch_bounds__(HIST_THREADS, 1) hist_kernel(const HistKArgs k) {
  constexpr bool TRAVERSE = MODE != 0;
  extern __shared__ __align__(128) unsigned char smem[];
  const HistLaunch& a = k.a;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);  // 2 barriers
  unsigned char* tiles = smem + 128;
  const size_t bufBytes = (size_t)a.F * k.pitch;
  const int tid = threadIdx.x;

  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  auto issue = [&](int64_t tile, int buf) {
    // warp 0: one bulk copy per feature column segment
    const int64_t r0 = tile * k.R;
    int64_t rem = a.n_rows - r0;
    uint32_t nb = (uint32_t)(rem < k.R ? ((rem + 15) & ~(int64_t)15) : k.R) * (uint32_t)a.codeBytes;   // bytes per column segment
    if (tid == 0) mbar_expect_tx(&bars[buf], nb * (uint32_t)a.F);
    __syncwarp();
    for (int f = tid; f < a.F; f += 32)
      bulk_g2s(tiles + buf * bufBytes + (size_t)f * k.pitch, a.bins + ((size_t)f * a.ldb + r0) * a.codeBytes, nb, &bars[buf]);
  };

  int64_t it = 0;
  const int64_t first = blockIdx.x;
  if (first < k.numTiles && tid < 32) issue(first, 0);
  for (int64_t tile = first; tile < k.numTiles; tile += gridDim.x, ++it) {
    const int buf = (k.nbuf == 2) ? (int)(it & 1) : 0;
    const int64_t nxt = tile + gridDim.x;
    if (k.nbuf == 2 && nxt < k.numTiles && tid < 32) issue(nxt, buf ^ 1);
    const uint32_t parity = (k.nbuf == 2) ? (uint32_t)((it >> 1) & 1) : (uint32_t)(it & 1);
    mbar_wait(&bars[buf], parity);
    const unsigned char* tb = tiles + buf * bufBytes;
    const int64_t r0 = tile * k.R;

    double yv[HIST_RPT];
#pragma unroll
    for (int q = 0; q < HIST_RPT; q++) {
      const int r = tid + q * HIST_THREADS;
      const int64_t row = r0 + r;
      yv[q] = (r < k.R && row < a.n_rows) ? a.y[row] : 0.0;
    }
    for (int g = 0; g < a.Tg; g++) {
      const int t = a.tree_ids[g];
#pragma unroll
      for (int q = 0; q < HIST_RPT; q++) {
        const int r = tid + q * HIST_THREADS;
        const int64_t row = r0 + r;
        int slot = -1;
        if (MODE == 2) {
          // convergent form (no early exits): the lanes of a warp aggregate their per-slot counts before touching memory
          if (r < k.R && row < a.n_rows) {
            int32_t* np = a.node + (size_t)t * a.ldn + row;
            int n = *np;
            if (n >= 0) {
              const DNode* tn = a.dnodes + (size_t)t * a.nodeStride;
              DNode d = tn[n];
              bool moved = false;
              while (d.feature >= 0) {
                const uint32_t bin = load_code(tb + (size_t)d.feature * k.pitch, r, a.codeBytes);
                bool left;
                if (d.aux & 0x80000000u) left = in_cat_mask(a.catmasks, d.aux, a.maskWords, bin);
                else left = bin <= d.aux;
                n = d.left + (left ? 0 : 1);
                d = tn[n];
                moved = true;
              }
              if (d.slot == SLOT_DEAD) *np = -1;
              else { if (moved) *np = n; slot = d.slot; }
            }
          }
          const unsigned peers = __match_any_sync(0xffffffffu, slot);
          if (slot >= 0 && (int)(__ffs(peers) - 1) == (tid & 31)) atomicAdd(&a.slot_count[slot], (uint32_t)__popc(peers));
          continue;
        }
        if (r >= k.R || row >= a.n_rows) continue;
        if (TRAVERSE) {
          int32_t* np = a.node + (size_t)t * a.ldn + row;
          int n = *np;
          if (n < 0) continue;
          const DNode* tn = a.dnodes + (size_t)t * a.nodeStride;
          DNode d = tn[n];
          bool moved = false;
          while (d.feature >= 0) {  // K6: LearningNode.predictImpl on binned features, resumed from the cached node
            const uint32_t bin = load_code(tb + (size_t)d.feature * k.pitch, r, a.codeBytes);
            bool left;
            if (d.aux & 0x80000000u) left = in_cat_mask(a.catmasks, d.aux, a.maskWords, bin);
            else left = bin <= d.aux;
            n = d.left + (left ? 0 : 1);
            d = tn[n];
            moved = true;
          }
          if (d.slot == SLOT_DEAD) { *np = -1; continue; }
          if (moved) *np = n;
          slot = d.slot;
          if (slot < 0) continue;
        } else {
          slot = a.node[(size_t)g * a.ldn + row];
          if (slot < 0) continue;
        }
        const uint32_t wv = a.w[(size_t)t * a.ldw + row];
        if (wv == 0) continue;
        const double dw = (double)wv;
        const double wy = dw * yv[q];
        const double wyy = wy * yv[q];
        const uint2* fo = a.slot_fo + (size_t)slot * a.Fsub;
        double* hb = a.hist + (size_t)a.slot_base[slot] * 4;
        for (int j = 0; j < a.Fsub; j++) {
          const uint2 e = fo[j];
          const uint32_t bin = load_code(tb + (size_t)e.x * k.pitch, r, a.codeBytes);
          double* c = hb + ((size_t)e.y + bin) * 4;
          atomicAdd(c, dw);
          atomicAdd(c + 1, wy);
          atomicAdd(c + 2, wyy);
        }
      }
    }
    __syncthreads();  // tile fully consumed before its buffer is refilled
    if (k.nbuf == 1 && nxt < k.numTiles && tid < 32) issue(nxt, 0);
  }
}

cudaError_t launch_hist(const HistLaunch& a, int mode, cudaStream_t s, std::string* err) {
  if (a.n_rows <= 0 || a.Tg <= 0) return cudaSuccess;
  HistKArgs k;
  k.a = a;
  // tile geometry: largest R (multiple of 128, <= 1024) whose staging fits; double-buffer when possible
  const size_t budget = 200 * 1024;
  int R = HIST_R, nbuf = 2;
  const int cb = a.codeBytes;
  while (R > 128 && (size_t)2 * a.F * (R * cb + HIST_PAD) > budget) R -= 128;
  if ((size_t)2 * a.F * (R * cb + HIST_PAD) > budget) { nbuf = 1; R = HIST_R; while (R > 128 && (size_t)a.F * (R * cb + HIST_PAD) > budget) R -= 128; }
  if ((size_t)nbuf * a.F * (R * cb + HIST_PAD) > budget) {
    if (err) *err = "hist: too many features for one shared-memory tile (F > 1400 unsupported)";
    return cudaErrorInvalidValue;
  }
  k.R = R; k.pitch = R * cb + HIST_PAD; k.nbuf = nbuf;
  k.numTiles = (a.n_rows + R - 1) / R;
  const size_t smem = 128 + (size_t)nbuf * a.F * k.pitch;
  auto kern = mode == 0 ? hist_kernel<0> : (mode == 1 ? hist_kernel<1> : hist_kernel<2>);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int64_t grid = device_sm_count();
  if (grid > k.numTiles) grid = k.numTiles;
  kern<<<(unsigned)grid, HIST_THREADS, smem, s>>>(k);
  return cudaGetLastError();
}


// =================================================================================================
// Bucket mode (K4'): rows of each node slot are gathered into a per-group row list by a counting sort
// (count in hist_kernel<2>, exclusive scan, scatter), then one CTA per (slot, row chunk) accumulates that
// slot's histogram in SHARED memory (u32 count by native ATOMS, f64 sums by CAS) with lane == feature, so the
// lanes of a warp never hit the same cell; the CTA flushes its non-empty cells with global RED.
// =================================================================================================
__global__ void __launch_bounds__(1024) scan_slots_kernel(const uint32_t* cnt, int nSlots, int CH, uint64_t* start, uint32_t* chunkPrefix,
                                                           uint32_t* cursor) {
  __shared__ unsigned long long s_rows[1024];
  __shared__ uint32_t s_chunks[1024];
  const int tid = threadIdx.x;
  const int per = (nSlots + 1023) / 1024;
  const int b = tid * per, e = min(nSlots, b + per);
  unsigned long long r = 0; uint32_t c = 0;
  for (int i = b; i < e; i++) { r += cnt[i]; c += (cnt[i] + CH - 1) / CH; }
  s_rows[tid] = r; s_chunks[tid] = c;
  __syncthreads();
  if (tid == 0) {
    unsigned long long ar = 0; uint32_t ac = 0;
    for (int i = 0; i < 1024; i++) { const unsigned long long tr = s_rows[i]; const uint32_t tc = s_chunks[i]; s_rows[i] = ar; s_chunks[i] = ac; ar += tr; ac += tc; }
    start[nSlots] = ar; chunkPrefix[nSlots] = ac;
  }
  __syncthreads();
  r = s_rows[tid]; c = s_chunks[tid];
  for (int i = b; i < e; i++) { start[i] = r; chunkPrefix[i] = c; cursor[i] = 0; r += cnt[i]; c += (cnt[i] + CH - 1) / CH; }
}

__global__ void __launch_bounds__(256) scatter_rows_kernel(const int32_t* node, int64_t ldn, const uint8_t* w, int64_t ldw, const int32_t* tree_ids,
                                                            int Tg, int64_t n_rows, const DNode* dnodes, int64_t nodeStride, const uint64_t* start,
                                                            uint32_t* cursor, uint32_t* rowlist) {
  const int g = blockIdx.y;
  const int t = tree_ids[g];
  const DNode* tn = dnodes + (size_t)t * nodeStride;
  const int lane = threadIdx.x & 31;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t n_pad = (n_rows + 31) & ~(int64_t)31;   // whole warps stay in the loop together
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < n_pad; row += stride) {
    int slot = -1;
    if (row < n_rows) {
      const int n = node[(size_t)t * ldn + row];
      if (n >= 0) slot = tn[n].slot;
    }
    // one cursor bump per (warp, slot): lanes that share a slot take consecutive places
    const unsigned peers = __match_any_sync(0xffffffffu, slot);
    const int leader = __ffs(peers) - 1;
    uint32_t base = 0;
    if (slot >= 0 && lane == leader) base = atomicAdd(&cursor[slot], (uint32_t)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (slot >= 0) {
      uint32_t wv = w[(size_t)t * ldw + row];
      if (wv > 31u) wv = 31u;   // packed in 5 bits; a Poisson(<=1) draw above 31 has probability < 1e-34
      const uint32_t pos = base + (uint32_t)__popc(peers & ((1u << lane) - 1u));
      rowlist[start[slot] + pos] = (uint32_t)row | (wv << 27);
    }
  }
}

constexpr int BUCKET_THREADS = 256;   // default; tasks whose histogram leaves room for only one or two CTAs per SM run 512 threads
// ---- dedicated instantiation for the common case: uint8 codes, at most 32 features per node, lane == feature ----
// (kept as its own small kernel: the general kernel below handles feature groups, uint16 codes and 512-thread CTAs, and
// folding this case into it cost 5-10 % of K4 time in same-box A/B runs: 0.657 s vs 0.69-0.73 s per C3 fit)
// Rows [0, n) of one (slot, chunk) task, one warp row-block at a time.  Lane k of a warp owns row k of the current 32-row
// block: it fetches the packed entry and the label once and forms (w, w*y, w*y*y); the values are then broadcast row by row
// while every lane (== feature) gathers its own bin code and updates its own histogram column.
template <bool DIRECT>
__device__ __forceinline__ void bucket_rows32(const BucketLaunch& a, const uint32_t* rl, int n, int warp, int lane, bool active, uint32_t featCol,
                                            double* hb, uint32_t gbase, double* ssum, double* ssq, unsigned* scnt, uint32_t sbase, uint32_t sstep) {
  constexpr int W = BUCKET_THREADS / 32;
  const unsigned FULL = 0xffffffffu;
  for (int base = warp * 32; base < n; base += W * 32) {
    const uint32_t mine = (base + lane < n) ? rl[base + lane] : 0u;          // w == 0 marks padding
    const uint32_t myw = mine >> 27;
    const double myy = myw ? a.y[mine & 0x7ffffffu] : 0.0;
    const double mywy = (double)myw * myy;                                     // MLlib: instanceWeight * label * label == (w*y)*y
    const double mywyy = mywy * myy;
    const int nb = min(32, n - base);
    for (int q = 0; q < nb; q += 4) {
      uint32_t bin[4], wv[4]; double wy[4], wyy[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const uint32_t p = __shfl_sync(FULL, mine, (q + u) & 31);
        wy[u] = __shfl_sync(FULL, mywy, (q + u) & 31);
        wyy[u] = __shfl_sync(FULL, mywyy, (q + u) & 31);
        wv[u] = (q + u) < nb ? (p >> 27) : 0u;
        bin[u] = (wv[u] && active) ? a.bins_rm[(size_t)(p & 0x7ffffffu) * a.ldr + featCol] : 0u;
      }
#pragma unroll
      for (int u = 0; u < 4; u++) {
        if (wv[u] != 0u && active) {
          if (DIRECT) {
            double* c = hb + (size_t)(gbase + bin[u]) * 4;
            atomicAdd(c, (double)wv[u]); atomicAdd(c + 1, wy[u]); atomicAdd(c + 2, wyy[u]);
          } else {
            const uint32_t cell = sbase + bin[u] * sstep;
            atomicAdd(&scnt[cell], wv[u]); atomicAdd(&ssum[cell], wy[u]); atomicAdd(&ssq[cell], wyy[u]);
          }
        }
      }
    }
  }
}

__global__ void __launch_bounds__(BUCKET_THREADS) bucket_hist32_kernel(const BucketLaunch a) {
  extern __shared__ __align__(16) unsigned char smraw[];
  double* ssum = reinterpret_cast<double*>(smraw);
  double* ssq = ssum + a.maxCells;
  unsigned* scnt = reinterpret_cast<unsigned*>(ssq + a.maxCells);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t total = a.chunk_prefix[a.nSlots];
  for (uint32_t task = blockIdx.x; task < total; task += gridDim.x) {
    int lo = 0, hi = a.nSlots;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (a.chunk_prefix[mid] <= task) lo = mid; else hi = mid; }
    const int slot = lo;
    const uint32_t chunk = task - a.chunk_prefix[slot];
    uint32_t cnt; const uint32_t* rl;
    const uint32_t begin = chunk * (uint32_t)a.CH;
    if (a.dseg != nullptr) {
      const DSeg sg = a.dseg[a.slot_pos[slot]];
      cnt = sg.count;
      rl = (sg.buf ? a.idx1 : a.idx0) + sg.start + begin;
    } else {
      cnt = a.slot_count[slot];
      rl = a.rowlist + a.slot_start[slot] + begin;
    }
    const int n = (int)min((uint32_t)a.CH, cnt - begin);
    const int ncell = a.slot_cells[slot];
    const uint2* fo = a.slot_fo + (size_t)slot * a.Fsub;
    double* hb = a.hist + (size_t)a.slot_base[slot] * 4;
    // Shared layout.  padded (padB > 0): cell (feature j, bin b) sits at ((j/32)*padB + b)*32 + j%32, i.e. bank == lane: the
    // lanes of a warp (lane == feature) can never conflict, whatever their bins.  compact: cell = feature offset + bin.
    const int padB = a.padB;
    const int scells = padB > 0 ? ((a.Fsub + 31) / 32) * padB * 32 : ncell;
    const bool direct = (int64_t)n * a.Fsub < (int64_t)ncell || scells > a.maxCells;   // few rows: skip the shared-memory round trip
    if (!direct) {
      for (int c = tid; c < scells; c += BUCKET_THREADS) { ssum[c] = 0.0; ssq[c] = 0.0; scnt[c] = 0u; }
      __syncthreads();
    }
    for (int j0 = 0; j0 < a.Fsub; j0 += 32) {
      const int j = j0 + lane;
      const bool active = j < a.Fsub;
      const uint2 e = active ? fo[j] : make_uint2(0u, 0u);
      const uint32_t sbase = padB > 0 ? (uint32_t)((j0 >> 5) * padB * 32 + lane) : e.y;   // shared-memory cell of bin 0
      const uint32_t sstep = padB > 0 ? 32u : 1u;
      if (direct) bucket_rows32<true>(a, rl, n, warp, lane, active, e.x, hb, e.y, ssum, ssq, scnt, sbase, sstep);
      else bucket_rows32<false>(a, rl, n, warp, lane, active, e.x, hb, e.y, ssum, ssq, scnt, sbase, sstep);
    }
    if (!direct) {
      __syncthreads();
      for (int c = tid; c < scells; c += BUCKET_THREADS) {
        const unsigned k = scnt[c];
        if (!k) continue;
        size_t gc = (size_t)c;
        if (padB > 0) { const int jj = (c / (padB * 32)) * 32 + (c & 31); gc = (size_t)fo[jj].y + (size_t)((c >> 5) % padB); }   // untouched cells of idle lanes stay 0
        double* g = hb + gc * 4;
        atomicAdd(g, (double)k); atomicAdd(g + 1, ssum[c]); atomicAdd(g + 2, ssq[c]);
      }
      __syncthreads();
    }
  }
}

// Rows [0, n) of one (slot, chunk) task, one warp row-block at a time.  Lane k of a warp owns row k of the current 32-row
// block: it fetches the packed entry and the label once and forms (w, w*y, w*y*y); the values are then broadcast row by row
// while every lane (row slot r = lane / G, feature fl = lane % G) gathers its own bin code and updates its own cell.
// ONE_ROW (G == 32) and the code width CB are compile-time in the hot instantiations; CB == 0 reads a.codeBytes.
template <bool DIRECT, int THREADS, bool ONE_ROW, int CB>
__device__ __forceinline__ void bucket_rows_impl(const BucketLaunch& a, const uint32_t* rl, int n, int warp, int lane, bool active, uint32_t featCol,
                                                 double* hb, uint32_t gbase, double* ssum, double* ssq, unsigned* scnt, uint32_t sbase, uint32_t sstep,
                                                 int Rrt, int rrt) {
  const int R = ONE_ROW ? 1 : Rrt, r = ONE_ROW ? 0 : rrt;
  constexpr int W = THREADS / 32;
  const unsigned FULL = 0xffffffffu;
  for (int base = warp * 32; base < n; base += W * 32) {
    const uint32_t mine = (base + lane < n) ? rl[base + lane] : 0u;          // w == 0 marks padding
    const uint32_t myw = mine >> 27;
    const double myy = myw ? a.y[mine & 0x7ffffffu] : 0.0;
    const double mywy = (double)myw * myy;                                     // MLlib: instanceWeight * label * label == (w*y)*y
    const double mywyy = mywy * myy;
    const int nb = min(32, n - base);
    for (int q = 0; q < nb; q += 4 * R) {
      uint32_t bin[4], wv[4]; double wy[4], wyy[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int ri = q + u * R + r;
        const uint32_t p = __shfl_sync(FULL, mine, ri & 31);
        wy[u] = __shfl_sync(FULL, mywy, ri & 31);
        wyy[u] = __shfl_sync(FULL, mywyy, ri & 31);
        wv[u] = ri < nb ? (p >> 27) : 0u;
        bin[u] = (wv[u] && active) ? load_code(a.bins_rm, (size_t)(p & 0x7ffffffu) * a.ldr + featCol, CB ? CB : a.codeBytes) : 0u;
      }
#pragma unroll
      for (int u = 0; u < 4; u++) {
        if (wv[u] != 0u && active) {
          if (DIRECT) {
            double* c = hb + (size_t)(gbase + bin[u]) * 4;
            atomicAdd(c, (double)wv[u]); atomicAdd(c + 1, wy[u]); atomicAdd(c + 2, wyy[u]);
          } else {
            const uint32_t cell = sbase + bin[u] * sstep;
            atomicAdd(&scnt[cell], wv[u]); atomicAdd(&ssum[cell], wy[u]); atomicAdd(&ssq[cell], wyy[u]);
          }
        }
      }
    }
  }
}
// One instantiation per (block size, rows per warp instruction, code width, shared layout): each kernel carries exactly two
// copies of the row loop (shared-memory path, direct path), which keeps the hot loop's instruction footprint small.
template <int THREADS, bool ONE_ROW, int CB, bool PADDED>
__global__ void __launch_bounds__(THREADS) bucket_hist_kernel(const BucketLaunch a) {
  extern __shared__ __align__(16) unsigned char smraw[];
  double* ssum = reinterpret_cast<double*>(smraw);
  double* ssq = ssum + a.maxCells;
  unsigned* scnt = reinterpret_cast<unsigned*>(ssq + a.maxCells);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // Shared layout.  padded (padB > 0): the slot's features are cut into groups of G (a power of two <= 32); one task handles one
  // group and keeps cell (feature fl of the group, bin b) at b*G + fl.  A warp instruction then covers 32/G rows x G features
  // with bank == lane for G == 32 (no conflict possible) and at most 32/G-way conflicts otherwise, and the task's histogram is
  // padB*G cells whatever the feature count, so several CTAs fit per SM even for 256 bins.  compact (padB == 0): one task per
  // (slot, chunk) with cell = feature offset + bin.
  const int padB = PADDED ? a.padB : 0;
  const int G = PADDED ? (ONE_ROW ? 32 : a.G) : 32;
  const int nG = PADDED ? (a.Fsub + G - 1) / G : 1;
  const uint32_t total = a.chunk_prefix[a.nSlots] * (uint32_t)nG;
  for (uint32_t task = blockIdx.x; task < total; task += gridDim.x) {
    const uint32_t bt = task / (uint32_t)nG;
    const int g = (int)(task % (uint32_t)nG);
    int lo = 0, hi = a.nSlots;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (a.chunk_prefix[mid] <= bt) lo = mid; else hi = mid; }
    const int slot = lo;
    const uint32_t chunk = bt - a.chunk_prefix[slot];
    uint32_t cnt; const uint32_t* rl;
    const uint32_t begin = chunk * (uint32_t)a.CH;
    if (a.dseg != nullptr) {
      const DSeg sg = a.dseg[a.slot_pos[slot]];
      cnt = sg.count;
      rl = (sg.buf ? a.idx1 : a.idx0) + sg.start + begin;
    } else {
      cnt = a.slot_count[slot];
      rl = a.rowlist + a.slot_start[slot] + begin;
    }
    const int n = (int)min((uint32_t)a.CH, cnt - begin);
    const int ncell = a.slot_cells[slot];
    const uint2* fo = a.slot_fo + (size_t)slot * a.Fsub;
    double* hb = a.hist + (size_t)a.slot_base[slot] * 4;
    const int scells = PADDED ? padB * G : ncell;
    const bool direct = (int64_t)n * a.Fsub < (int64_t)ncell || scells > a.maxCells;   // few rows: skip the shared-memory round trip
    if (!direct) {
      for (int c = tid; c < scells; c += THREADS) { ssum[c] = 0.0; ssq[c] = 0.0; scnt[c] = 0u; }
      __syncthreads();
    }
    if (PADDED) {
      const int fl = lane & (G - 1), r = lane / G;
      const int j = g * G + fl;
      const bool active = j < a.Fsub;
      const uint2 e = active ? fo[j] : make_uint2(0u, 0u);
      if (direct) bucket_rows_impl<true, THREADS, ONE_ROW, CB>(a, rl, n, warp, lane, active, e.x, hb, e.y, ssum, ssq, scnt, 0u, 0u, 32 / G, r);
      else bucket_rows_impl<false, THREADS, ONE_ROW, CB>(a, rl, n, warp, lane, active, e.x, hb, e.y, ssum, ssq, scnt, (uint32_t)fl, (uint32_t)G, 32 / G, r);
    } else {
      for (int j0 = 0; j0 < a.Fsub; j0 += 32) {
        const int j = j0 + lane;
        const bool active = j < a.Fsub;
        const uint2 e = active ? fo[j] : make_uint2(0u, 0u);
        if (direct) bucket_rows_impl<true, THREADS, true, CB>(a, rl, n, warp, lane, active, e.x, hb, e.y, ssum, ssq, scnt, 0u, 0u, 1, 0);
        else bucket_rows_impl<false, THREADS, true, CB>(a, rl, n, warp, lane, active, e.x, hb, e.y, ssum, ssq, scnt, e.y, 1u, 1, 0);
      }
    }
    if (!direct) {
      __syncthreads();
      for (int c = tid; c < scells; c += THREADS) {
        const unsigned k = scnt[c];
        if (!k) continue;
        size_t gc = (size_t)c;
        if (PADDED) gc = (size_t)fo[g * G + (c & (G - 1))].y + (size_t)(c / G);   // cells of idle lanes are never touched, so k == 0 there
        double* gp = hb + gc * 4;
        atomicAdd(gp, (double)k); atomicAdd(gp + 1, ssum[c]); atomicAdd(gp + 2, ssq[c]);
      }
      __syncthreads();
    }
  }
}

// picks the instantiation whose hot loop matches the launch (block size, one row per instruction or several, code width, layout)
static cudaError_t launch_bucket_kernel(const BucketLaunch& b, int grid, int threads, size_t smem, cudaStream_t s) {
  if (b.padB > 0 && b.G == 32 && b.Fsub <= 32 && b.codeBytes == 1 && threads == BUCKET_THREADS) {
    cudaError_t e = cudaFuncSetAttribute(bucket_hist32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    bucket_hist32_kernel<<<grid, BUCKET_THREADS, smem, s>>>(b);
    return cudaGetLastError();
  }
  const bool padded = b.padB > 0;
  const bool one = padded && b.G == 32;
  const int cb = b.codeBytes;
#define B2RF_LAUNCH_BUCKET(T, O, C, P)                                                                                 \
  do {                                                                                                                 \
    cudaError_t e_ = cudaFuncSetAttribute(bucket_hist_kernel<T, O, C, P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e_ != cudaSuccess) return e_;                                                                                  \
    bucket_hist_kernel<T, O, C, P><<<grid, T, smem, s>>>(b);                                                           \
    return cudaGetLastError();                                                                                         \
  } while (0)
#define B2RF_PICK(T)                                                          \
  do {                                                                        \
    if (!padded) { if (cb == 1) B2RF_LAUNCH_BUCKET(T, true, 1, false); B2RF_LAUNCH_BUCKET(T, true, 2, false); } \
    if (one && cb == 1) B2RF_LAUNCH_BUCKET(T, true, 1, true);                 \
    if (one) B2RF_LAUNCH_BUCKET(T, true, 2, true);                            \
    if (cb == 1) B2RF_LAUNCH_BUCKET(T, false, 1, true);                       \
    B2RF_LAUNCH_BUCKET(T, false, 2, true);                                    \
  } while (0)
  if (threads == 512) B2RF_PICK(512);
  B2RF_PICK(256);
#undef B2RF_PICK
#undef B2RF_LAUNCH_BUCKET
}

// ---- segment mode: task table from the nodes' row segments, then the same shared-memory histogram kernel ----
__global__ void __launch_bounds__(1024) seg_tasks_kernel(const DSeg* dseg, const int64_t* pos, int n, int CH, uint32_t* chunkPrefix) {
  __shared__ uint32_t s_chunks[1024];
  const int tid = threadIdx.x;
  const int per = (n + 1023) / 1024;
  const int b = tid * per, e = min(n, b + per);
  uint32_t c = 0;
  for (int i = b; i < e; i++) c += (dseg[pos[i]].count + CH - 1) / CH;
  s_chunks[tid] = c;
  __syncthreads();
  if (tid == 0) {
    uint32_t ac = 0;
    for (int i = 0; i < 1024; i++) { const uint32_t tc = s_chunks[i]; s_chunks[i] = ac; ac += tc; }
    chunkPrefix[n] = ac;
  }
  __syncthreads();
  c = s_chunks[tid];
  for (int i = b; i < e; i++) { chunkPrefix[i] = c; c += (dseg[pos[i]].count + CH - 1) / CH; }
}

cudaError_t launch_segment_hist(const BucketLaunch& b, cudaStream_t s) {
  seg_tasks_kernel<<<1, 1024, 0, s>>>(b.dseg, b.slot_pos, b.nSlots, b.CH, const_cast<uint32_t*>(b.chunk_prefix));
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const size_t smem = (size_t)b.maxCells * 20 + 16;
  int perSM = (int)((size_t)200 * 1024 / (smem + 1024));
  if (perSM < 1) perSM = 1;
  if (perSM > 4) perSM = 4;                                   // 64 registers x 256 threads: four CTAs fill the register file
  return launch_bucket_kernel(b, device_sm_count() * perSM, perSM <= 2 ? 512 : BUCKET_THREADS, smem, s);   // big histograms: fewer CTAs, 16 warps each
}

// ---- initial row lists: per tree, the in-bag rows in ascending order, packed row | weight << 27 ----
__global__ void __launch_bounds__(256) compact_count_kernel(const uint8_t* w, int64_t ldw, int64_t n_rows, uint32_t* blockCnt, int nb) {
  const int t = blockIdx.y, b = blockIdx.x;
  const int64_t r0 = (int64_t)b * COMPACT_ROWS;
  uint32_t c = 0;
  for (int i = threadIdx.x; i < COMPACT_ROWS; i += 256) { const int64_t r = r0 + i; c += (r < n_rows && w[(size_t)t * ldw + r] != 0); }
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  __shared__ uint32_t sw[8];
  if ((threadIdx.x & 31) == 0) sw[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) { uint32_t s = 0; for (int i = 0; i < 8; i++) s += sw[i]; blockCnt[(size_t)t * nb + b] = s; }
}
__global__ void compact_scan_kernel(uint32_t* blockCnt, int nb, unsigned long long* totals) {
  const int t = blockIdx.x;
  if (threadIdx.x != 0) return;
  unsigned long long run = 0;
  for (int b = 0; b < nb; b++) { const uint32_t c = blockCnt[(size_t)t * nb + b]; blockCnt[(size_t)t * nb + b] = (uint32_t)run; run += c; }
  totals[t] = run;
}
__global__ void __launch_bounds__(256) compact_write_kernel(const uint8_t* w, int64_t ldw, int64_t n_rows, const uint32_t* blockOff, int nb,
                                                             const uint64_t* treeBase, uint32_t* idx) {
  const int t = blockIdx.y, b = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ uint32_t sw[8];
  uint32_t run = 0;
  uint32_t* out = idx + treeBase[t] + blockOff[(size_t)t * nb + b];
  for (int it = 0; it < COMPACT_ROWS / 256; it++) {
    const int64_t r = (int64_t)b * COMPACT_ROWS + it * 256 + threadIdx.x;
    uint32_t wv = (r < n_rows) ? w[(size_t)t * ldw + r] : 0u;
    const unsigned bal = __ballot_sync(0xffffffffu, wv != 0);
    if (lane == 0) sw[warp] = __popc(bal);
    __syncthreads();
    uint32_t before = 0, total = 0;
    for (int i = 0; i < 8; i++) { if (i < warp) before += sw[i]; total += sw[i]; }
    if (wv != 0) {
      if (wv > 31u) wv = 31u;   // packed in 5 bits
      out[run + before + __popc(bal & ((1u << lane) - 1u))] = (uint32_t)r | (wv << 27);
    }
    run += total;
    __syncthreads();
  }
}
cudaError_t launch_compact_count(const uint8_t* w, int64_t ldw, int T, int64_t n_rows, uint32_t* blockCnt, unsigned long long* totals, cudaStream_t s) {
  const int nb = (int)((n_rows + COMPACT_ROWS - 1) / COMPACT_ROWS);
  if (nb <= 0 || T <= 0) return cudaSuccess;
  compact_count_kernel<<<dim3((unsigned)nb, (unsigned)T), 256, 0, s>>>(w, ldw, n_rows, blockCnt, nb);
  compact_scan_kernel<<<T, 32, 0, s>>>(blockCnt, nb, totals);
  return cudaGetLastError();
}
cudaError_t launch_compact_write(const uint8_t* w, int64_t ldw, int T, int64_t n_rows, const uint32_t* blockOff, const uint64_t* treeBase, uint32_t* idx,
                                 cudaStream_t s) {
  const int nb = (int)((n_rows + COMPACT_ROWS - 1) / COMPACT_ROWS);
  if (nb <= 0 || T <= 0) return cudaSuccess;
  compact_write_kernel<<<dim3((unsigned)nb, (unsigned)T), 256, 0, s>>>(w, ldw, n_rows, blockOff, nb, treeBase, idx);
  return cudaGetLastError();
}
__global__ void regrow_segs_kernel(const DSeg* src, int64_t oldStride, DSeg* dst, int64_t newStride) {
  const int t = blockIdx.y;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < oldStride; i += (int64_t)gridDim.x * blockDim.x)
    dst[(size_t)t * newStride + i] = src[(size_t)t * oldStride + i];
}
cudaError_t launch_regrow_segs(const DSeg* src, int64_t oldStride, DSeg* dst, int64_t newStride, int T, cudaStream_t s) {
  regrow_segs_kernel<<<dim3(64, (unsigned)T), 256, 0, s>>>(src, oldStride, dst, newStride);
  return cudaGetLastError();
}

// ---- K6 in segment mode: stable partition of each split node's row segment into its children's segments ----
__device__ __forceinline__ int find_task_owner(const uint32_t* prefix, int n, uint32_t task) {
  int lo = 0, hi = n;
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (prefix[mid] <= task) lo = mid; else hi = mid; }
  return lo;
}

__global__ void __launch_bounds__(256) part_count_kernel(const PartLaunch p) {
  __shared__ uint32_t sw[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t total = p.chunk_prefix[p.nSplits];
  for (uint32_t task = blockIdx.x; task < total; task += gridDim.x) {
    const int k = find_task_owner(p.chunk_prefix, p.nSplits, task);
    const int64_t pos = p.split_pos[k];
    const DSeg sg = p.dseg[pos];
    const DNode d = p.dnodes[pos];
    const uint32_t c = task - p.chunk_prefix[k];
    const uint32_t* src = (sg.buf ? p.idx1 : p.idx0) + sg.start + (size_t)c * SEG_CH;
    const int n = (int)min((uint32_t)SEG_CH, sg.count - c * SEG_CH);
    const uint8_t* col = p.bins + (size_t)d.feature * p.ldb * p.codeBytes;
    const bool cat = (d.aux & 0x80000000u) != 0;
    const uint32_t* mask = cat ? p.catmasks + (size_t)(d.aux & 0x7fffffffu) * p.maskWords : nullptr;
    uint32_t cnt = 0;
    for (int i = tid; i < SEG_CH; i += 256) {
      bool left = false;
      if (i < n) {
        const uint32_t bin = load_code(col, src[i] & 0x7ffffffu, p.codeBytes);
        left = cat ? ((mask[bin >> 5] >> (bin & 31)) & 1u) : (bin <= d.aux);
      }
      const unsigned bal = __ballot_sync(0xffffffffu, left);
      if (lane == 0) { p.predbits[(size_t)task * (SEG_CH / 32) + (i >> 5)] = bal; cnt += __popc(bal); }
    }
    if (lane == 0) sw[warp] = cnt;
    __syncthreads();
    if (tid == 0) { uint32_t s = 0; for (int i = 0; i < 8; i++) s += sw[i]; p.left_cnt[task] = s; }
    __syncthreads();
  }
}

__global__ void part_scan_kernel(const PartLaunch p) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= p.nSplits) return;
  const int64_t pos = p.split_pos[k];
  const DSeg sg = p.dseg[pos];
  uint32_t run = 0;
  for (uint32_t task = p.chunk_prefix[k]; task < p.chunk_prefix[k + 1]; task++) { p.left_off[task] = run; run += p.left_cnt[task]; }
  DSeg L, R;
  L.start = sg.start; L.count = run; L.buf = sg.buf ^ 1u;
  R.start = sg.start + run; R.count = sg.count - run; R.buf = sg.buf ^ 1u;
  // the left child's flat position is stored right behind the parents in split_pos (see launch_partition); right = left + 1
  const int64_t lpos = p.split_pos[p.nSplits + k];
  p.dseg[lpos] = L;
  p.dseg[lpos + 1] = R;
}

__global__ void __launch_bounds__(256) part_scatter_kernel(const PartLaunch p) {
  __shared__ uint32_t sbits[SEG_CH / 32];
  __shared__ uint32_t spref[SEG_CH / 32];
  __shared__ uint32_t swarp[4];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t total = p.chunk_prefix[p.nSplits];
  for (uint32_t task = blockIdx.x; task < total; task += gridDim.x) {
    const int k = find_task_owner(p.chunk_prefix, p.nSplits, task);
    const int64_t pos = p.split_pos[k];
    const DSeg sg = p.dseg[pos];
    const uint32_t c = task - p.chunk_prefix[k];
    const uint32_t* src = (sg.buf ? p.idx1 : p.idx0) + sg.start + (size_t)c * SEG_CH;
    uint32_t* dst = (sg.buf ? p.idx0 : p.idx1) + sg.start;
    const int n = (int)min((uint32_t)SEG_CH, sg.count - c * SEG_CH);
    const uint32_t nLeft = p.dseg[p.split_pos[p.nSplits + k]].count;
    const uint32_t lbase = p.left_off[task];
    const uint32_t rbase = nLeft + (c * SEG_CH - lbase);
    if (tid < SEG_CH / 32) {   // 128 words: exclusive prefix of popcounts (4 warps)
      const uint32_t b = p.predbits[(size_t)task * (SEG_CH / 32) + tid];
      sbits[tid] = b;
      uint32_t v = __popc(b), incl = v;
      for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += u; }
      spref[tid] = incl - v;
      if (lane == 31) swarp[warp] = incl;
    }
    __syncthreads();
    if (tid < SEG_CH / 32) { uint32_t add = 0; for (int i = 0; i < warp; i++) add += swarp[i]; spref[tid] += add; }
    __syncthreads();
    for (int i = tid; i < n; i += 256) {
      const uint32_t bits = sbits[i >> 5];
      const uint32_t lrank = spref[i >> 5] + __popc(bits & ((1u << (i & 31)) - 1u));
      const uint32_t e = src[i];
      if ((bits >> (i & 31)) & 1u) dst[lbase + lrank] = e;
      else dst[rbase + ((uint32_t)i - lrank)] = e;
    }
    __syncthreads();
  }
}

// p.split_pos holds 2 * nSplits entries: parents' flat positions, then the left children's flat positions.
cudaError_t launch_partition(const PartLaunch& p, cudaStream_t s) {
  if (p.nSplits <= 0) return cudaSuccess;
  seg_tasks_kernel<<<1, 1024, 0, s>>>(p.dseg, p.split_pos, p.nSplits, SEG_CH, p.chunk_prefix);
  const int grid = device_sm_count() * 8;
  part_count_kernel<<<grid, 256, 0, s>>>(p);
  part_scan_kernel<<<(p.nSplits + 127) / 128, 128, 0, s>>>(p);
  part_scatter_kernel<<<grid, 256, 0, s>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_bucket_pipeline(const HistLaunch& h, const BucketLaunch& b, uint32_t* cursor, cudaStream_t s, std::string* err) {
  // pass 1: route rows (K6) and count rows per slot
  cudaError_t e = launch_hist(h, 2, s, err);
  if (e != cudaSuccess) return e;
  scan_slots_kernel<<<1, 1024, 0, s>>>(b.slot_count, b.nSlots, b.CH, const_cast<uint64_t*>(b.slot_start), const_cast<uint32_t*>(b.chunk_prefix), cursor);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  // pass 2: scatter packed (row, weight) entries into per-slot segments
  {
    int64_t bx = (h.n_rows + 255) / 256;
    const int64_t cap = (int64_t)device_sm_count() * 16;
    if (bx > cap) bx = cap;
    dim3 grid((unsigned)bx, (unsigned)h.Tg);
    scatter_rows_kernel<<<grid, 256, 0, s>>>(h.node, h.ldn, h.w, h.ldw, h.tree_ids, h.Tg, h.n_rows, h.dnodes, h.nodeStride, b.slot_start, cursor,
                                             const_cast<uint32_t*>(b.rowlist));
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
  }
  // pass 3: per-(slot, chunk) shared-memory histograms
  const size_t smem = (size_t)b.maxCells * 20 + 16;
  int perSM = (int)((size_t)200 * 1024 / (smem + 1024));
  if (perSM < 1) perSM = 1;
  if (perSM > 4) perSM = 4;                                   // 64 registers x 256 threads: four CTAs fill the register file
  return launch_bucket_kernel(b, device_sm_count() * perSM, perSM <= 2 ? 512 : BUCKET_THREADS, smem, s);   // big histograms: fewer CTAs, 16 warps each
}

__global__ void __launch_bounds__(256) count_active_kernel(const int32_t* node, int64_t ldn, int64_t n_rows, unsigned long long* out) {
  const int t = blockIdx.y;
  unsigned long long c = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_rows; i += (int64_t)gridDim.x * blockDim.x) c += node[(size_t)t * ldn + i] >= 0;
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&out[t], c);
}
cudaError_t launch_count_active(const int32_t* node, int64_t ldn, int T, int64_t n_rows, unsigned long long* out, cudaStream_t s) {
  if (n_rows <= 0 || T <= 0) return cudaSuccess;
  int64_t bx = (n_rows + 255) / 256;
  if (bx > 1024) bx = 1024;
  count_active_kernel<<<dim3((unsigned)bx, (unsigned)T), 256, 0, s>>>(node, ldn, n_rows, out);
  return cudaGetLastError();
}

__global__ void scatter_dnodes_kernel(DNode* dn, const DNodeUpdate* u, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dn[u[i].pos] = u[i].node;
}
cudaError_t launch_scatter_dnodes(DNode* dnodes, const DNodeUpdate* upd, int64_t n, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  scatter_dnodes_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(dnodes, upd, n);
  return cudaGetLastError();
}

// =================================================================================================
// K5: binsToBestSplit — one warp per node slot
// =================================================================================================
constexpr int SPLIT_WARPS = 4;   // at most; fewer when a warp's shared arrays (58 B per bin) would not fit
// per-warp shared arrays, carved from dynamic shared memory: raw[maxB][3], cum[maxB][3], key[maxB] (f64), mask[maskWords] (u32),
// order[maxB] (u16)
__host__ __device__ inline size_t split_warp_bytes(int maxB, int maskWords) {
  size_t b = (size_t)maxB * (3 + 3 + 1) * 8 + (size_t)maskWords * 4 + (size_t)maxB * 2;
  return (b + 15) & ~(size_t)15;
}

__device__ __forceinline__ double variance_calc(double cnt, double sum, double sq) {  // Variance.calculate
  if (cnt == 0) return 0.0;
  const double sl = sq - (sum * sum) / cnt;
  return sl / cnt;
}

struct SplitView {   // one warp's shared arrays
  double (*raw)[3];
  double (*cum)[3];
  double* key;
  uint32_t* mask;
  uint16_t* order;
};

__global__ void __launch_bounds__(SPLIT_WARPS * 32) best_split_kernel(const SplitLaunch a) {
  extern __shared__ __align__(16) unsigned char smraw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int warpsPerBlock = blockDim.x >> 5;
  const int slot = blockIdx.x * warpsPerBlock + warp;
  if (slot >= a.nSlots) return;
  SplitView sm;
  {
    unsigned char* base = smraw + (size_t)warp * split_warp_bytes(a.maxB, a.maskWords);
    sm.raw = reinterpret_cast<double (*)[3]>(base);
    sm.cum = reinterpret_cast<double (*)[3]>(base + (size_t)a.maxB * 24);
    sm.key = reinterpret_cast<double*>(base + (size_t)a.maxB * 48);
    sm.mask = reinterpret_cast<uint32_t*>(base + (size_t)a.maxB * 56);
    sm.order = reinterpret_cast<uint16_t*>(base + (size_t)a.maxB * 56 + (size_t)a.maskWords * 4);
  }
  const unsigned FULL = 0xffffffffu;

  double P[3] = {0, 0, 0};
  double Pimp = 0;
  bool haveP = a.hasParent[slot] != 0;
  if (haveP) {
    P[0] = a.parent4[(size_t)slot * 4]; P[1] = a.parent4[(size_t)slot * 4 + 1]; P[2] = a.parent4[(size_t)slot * 4 + 2];
    Pimp = variance_calc(P[0], P[1], P[2]);
  }
  bool any = false;
  double bestGain = 0; int bestF = -1, bestS = -1, bestJ = -1;
  double bL[3] = {0, 0, 0}, bR[3] = {0, 0, 0};
  for (int q = lane; q < a.maskWords; q += 32) sm.mask[q] = 0u;
  int firstF = -1;
  const uint2* fo = a.slot_fo + (size_t)slot * a.Fsub;
  const double* hb = a.hist + (size_t)a.slot_base[slot] * 4;

  for (int j = 0; j < a.Fsub; j++) {
    const uint2 e = fo[j];
    const int f = (int)e.x;
    if (firstF < 0) firstF = f;
    const int nb = a.numBins[f];
    const int ns = nb - 1;
    if (ns <= 0) continue;
    const double* cells = hb + (size_t)e.y * 4;
    __syncwarp();
    for (int b = lane; b < nb; b += 32) {
      sm.raw[b][0] = cells[(size_t)b * 4]; sm.raw[b][1] = cells[(size_t)b * 4 + 1]; sm.raw[b][2] = cells[(size_t)b * 4 + 2];
    }
    __syncwarp();
    const bool cat = a.isCat[f] != 0;
    if (cat) {
      for (int b = lane; b < nb; b += 32) sm.key[b] = sm.raw[b][0] != 0 ? sm.raw[b][1] / sm.raw[b][0] : DBL_MAX;
      __syncwarp();
      for (int b = lane; b < nb; b += 32) {  // stable rank by counting == List.sortBy(_._2)
        const double kb = sm.key[b];
        int rank = 0;
        for (int c = 0; c < nb; c++) { const double kc = sm.key[c]; rank += (kc < kb) || (kc == kb && c < b); }
        sm.order[rank] = (uint16_t)b;
      }
    } else {
      for (int b = lane; b < nb; b += 32) sm.order[b] = (uint16_t)b;
    }
    __syncwarp();
    if (lane == 0) {  // sequential prefix in MLlib's association order: h[next] += h[cur]
      int o = sm.order[0];
      double c0 = sm.raw[o][0], c1 = sm.raw[o][1], c2 = sm.raw[o][2];
      sm.cum[0][0] = c0; sm.cum[0][1] = c1; sm.cum[0][2] = c2;
      for (int q = 1; q < nb; q++) {
        o = sm.order[q];
        c0 = sm.raw[o][0] + c0; c1 = sm.raw[o][1] + c1; c2 = sm.raw[o][2] + c2;
        sm.cum[q][0] = c0; sm.cum[q][1] = c1; sm.cum[q][2] = c2;
      }
    }
    __syncwarp();
    const double T0 = sm.cum[nb - 1][0], T1 = sm.cum[nb - 1][1], T2 = sm.cum[nb - 1][2];
    if (!haveP) {  // root: parent = left + right of the first evaluated candidate (calculateImpurityStats, stats == null)
      const double l0 = sm.cum[0][0], l1 = sm.cum[0][1], l2 = sm.cum[0][2];
      P[0] = l0 + (T0 - l0); P[1] = l1 + (T1 - l1); P[2] = l2 + (T2 - l2);
      Pimp = variance_calc(P[0], P[1], P[2]);
      haveP = true;
    }
    double lg = 0; int ls = -1;
    for (int s = lane; s < ns; s += 32) {
      const double L0 = sm.cum[s][0], L1 = sm.cum[s][1], L2 = sm.cum[s][2];
      const double R0 = T0 - L0, R1 = T1 - L1, R2 = T2 - L2;
      double gain;
      if (L0 < a.minInstances || R0 < a.minInstances || L0 < a.minWeight || R0 < a.minWeight) gain = -DBL_MAX;
      else {
        const double total = L0 + R0;
        const double li = variance_calc(L0, L1, L2), ri = variance_calc(R0, R1, R2);
        const double lw = L0 / total, rw = R0 / total;
        gain = Pimp - lw * li - rw * ri;
        if (gain < a.minInfoGain) gain = -DBL_MAX;
      }
      if (ls < 0 || gain > lg) { lg = gain; ls = s; }
    }
    // warp argmax: larger gain wins, ties -> lower split index (maxBy keeps the first maximum)
    for (int o = 16; o > 0; o >>= 1) {
      const double og = __shfl_xor_sync(FULL, lg, o);
      const int os = __shfl_xor_sync(FULL, ls, o);
      if (os >= 0 && (ls < 0 || og > lg || (og == lg && os < ls))) { lg = og; ls = os; }
    }
    if (!any || lg > bestGain) {  // strict >: the earliest feature in subset order wins ties
      any = true; bestGain = lg; bestF = f; bestS = ls; bestJ = j;
      bL[0] = sm.cum[ls][0]; bL[1] = sm.cum[ls][1]; bL[2] = sm.cum[ls][2];
      bR[0] = T0 - bL[0]; bR[1] = T1 - bL[1]; bR[2] = T2 - bL[2];
      // left set of the split = the first ls + 1 categories in centroid order (kept as a bit mask in shared memory)
      __syncwarp();
      for (int q = lane; q < a.maskWords; q += 32) sm.mask[q] = 0u;
      __syncwarp();
      if (cat) for (int q = lane; q <= ls; q += 32) { const int c = sm.order[q]; atomicOr(&sm.mask[c >> 5], 1u << (c & 31)); }
      __syncwarp();
    }
  }
  (void)bestJ;
  if (lane == 0) {
    SplitRec r;
    memset(&r, 0, sizeof r);
    if (!any) {  // no feature with candidate splits: dummy split, invalid stats on the node totals
      if (!haveP && firstF >= 0) {
        const double* cells = hb + (size_t)fo[0].y * 4;
        const int nb = a.numBins[firstF];
        for (int b = 0; b < nb; b++) { P[0] += cells[(size_t)b * 4]; P[1] += cells[(size_t)b * 4 + 1]; P[2] += cells[(size_t)b * 4 + 2]; }
        Pimp = variance_calc(P[0], P[1], P[2]);
      }
      r.gain = -DBL_MAX; r.valid = 0; r.feature = firstF < 0 ? 0 : firstF; r.splitIdx = 0;
    } else {
      r.gain = bestGain; r.valid = bestGain != -DBL_MAX; r.feature = bestF; r.splitIdx = bestS;
      for (int q = 0; q < 3; q++) { r.left[q] = bL[q]; r.right[q] = bR[q]; }
      r.left[3] = bL[0]; r.right[3] = bR[0];
      r.leftImp = variance_calc(bL[0], bL[1], bL[2]);
      r.rightImp = variance_calc(bR[0], bR[1], bR[2]);
    }
    r.impurity = Pimp;
    r.parent[0] = P[0]; r.parent[1] = P[1]; r.parent[2] = P[2]; r.parent[3] = P[0];
    a.out[slot] = r;
  }
  __syncwarp();
  for (int q = lane; q < a.maskWords; q += 32) a.catmask_out[(size_t)slot * a.maskWords + q] = any ? sm.mask[q] : 0u;
}

cudaError_t launch_best_split(const SplitLaunch& a, cudaStream_t s) {
  if (a.nSlots <= 0) return cudaSuccess;
  const size_t per = split_warp_bytes(a.maxB, a.maskWords);
  int warps = (int)(((size_t)200 * 1024) / per);
  if (warps > SPLIT_WARPS) warps = SPLIT_WARPS;
  if (warps < 1) return cudaErrorInvalidValue;   // maxBins too large for one warp's shared arrays (checked by the trainer)
  const size_t smem = per * warps;
  cudaError_t e = cudaFuncSetAttribute(best_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  best_split_kernel<<<(a.nSlots + warps - 1) / warps, warps * 32, smem, s>>>(a);
  return cudaGetLastError();
}

// =================================================================================================
// K1: findSplits
// =================================================================================================
__host__ __device__ inline uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}

__global__ void sample_rows_kernel(const double* X, int64_t ldx, int64_t n_rows, int64_t row0, int F, uint64_t seedWord, double fraction,
                                   double* out, int64_t cap, unsigned long long* count) {
  // one warp per row: lane 0 decides, the warp copies the row
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= n_rows) return;
  const uint64_t h = mix64(seedWord + (uint64_t)(row0 + row) * 0x9E3779B97F4A7C15ULL);
  const bool keep = ((double)(h >> 11) * 0x1.0p-53) < fraction;
  if (!keep) return;
  unsigned long long pos = 0;
  if (lane == 0) pos = atomicAdd(count, 1ULL);
  pos = __shfl_sync(0xffffffffu, pos, 0);
  if ((int64_t)pos >= cap) return;
  for (int f = lane; f < F; f += 32) out[(size_t)pos * F + f] = X[(size_t)row * ldx + f];
}

// One CTA per feature.  Workspace: F * n_pad doubles (sorted non-zero sample values, +inf padded).
__global__ void __launch_bounds__(1024) find_splits_kernel(const double* sample, int64_t n_s, int F, const int32_t* arity, int maxSplits,
                                                            double wns, double* thresholds, int32_t* nsplits, double* ws, int64_t n_pad) {
  const int f = blockIdx.x;
  if (arity[f] != 0) { if (threadIdx.x == 0) nsplits[f] = 0; return; }
  double* v = ws + (size_t)f * n_pad;
  __shared__ unsigned long long s_cnt;
  __shared__ unsigned long long s_distinct;
  if (threadIdx.x == 0) { s_cnt = 0; s_distinct = 0; }
  __syncthreads();
  unsigned long long mycnt = 0;
  for (int64_t i = threadIdx.x; i < n_pad; i += blockDim.x) {
    double x = INFINITY;
    if (i < n_s) { const double xv = sample[(size_t)i * F + f]; if (xv != 0.0) { x = xv; mycnt++; } }
    v[i] = x;
  }
  atomicAdd(&s_cnt, mycnt);
  __syncthreads();
  // bitonic sort ascending (n_pad is a power of two)
  for (int64_t kk = 2; kk <= n_pad; kk <<= 1) {
    for (int64_t jj = kk >> 1; jj > 0; jj >>= 1) {
      for (int64_t i = threadIdx.x; i < n_pad; i += blockDim.x) {
        const int64_t ixj = i ^ jj;
        if (ixj > i) {
          const double x = v[i], y = v[ixj];
          const bool up = (i & kk) == 0;
          if ((x > y) == up) { v[i] = y; v[ixj] = x; }
        }
      }
      __syncthreads();
    }
  }
  const int64_t c = (int64_t)s_cnt;  // number of non-zero sample values
  unsigned long long myd = 0;
  for (int64_t i = threadIdx.x; i < c; i += blockDim.x) if (i == 0 || v[i] != v[i - 1]) myd++;
  atomicAdd(&s_distinct, myd);
  __syncthreads();
  if (threadIdx.x != 0) return;
  double* out = thresholds + (size_t)f * maxSplits;
  if (c == 0) { nsplits[f] = 0; return; }
  const double partNum = (double)c;
  const double tol = 0x1.0p-52 * (double)c * 100;
  const bool hasZero = (wns - partNum) > tol;
  const double zeroCnt = wns - partNum;
  const int64_t distinct = (int64_t)s_distinct + (hasZero ? 1 : 0);
  const int64_t possible = distinct - 1;
  if (possible == 0) { nsplits[f] = 0; return; }
  const bool all = possible <= maxSplits;
  const double stride = wns / (double)(maxSplits + 1);
  // run iterator over the merged ascending sequence (negatives, [virtual zero], positives)
  int64_t i = 0;
  bool zeroDone = !hasZero;
  int nout = 0;
  bool firstRun = true;
  double prevVal = 0, cur = 0, target = stride;
  while (i < c || !zeroDone) {
    double val, cnt;
    if (!zeroDone && (i >= c || v[i] > 0.0)) { val = 0.0; cnt = zeroCnt; zeroDone = true; }
    else {
      val = v[i];
      int64_t e = i + 1;
      while (e < c && v[e] == val) e++;
      cnt = (double)(e - i);
      i = e;
    }
    if (firstRun) { cur = cnt; firstRun = false; }
    else {
      const double prev = cur;
      cur += cnt;
      if (all) { if (nout < maxSplits) out[nout++] = (prevVal + val) / 2.0; }
      else if (fabs(prev - target) < fabs(cur - target)) {
        if (nout < maxSplits) out[nout++] = (prevVal + val) / 2.0;
        target += stride;
      }
    }
    prevVal = val;
  }
  nsplits[f] = nout;
}

// =================================================================================================
// K2: binning (row-major f64 -> column-major u8)
// =================================================================================================
constexpr int BIN_FT = 32;    // features per block (lanes); large maxBins use fewer so the thresholds still fit shared memory
constexpr int BIN_RT = 128;   // rows per block
template <typename CodeT>
__global__ void __launch_bounds__(256) bin_kernel(const double* X, int64_t ldx, int64_t n_rows, int F, const double* thresholds,
                                                   const int32_t* nsplits, const int32_t* arity, int maxSplits, CodeT* bins, int64_t ldb,
                                                   int64_t row0, int32_t* err, CodeT* bins_rm, int64_t ldr, int FT) {
  extern __shared__ __align__(16) unsigned char smraw[];
  double* thr = reinterpret_cast<double*>(smraw);                                                  // [FT][maxSplits]
  CodeT* tile = reinterpret_cast<CodeT*>(smraw + (size_t)FT * (maxSplits > 0 ? maxSplits : 1) * 8);  // [FT][BIN_RT]
  const int f0 = blockIdx.y * FT;
  const int64_t r0 = (int64_t)blockIdx.x * BIN_RT;
  const int nf = min(FT, F - f0);
  for (int i = threadIdx.x; i < nf * maxSplits; i += blockDim.x) {
    const int ff = i / maxSplits, q = i % maxSplits;
    thr[ff * maxSplits + q] = (q < nsplits[f0 + ff]) ? thresholds[(size_t)(f0 + ff) * maxSplits + q] : INFINITY;
  }
  __syncthreads();
  const int fl = threadIdx.x % FT;   // feature lane
  const int rl = threadIdx.x / FT;   // 256 / FT rows per pass
  const int rstep = 256 / FT;
  for (int rr = rl; rr < BIN_RT; rr += rstep) {
    const int64_t row = r0 + rr;
    CodeT b = 0;
    if (row < n_rows && fl < nf) {
      const int f = f0 + fl;
      const double x = X[(size_t)row * ldx + f];
      const int ar = arity[f];
      if (ar == 0) {
        // TreePoint.findBin: Arrays.binarySearch -> number of thresholds strictly below x
        const double* t = thr + fl * maxSplits;
        int lo = 0, hi = nsplits[f];
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (t[mid] < x) lo = mid + 1; else hi = mid; }
        b = (CodeT)lo;
      } else {
        if (x < 0 || x >= (double)ar) atomicMax(err, f + 1);
        b = (CodeT)(int)x;
      }
    }
    tile[fl * BIN_RT + rr] = b;
    if (bins_rm != nullptr && row < n_rows && fl < nf) bins_rm[(size_t)(row0 + row) * ldr + f0 + fl] = b;   // row-major twin for gathers
  }
  __syncthreads();
  // coalesced column-major stores: 32-bit words
  constexpr int PER = 4 / (int)sizeof(CodeT);   // codes per word
  const int words = nf * (BIN_RT / PER);
  for (int i = threadIdx.x; i < words; i += blockDim.x) {
    const int ff = i / (BIN_RT / PER), wq = i % (BIN_RT / PER);
    const int64_t row = r0 + wq * PER;
    if (row + PER - 1 < n_rows && ((row0 + row) % PER) == 0) {
      *reinterpret_cast<uint32_t*>(bins + (size_t)(f0 + ff) * ldb + row0 + row) = *reinterpret_cast<const uint32_t*>(tile + ff * BIN_RT + wq * PER);
    } else {
      for (int q = 0; q < PER; q++) if (row + q < n_rows) bins[(size_t)(f0 + ff) * ldb + row0 + row + q] = tile[ff * BIN_RT + wq * PER + q];
    }
  }
}

// =================================================================================================
// K3: bagging.  One thread per emulated Spark partition; generator state in local memory.
// =================================================================================================
constexpr int BAG_THREADS = 64;
// Well19937c with its 624-word state in shared memory, laid out [word][thread] so that a thread's word w lives in bank
// (thread % 32) whatever w is: the per-thread indices diverge (variable draws per sample) without bank conflicts.
struct Well19937cSm {
  uint32_t* v;   // &state[0][threadIdx.x]; word w is v[w * BAG_THREADS]
  int index;
  // The recurrence's loop-carried values stay in registers: c0 = v[index] (the z4 of the previous step) and cTop = the word
  // whose top bit feeds z0.  Of the three words a step writes, two are overwritten before anything reads them, so only z3
  // is stored; the four old words a step needs are fetched one step ahead (p*), off the dependent chain.
  uint32_t c0, cTop, pM1, pM2, pM3, pR2;
  __device__ __forceinline__ uint32_t& at(int w) { return v[w * BAG_THREADS]; }
  __device__ __forceinline__ void prefetch() {
    int i1 = index + 70; if (i1 >= 624) i1 -= 624;
    int i2 = index + 179; if (i2 >= 624) i2 -= 624;
    int i3 = index + 449; if (i3 >= 624) i3 -= 624;
    int r2 = index - 2; if (r2 < 0) r2 += 624;
    pM1 = at(i1); pM2 = at(i2); pM3 = at(i3); pR2 = at(r2);
  }
  __device__ void seed(int64_t sd) {
    uint32_t a = (uint32_t)((uint64_t)sd >> 32), b = (uint32_t)((uint64_t)sd & 0xffffffffULL);
    at(0) = a; at(1) = b;
    for (int i = 2; i < 624; i++) {
      const int64_t l = (int64_t)(int32_t)a;
      const uint32_t c = (uint32_t)((1812433253LL * (l ^ (l >> 30)) + i) & 0xffffffffLL);
      at(i) = c;
      a = b; b = c;
    }
    index = 0;
    c0 = at(0); cTop = at(623);
    prefetch();
  }
  __device__ __forceinline__ uint32_t next(int bits) {
    const uint32_t v0 = c0, vM1 = pM1, vM2 = pM2, vM3 = pM3, vR2 = pR2;
    const uint32_t z0 = (0x80000000u & cTop) ^ (0x7FFFFFFFu & vR2);
    const uint32_t z1 = (v0 ^ (v0 << 25)) ^ (vM1 ^ (vM1 >> 27));
    const uint32_t z2 = (vM2 >> 9) ^ (vM3 ^ (vM3 >> 1));
    const uint32_t z3 = z1 ^ z2;
    uint32_t z4 = z0 ^ (z1 ^ (z1 << 9)) ^ (z2 ^ (z2 << 21)) ^ (z3 ^ (z3 >> 21));
    at(index) = z3;               // v[index-1] = z4 and v[index-2] &= 0x80000000 are dead stores: both words are rewritten
    c0 = z4;                      // (as z3 of the next two steps) before they are read from memory
    cTop = vR2;
    index = index == 0 ? 623 : index - 1;
    prefetch();
    z4 ^= (z4 << 7) & 0xe46e1700u;
    z4 ^= (z4 << 15) & 0x9b868000u;
    return z4 >> (32 - bits);
  }
  __device__ __forceinline__ double nextDouble() {
    const int64_t high = (int64_t)next(26) << 26;
    const int64_t low = (int64_t)next(26);
    return (double)(high | low) * 0x1.0p-52;
  }
};

__device__ inline uint32_t rotl32d(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
__device__ uint32_t murmur3_8bytes(const uint8_t* data, uint32_t seed) {
  uint32_t h = seed;
  for (int i = 0; i < 8; i += 4) {
    uint32_t k = (uint32_t)data[i] | ((uint32_t)data[i + 1] << 8) | ((uint32_t)data[i + 2] << 16) | ((uint32_t)data[i + 3] << 24);
    k *= 0xcc9e2d51u; k = rotl32d(k, 15); k *= 0x1b873593u;
    h ^= k; h = rotl32d(h, 13); h = h * 5 + 0xe6546b64u;
  }
  h ^= 8u;
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
struct XorShiftDev {
  uint64_t s;
  __device__ void seed(int64_t init) {
    uint8_t b[8];
    for (int i = 0; i < 8; i++) b[i] = (uint8_t)((uint64_t)init >> (56 - 8 * i));
    const uint32_t lo = murmur3_8bytes(b, 0x3c074a61u);
    const uint32_t hi = murmur3_8bytes(b, lo);
    s = ((uint64_t)hi << 32) | (uint64_t)lo;
  }
  __device__ __forceinline__ int32_t next(int bits) {
    uint64_t n = s ^ (s << 21);
    n ^= n >> 35;
    n ^= n << 4;
    s = n;
    return (int32_t)(n & ((1ULL << bits) - 1));
  }
  __device__ __forceinline__ double nextDouble() { return (double)(((int64_t)next(26) << 27) + next(27)) * 0x1.0p-53; }
};

// mode 0: Poisson (with replacement); 1: Bernoulli (without replacement)
__global__ void __launch_bounds__(BAG_THREADS) bag_kernel(int64_t seed, const int64_t* part_offsets, int P, int part0, int T, double rate, double expNegRate,
                                                          int mode, uint8_t* w, int64_t ldw, int32_t* node, int64_t ldn) {
  extern __shared__ uint32_t bag_state[];   // [624][BAG_THREADS]
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const int64_t r0 = part_offsets[p], r1 = part_offsets[p + 1];
  if (mode == 0) {
    Well19937cSm rng;
    rng.v = bag_state + threadIdx.x;
    rng.seed(seed + (int64_t)(part0 + p) + 1);
    const double lim = 1000 * rate;
    // One uniform per loop trip for every lane: PoissonDistribution.sample() consumes a data-dependent number of uniforms, and
    // a per-sample inner loop would make each warp pay for its slowest lane on every sample.  Flattened, lanes only diverge
    // on the short "sample finished" branch.
    int64_t i = r0;
    int t = 0;
    long n = 0;
    double r = 1.0;
    while (i < r1) {
      bool finished = true;
      if ((double)n < lim) {
        r *= rng.nextDouble();
        if (r >= expNegRate) { n++; finished = false; }
      }
      if (finished) {
        const int c = n > 255 ? 255 : (int)n;
        w[(size_t)t * ldw + i] = (uint8_t)c;
        if (node != nullptr) node[(size_t)t * ldn + i] = c > 0 ? 0 : -1;
        n = 0; r = 1.0;
        if (++t == T) { t = 0; i++; }
      }
    }
  } else if (mode == 1) {
    XorShiftDev rng;
    rng.seed(seed + (int64_t)(part0 + p) + 1);
    for (int64_t i = r0; i < r1; i++)
      for (int t = 0; t < T; t++) {
        const int c = rng.nextDouble() < rate ? 1 : 0;
        w[(size_t)t * ldw + i] = (uint8_t)c;
        if (node != nullptr) node[(size_t)t * ldn + i] = c > 0 ? 0 : -1;
      }
  }
}
// ---------------------------------------------------------------------------------------------------------------------
// K3, warp-parallel form: ONE WARP per emulated partition instead of one thread.
//
// Well19937c is serial only through one 32-bit word: with x_k the word the step at ring index i-k reads from v[i-k],
//     x_{k+1} = M(x_k) ^ d_k,   M(x) = L(x ^ (x << 25)),  L(t) = (t << 9) ^ (t >>> 21),
//     d_k = L(a_k) ^ z0_k ^ (z2_k << 21) ^ (z2_k >>> 21)        (the z1 and z2 terms of z3 ^ (z3 >>> 21) cancel in z4)
// where a_k, z0_k, z2_k only involve ring words written at least 70 steps earlier.  M is linear over GF(2), so 32 consecutive
// steps are a prefix scan of affine maps: five shuffle stages, stage s applying M^(2^s) through four 256-entry byte tables.
// The Poisson sampler consumes a data-dependent number of uniforms per sample; it is parallelised exactly (same products in
// the same order as PoissonDistribution.sample): every lane computes the sample that WOULD start at its uniform, lane 0
// finishes the sample carried in from the previous window and walks the start positions, marked lanes emit.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int BAGW_WARPS = 4;
__device__ __forceinline__ uint32_t well_M(uint32_t x) { const uint32_t t = x ^ (x << 25); return (t << 9) ^ (t >> 21); }

struct BagWarpSmem {
  uint32_t v[624];
  double u[32];
  uint32_t len[32];
};

__device__ __forceinline__ uint32_t well_apply(const uint32_t* tab /* [4][256] */, uint32_t y) {
  return tab[y & 255u] ^ tab[256 + ((y >> 8) & 255u)] ^ tab[512 + ((y >> 16) & 255u)] ^ tab[768 + (y >> 24)];
}

// 32 generator steps; returns this lane's tempered 32-bit output (step `lane`), updates ring, index and the carried word.
__device__ __forceinline__ uint32_t well_batch32(uint32_t* v, int& index, uint32_t& c0, const uint32_t* tabs, int lane) {
  int i = index - lane; if (i < 0) i += 624;
  int i1 = i + 70;  if (i1 >= 624) i1 -= 624;
  int i2 = i + 179; if (i2 >= 624) i2 -= 624;
  int i3 = i + 449; if (i3 >= 624) i3 -= 624;
  int r1 = i - 1;   if (r1 < 0) r1 += 624;
  int r2 = i - 2;   if (r2 < 0) r2 += 624;
  const uint32_t vM1 = v[i1], vM2 = v[i2], vM3 = v[i3];
  const uint32_t z0 = (0x80000000u & v[r1]) ^ (0x7FFFFFFFu & v[r2]);
  const uint32_t a = vM1 ^ (vM1 >> 27);
  const uint32_t z2 = (vM2 >> 9) ^ (vM3 ^ (vM3 >> 1));
  const uint32_t la = (a << 9) ^ (a >> 21);
  uint32_t b = la ^ z0 ^ (z2 << 21) ^ (z2 >> 21);          // d_k
  if (lane == 0) b ^= well_M(c0);                           // fold the carried word into the first element
#pragma unroll
  for (int st = 0; st < 5; st++) {
    const int o = 1 << st;
    const uint32_t y = __shfl_up_sync(0xffffffffu, b, o);
    if (lane >= o) b ^= well_apply(tabs + st * 1024, y);
  }
  // b == x_{k+1} == z4 of step k
  uint32_t x = __shfl_up_sync(0xffffffffu, b, 1);
  if (lane == 0) x = c0;
  const uint32_t z1 = (x ^ (x << 25)) ^ a;
  __syncwarp();
  v[i] = z1 ^ z2;                                            // z3; the other two words a step writes are dead (see Well19937cSm)
  __syncwarp();
  c0 = __shfl_sync(0xffffffffu, b, 31);
  index -= 32; if (index < 0) index += 624;
  uint32_t z4 = b;
  z4 ^= (z4 << 7) & 0xe46e1700u;
  z4 ^= (z4 << 15) & 0x9b868000u;
  return z4;
}

__global__ void __launch_bounds__(BAGW_WARPS * 32) bag_warp_kernel(int64_t seed, const int64_t* part_offsets, int P, int part0, int T, double rate,
                                                                   double expNegRate, uint8_t* w, int64_t ldw, int32_t* node, int64_t ldn) {
  __shared__ uint32_t tabs[5 * 1024];              // M^(2^s), s = 0..4, as 4 x 256 byte tables
  __shared__ BagWarpSmem sm_all[BAGW_WARPS];
  // tables: entry (s, byte, val) = M^(2^s)(val << 8*byte)
  for (int e = threadIdx.x; e < 5 * 1024; e += blockDim.x) {
    const int st = e >> 10, byte = (e >> 8) & 3, val = e & 255;
    uint32_t x = (uint32_t)val << (8 * byte);
    for (int r = 0; r < (1 << st); r++) x = well_M(x);
    tabs[e] = x;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int p = blockIdx.x * BAGW_WARPS + warp;
  if (p >= P) return;
  BagWarpSmem& sm = sm_all[warp];
  const unsigned FULL = 0xffffffffu;
  // AbstractWell.setSeed: v[0..1] = (hi, lo); v[i] = 1812433253 * (l ^ (l >> 30)) + i with l = (long) v[i - 2]: two serial chains
  {
    const int64_t sd = seed + (int64_t)(part0 + p) + 1;
    if (lane < 2) {
      uint32_t cur = lane == 0 ? (uint32_t)((uint64_t)sd >> 32) : (uint32_t)((uint64_t)sd & 0xffffffffULL);
      sm.v[lane] = cur;
      for (int i = lane + 2; i < 624; i += 2) {
        const int64_t l = (int64_t)(int32_t)cur;
        cur = (uint32_t)((1812433253LL * (l ^ (l >> 30)) + i) & 0xffffffffLL);
        sm.v[i] = cur;
      }
    }
  }
  __syncwarp();
  int index = 0;
  uint32_t c0 = sm.v[0];
  const int64_t r0 = part_offsets[p], r1 = part_offsets[p + 1];
  const int64_t total = (r1 - r0) * (int64_t)T;     // samples of this partition, row-major over trees
  const long limI = (long)ceil(1000 * rate);
  int64_t ord = 0;                                   // samples emitted so far == ordRow * T + ordT (kept split: no 64-bit divisions)
  int64_t ordRow = 0; uint32_t ordT = 0;
  double rc = 1.0; long nc = 0;                      // the sample that is in progress across windows
  auto emit = [&](uint32_t k, long n) {              // sample number ord + k (k <= 32)
    const uint32_t tt = ordT + k;
    const int64_t row = r0 + ordRow + (int64_t)(tt / (uint32_t)T);
    const uint32_t t = tt % (uint32_t)T;
    const int c = n > 255 ? 255 : (int)n;
    w[(size_t)t * ldw + row] = (uint8_t)c;
    if (node != nullptr) node[(size_t)t * ldn + row] = c > 0 ? 0 : -1;
  };
  while (ord < total) {
    // ---- 64 generator steps -> 32 uniforms (BitsStreamGenerator.nextDouble: two 26-bit halves) ----
#pragma unroll
    for (int half = 0; half < 2; half++) {
      const uint32_t out = well_batch32(sm.v, index, c0, tabs, lane) >> 6;
      const uint32_t hi = __shfl_sync(FULL, out, (2 * lane) & 31), lo = __shfl_sync(FULL, out, (2 * lane + 1) & 31);
      if (lane < 16) sm.u[half * 16 + lane] = (double)(((int64_t)hi << 26) | (int64_t)lo) * 0x1.0p-52;
    }
    __syncwarp();
    // ---- every lane: the sample that would START at uniform `lane` (r = 1; r *= u; ends when r < p) ----
    uint32_t mylen = 0; double myr = 1.0;            // mylen == 0: runs past the window (myr = product so far)
    for (int k = lane; k < 32; k++) { myr *= sm.u[k]; if (!(myr >= expNegRate)) { mylen = (uint32_t)(k - lane + 1); break; } }
    sm.len[lane] = mylen;
    __syncwarp();
    // ---- lane 0 finishes the carried-in sample; the start positions after it are found by pointer doubling ----
    int first_done = 0; long first_n = 0; int s0 = 32;
    if (lane == 0) {
      int pos = 0; bool open = true;
      while (pos < 32) {
        if (nc >= limI) { open = false; break; }      // iteration cap of sample(): ends without consuming (practically never)
        rc *= sm.u[pos]; pos++;
        if (rc >= expNegRate) nc++; else { open = false; break; }
      }
      if (!open) { first_done = 1; first_n = nc; rc = 1.0; nc = 0; s0 = pos; }
    }
    first_done = __shfl_sync(FULL, first_done, 0);
    s0 = __shfl_sync(FULL, s0, 0);
    // samples start at s0, s0 + len[s0], ... until a chain runs past the window.  J = position after this lane's chain
    // (>= 32: leaves the window, 64: chain incomplete); S = set of reached starts, doubled five times (up to 31 hops).
    uint32_t J = mylen == 0 ? 64u : (uint32_t)lane + mylen;
    uint32_t S = (first_done && s0 < 32) ? (1u << s0) : 0u;
#pragma unroll
    for (int rnd = 0; rnd < 5; rnd++) {
      const uint32_t contrib = (((S >> lane) & 1u) && J < 32u) ? (1u << J) : 0u;
      S |= __reduce_or_sync(FULL, contrib);
      const uint32_t Jn = __shfl_sync(FULL, J, J & 31u);
      if (J < 32u) J = Jn;
    }
    const uint32_t openMask = __ballot_sync(FULL, ((S >> lane) & 1u) && mylen == 0);   // at most one reached chain is incomplete
    const int carry_pos = openMask ? (__ffs(openMask) - 1) : -1;
    const uint32_t marks = S & ~openMask;
    // ---- emit ----
    if (first_done) {
      if (lane == 0 && ord < total) emit(0u, first_n);
      if ((marks >> lane) & 1u) {
        const uint32_t k = 1u + (uint32_t)__popc(marks & ((1u << lane) - 1u));
        if (ord + k < total) emit(k, (long)mylen - 1);
      }
      const uint32_t done = 1u + (uint32_t)__popc(marks);
      ord += done;
      ordT += done; ordRow += ordT / (uint32_t)T; ordT %= (uint32_t)T;
      // the sample left open at the window's end (if any) continues in the next window
      const double cr = __shfl_sync(FULL, myr, carry_pos < 0 ? 0 : carry_pos);
      if (carry_pos >= 0) { rc = cr; nc = 32 - carry_pos; } else { rc = 1.0; nc = 0; }
    }
    // lane 0 holds the authoritative in-progress sample in both cases (it alone advanced it when the carried-in sample did
    // not finish inside the window)
    rc = __shfl_sync(FULL, rc, 0);
    nc = (long)__shfl_sync(FULL, (long long)nc, 0);
    __syncwarp();
  }
}

__global__ void bag_ones_kernel(int64_t n, int T, uint8_t* w, int64_t ldw, int32_t* node, int64_t ldn) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int t = 0; t < T; t++) { w[(size_t)t * ldw + i] = 1; if (node != nullptr) node[(size_t)t * ldn + i] = 0; }
}

// =================================================================================================
// K7: predict
// =================================================================================================
__global__ void predict_kernel(const double* X, int64_t ldx, int64_t n_rows, int T, const int64_t* tree_off, const int32_t* feature,
                               const double* threshold, const int32_t* left, const int32_t* right, const double* prediction,
                               const int32_t* ncat, const int32_t* cat_off, const int32_t* cat_len, const double* cats, double* out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows) return;
  const double* x = X + (size_t)i * ldx;
  double acc = 0.0;
  for (int t = 0; t < T; t++) {
    const int64_t base = tree_off[t];
    int n = 0;
    int f;
    while ((f = feature[base + n]) >= 0) {
      const double xv = x[f];
      bool l;
      if (ncat[base + n] == 0) l = xv <= threshold[base + n];
      else {
        l = false;
        const int co = cat_off[base + n], cl = cat_len[base + n];
        for (int c = 0; c < cl; c++) if (cats[co + c] == xv) { l = true; break; }
      }
      n = l ? left[base + n] : right[base + n];
    }
    acc += prediction[base + n];
  }
  out[i] = acc / (double)T;
}

}  // namespace b2rf

// =================================================================================================
// C-ABI wrappers for the stand-alone kernels
// =================================================================================================
using namespace b2rf;

extern "C" {

int b2rf_version(void) { return 100; }
const char* b2rf_last_error(void) { return b2rf::last_error_cstr(); }

double b2rf_samples_fraction(int64_t numExamples, int32_t metaMaxBins) {
  const int64_t required = (int64_t)metaMaxBins * metaMaxBins > 10000 ? (int64_t)metaMaxBins * metaMaxBins : 10000;
  return required < numExamples ? (double)required / (double)numExamples : 1.0;
}

// XORShiftRandom(seed).nextInt() — host copy of the hashSeed chain, used to derive the row-sample seed word
static uint32_t h_rotl(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
static uint32_t h_murmur8(const uint8_t* d, uint32_t seed) {
  uint32_t h = seed;
  for (int i = 0; i < 8; i += 4) {
    uint32_t k = (uint32_t)d[i] | ((uint32_t)d[i + 1] << 8) | ((uint32_t)d[i + 2] << 16) | ((uint32_t)d[i + 3] << 24);
    k *= 0xcc9e2d51u; k = h_rotl(k, 15); k *= 0x1b873593u; h ^= k; h = h_rotl(h, 13); h = h * 5 + 0xe6546b64u;
  }
  h ^= 8u; h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
static int32_t xorshift_first_int(int64_t seed) {
  uint8_t b[8];
  for (int i = 0; i < 8; i++) b[i] = (uint8_t)((uint64_t)seed >> (56 - 8 * i));
  const uint32_t lo = h_murmur8(b, 0x3c074a61u), hi = h_murmur8(b, lo);
  uint64_t s = ((uint64_t)hi << 32) | lo;
  uint64_t n = s ^ (s << 21); n ^= n >> 35; n ^= n << 4;
  return (int32_t)(n & 0xffffffffULL);
}

int b2rf_sample_rows(const double* d_X, int64_t ldx, int64_t n_rows, int64_t row0_global, int32_t F, int64_t seed, double fraction,
                     double* d_sample_out, int64_t cap, int64_t* d_count, void* stream) {
  if (!d_X || !d_sample_out || !d_count || F <= 0) { set_error("b2rf_sample_rows: bad argument"); return B2RF_ERR_ARG; }
  if (n_rows <= 0) return B2RF_OK;
  const uint64_t seedWord = ((uint64_t)(uint32_t)xorshift_first_int(seed) << 32) | 0x9E3779B9u;
  const int64_t threads = n_rows * 32;
  sample_rows_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_X, ldx, n_rows, row0_global, F, seedWord, fraction,
                                                                                        d_sample_out, cap, (unsigned long long*)d_count);
  return cuda_ok(cudaGetLastError(), "b2rf_sample_rows") ? B2RF_OK : B2RF_ERR_CUDA;
}

int64_t b2rf_sample_indices(int64_t seed, int64_t row0_global, int64_t n_rows, double fraction, int64_t* h_idx_out, int64_t cap) {
  const uint64_t seedWord = ((uint64_t)(uint32_t)xorshift_first_int(seed) << 32) | 0x9E3779B9u;
  int64_t k = 0;
  for (int64_t i = 0; i < n_rows; i++) {
    const uint64_t h = mix64(seedWord + (uint64_t)(row0_global + i) * 0x9E3779B97F4A7C15ULL);
    if (((double)(h >> 11) * 0x1.0p-53) < fraction) { if (k < cap) h_idx_out[k] = i; k++; }
  }
  return k;
}

static int64_t next_pow2(int64_t n) { int64_t p = 1; while (p < n) p <<= 1; return p; }

size_t b2rf_find_splits_workspace_bytes(int64_t n_sample, int32_t F) { return (size_t)next_pow2(n_sample < 2 ? 2 : n_sample) * (size_t)F * 8; }

int b2rf_find_splits(const double* d_sample, int64_t n_sample, int32_t F, const int32_t* d_arity, int32_t maxPossibleBins,
                     double weightedNumSamples, double* d_thresholds, int32_t* d_nsplits, void* d_workspace, size_t workspace_bytes,
                     void* stream) {
  if (F <= 0 || n_sample <= 0 || maxPossibleBins < 1) { set_error("b2rf_find_splits: bad argument"); return B2RF_ERR_ARG; }
  if (workspace_bytes < b2rf_find_splits_workspace_bytes(n_sample, F)) { set_error("b2rf_find_splits: workspace too small"); return B2RF_ERR_ARG; }
  const int64_t n_pad = next_pow2(n_sample < 2 ? 2 : n_sample);
  find_splits_kernel<<<F, 1024, 0, (cudaStream_t)stream>>>(d_sample, n_sample, F, d_arity, maxPossibleBins - 1, weightedNumSamples, d_thresholds,
                                                            d_nsplits, (double*)d_workspace, n_pad);
  return cuda_ok(cudaGetLastError(), "b2rf_find_splits") ? B2RF_OK : B2RF_ERR_CUDA;
}

// ---- split finding without any allocation ----------------------------------------------------------------------
// Some hosts of this library cannot call the CUDA allocator at all: a Spark executor plugin (or any serving process) whose
// own memory pool has already reserved the whole GPU leaves nothing for cudaMalloc to hand out, and a latency-critical
// caller cannot accept the device-wide synchronisation that cudaMalloc / cudaFree imply.  Such a caller can still lend
// input and output buffers from its pool, but b2rf_find_splits also needs a sort workspace.  This variant keeps that
// workspace in a fixed array that is part of the loaded module itself.
//
//  * The array is B2RF_NOALLOC_WORKSPACE_BYTES (8 MiB) on EVERY GPU that loads libb2rf.so, from load until exit.
//  * Requests needing more (next_pow2(n_sample) * F * 8 bytes) are refused with B2RF_ERR_ARG; use b2rf_find_splits then.
//  * The first call on each device looks the array's device address up (one cudaGetSymbolAddress, no allocation) with
//    that device current.  Make that first call outside of any stream capture.
//  * All calls on a device share the one array and it keeps whatever the previous call left in it.  Nothing in it has
//    to start at zero: the kernel overwrites every element it later reads.  Two calls whose kernels overlap on the GPU
//    (different streams, or different host threads) would sort in the same memory and corrupt each other — serialise them.
static __device__ __align__(256) unsigned char g_noalloc_workspace[B2RF_NOALLOC_WORKSPACE_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_addr[kMaxDevices] = {};   // device address of g_noalloc_workspace, per device ordinal

// In host code `g_noalloc_workspace` is only a handle for the symbol, not a device address; the address a kernel can use
// comes from cudaGetSymbolAddress, queried once per device while that device is current.
int noalloc_workspace(void** out) {
  int dev = -1;
  if (!cuda_ok(cudaGetDevice(&dev), "b2rf_find_splits_noalloc: cudaGetDevice")) return B2RF_ERR_CUDA;
  if (dev < 0 || dev >= kMaxDevices) { set_error("b2rf_find_splits_noalloc: device ordinal out of range"); return B2RF_ERR_ARG; }
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_addr[dev] == nullptr) {
    void* ptr = nullptr;
    if (!cuda_ok(cudaGetSymbolAddress(&ptr, g_noalloc_workspace), "b2rf_find_splits_noalloc: cudaGetSymbolAddress")) return B2RF_ERR_CUDA;
    if (ptr == nullptr) { set_error("b2rf_find_splits_noalloc: symbol address is null"); return B2RF_ERR_CUDA; }
    g_noalloc_addr[dev] = ptr;
  }
  *out = g_noalloc_addr[dev];
  return B2RF_OK;
}
}  // namespace

size_t b2rf_find_splits_noalloc_max_bytes(void) { return (size_t)B2RF_NOALLOC_WORKSPACE_BYTES; }

int b2rf_find_splits_noalloc(const double* d_sample, int64_t n_sample, int32_t F, const int32_t* d_arity, int32_t maxPossibleBins,
                             double weightedNumSamples, double* d_thresholds, int32_t* d_nsplits, void* stream) {
  if (F <= 0 || n_sample <= 0 || maxPossibleBins < 1 || !d_sample || !d_arity || !d_thresholds || !d_nsplits) {
    set_error("b2rf_find_splits_noalloc: bad argument"); return B2RF_ERR_ARG;
  }
  const size_t need = b2rf_find_splits_workspace_bytes(n_sample, F);
  if (need > (size_t)B2RF_NOALLOC_WORKSPACE_BYTES) {
    char buf[256];
    snprintf(buf, sizeof buf, "b2rf_find_splits_noalloc: this sample needs a %zu-byte workspace but the built-in one holds %zu bytes "
             "(next_pow2(n_sample) * F * 8 must not exceed it); call b2rf_find_splits with a caller-provided workspace instead",
             need, (size_t)B2RF_NOALLOC_WORKSPACE_BYTES);
    set_error(buf);
    return B2RF_ERR_ARG;
  }
  void* ws = nullptr;
  const int rc = noalloc_workspace(&ws);
  if (rc != B2RF_OK) return rc;
  const int64_t n_pad = next_pow2(n_sample < 2 ? 2 : n_sample);
  find_splits_kernel<<<F, 1024, 0, (cudaStream_t)stream>>>(d_sample, n_sample, F, d_arity, maxPossibleBins - 1, weightedNumSamples, d_thresholds,
                                                            d_nsplits, static_cast<double*>(ws), n_pad);
  return cuda_ok(cudaGetLastError(), "b2rf_find_splits_noalloc") ? B2RF_OK : B2RF_ERR_CUDA;
}

int b2rf_bin(const double* d_X, int64_t ldx, int64_t n_rows, int32_t F, const double* d_thresholds, const int32_t* d_nsplits,
             const int32_t* d_arity, int32_t maxSplits, void* d_bins, int64_t ldb, int64_t row0, int32_t* d_err, void* d_bins_rm,
             int64_t ldr, int32_t code_bytes, void* stream) {
  if (F <= 0 || maxSplits < 0 || (code_bytes != 1 && code_bytes != 2) || (code_bytes == 1 && maxSplits > 255) || maxSplits > 65534) {
    set_error("b2rf_bin: code_bytes must be 1 (maxBins <= 256) or 2 (maxBins <= 65535)"); return B2RF_ERR_ARG;
  }
  if (n_rows <= 0) return B2RF_OK;
  const int ms = maxSplits > 0 ? maxSplits : 1;
  int FT = BIN_FT;
  while (FT > 1 && (size_t)FT * ms * 8 + (size_t)FT * BIN_RT * code_bytes > (size_t)160 * 1024) FT /= 2;
  const size_t smem = (size_t)FT * ms * 8 + (size_t)FT * BIN_RT * code_bytes;
  if (smem > (size_t)220 * 1024) { set_error("b2rf_bin: maxBins too large for the shared-memory threshold table"); return B2RF_ERR_ARG; }
  dim3 grid((unsigned)((n_rows + BIN_RT - 1) / BIN_RT), (unsigned)((F + FT - 1) / FT));
  cudaStream_t s = (cudaStream_t)stream;
  if (code_bytes == 1) {
    if (!cuda_ok(cudaFuncSetAttribute(bin_kernel<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "b2rf_bin attr")) return B2RF_ERR_CUDA;
    bin_kernel<uint8_t><<<grid, 256, smem, s>>>(d_X, ldx, n_rows, F, d_thresholds, d_nsplits, d_arity, ms, (uint8_t*)d_bins, ldb, row0, d_err,
                                               (uint8_t*)d_bins_rm, ldr, FT);
  } else {
    if (!cuda_ok(cudaFuncSetAttribute(bin_kernel<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "b2rf_bin attr")) return B2RF_ERR_CUDA;
    bin_kernel<uint16_t><<<grid, 256, smem, s>>>(d_X, ldx, n_rows, F, d_thresholds, d_nsplits, d_arity, ms, (uint16_t*)d_bins, ldb, row0, d_err,
                                                (uint16_t*)d_bins_rm, ldr, FT);
  }
  return cuda_ok(cudaGetLastError(), "b2rf_bin") ? B2RF_OK : B2RF_ERR_CUDA;
}

int b2rf_bag(int64_t seed, const int64_t* h_part_offsets, int32_t P, int32_t part0, int32_t T, double subsamplingRate, int32_t bootstrap,
             uint8_t* d_w, int64_t ldw, int32_t* d_node, int64_t ldn, void* stream) {
  if (P <= 0 || T <= 0 || !h_part_offsets) { set_error("b2rf_bag: bad argument"); return B2RF_ERR_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t n = h_part_offsets[P];
  if (!bootstrap && T == 1 && subsamplingRate == 1.0) {
    if (n > 0) bag_ones_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(n, T, d_w, ldw, d_node, ldn);
    return cuda_ok(cudaGetLastError(), "b2rf_bag") ? B2RF_OK : B2RF_ERR_CUDA;
  }
  int64_t* d_po = nullptr;
  if (!cuda_ok(cudaMallocAsync((void**)&d_po, (size_t)(P + 1) * 8, s), "b2rf_bag alloc")) return B2RF_ERR_CUDA;
  if (!cuda_ok(cudaMemcpyAsync(d_po, h_part_offsets, (size_t)(P + 1) * 8, cudaMemcpyHostToDevice, s), "b2rf_bag copy")) return B2RF_ERR_CUDA;
  // exp(-rate) is evaluated once on the host so that every consumer of the stream compares against the same constant
  static const bool serialOnly = [] { const char* m = std::getenv("B2RF_BAG_SERIAL"); return m && *m == '1'; }();
  bool ok = true;
  if (bootstrap && 1000 * subsamplingRate >= 64 && !serialOnly) {   // warp per partition (window of 32 uniforms < iteration cap)
    bag_warp_kernel<<<(P + BAGW_WARPS - 1) / BAGW_WARPS, BAGW_WARPS * 32, 0, s>>>(seed, d_po, P, part0, T, subsamplingRate, std::exp(-subsamplingRate),
                                                                                d_w, ldw, d_node, ldn);
    ok = cuda_ok(cudaGetLastError(), "b2rf_bag");
  } else {
    const size_t smem = (size_t)624 * BAG_THREADS * 4;
    ok = cuda_ok(cudaFuncSetAttribute(bag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "b2rf_bag attr");
    if (ok) {
      bag_kernel<<<(P + BAG_THREADS - 1) / BAG_THREADS, BAG_THREADS, smem, s>>>(seed, d_po, P, part0, T, subsamplingRate, std::exp(-subsamplingRate),
                                                                               bootstrap ? 0 : 1, d_w, ldw, d_node, ldn);
      ok = cuda_ok(cudaGetLastError(), "b2rf_bag");
    }
  }
  cudaFreeAsync(d_po, s);   // stream-ordered.  No host sync: cudaMemcpyAsync from pageable memory stages h_part_offsets before returning
  return ok ? B2RF_OK : B2RF_ERR_CUDA;
}

int b2rf_hist_build(const void* d_bins, int64_t ldb, int32_t code_bytes, int32_t F, const double* d_y, const uint8_t* d_w, int64_t ldw,
                    const int32_t* d_slot_of_row, int64_t ldn, const int32_t* d_tree_ids, int32_t Tg, int64_t n_rows,
                    const uint32_t* d_slot_fo, const int64_t* d_slot_base, int32_t Fsub, double* d_hist, void* stream) {
  if (code_bytes != 1 && code_bytes != 2) { set_error("b2rf_hist_build: code_bytes must be 1 or 2"); return B2RF_ERR_ARG; }
  if ((ldb & 15) != 0 || ldb < n_rows) { set_error("b2rf_hist_build: ldb must be a multiple of 16 and >= n_rows (128 recommended)"); return B2RF_ERR_ARG; }
  HistLaunch a;
  a.bins = static_cast<const uint8_t*>(d_bins); a.ldb = ldb; a.F = F; a.y = d_y; a.w = d_w; a.ldw = ldw; a.codeBytes = code_bytes; a.maskWords = 8;
  a.node = const_cast<int32_t*>(d_slot_of_row); a.ldn = ldn; a.tree_ids = d_tree_ids; a.Tg = Tg; a.n_rows = n_rows;
  a.dnodes = nullptr; a.nodeStride = 0; a.catmasks = nullptr;
  a.slot_fo = reinterpret_cast<const uint2*>(d_slot_fo); a.slot_base = d_slot_base; a.Fsub = Fsub; a.hist = d_hist; a.slot_count = nullptr;
  std::string err;
  cudaError_t e = launch_hist(a, 0, (cudaStream_t)stream, &err);
  if (e != cudaSuccess) { if (!err.empty()) set_error(err); else cuda_ok(e, "b2rf_hist_build"); return e == cudaErrorInvalidValue && !err.empty() ? B2RF_ERR_ARG : B2RF_ERR_CUDA; }
  return B2RF_OK;
}

int b2rf_predict(const double* d_X, int64_t ldx, int64_t n_rows, int32_t T, const int64_t* d_tree_off, const int32_t* d_feature,
                 const double* d_threshold, const int32_t* d_left, const int32_t* d_right, const double* d_prediction, const int32_t* d_ncat,
                 const int32_t* d_cat_off, const int32_t* d_cat_len, const double* d_cats, double* d_out, void* stream) {
  if (n_rows <= 0) return B2RF_OK;
  predict_kernel<<<(unsigned)((n_rows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_X, ldx, n_rows, T, d_tree_off, d_feature, d_threshold, d_left, d_right,
                                                                                     d_prediction, d_ncat, d_cat_off, d_cat_len, d_cats, d_out);
  return cuda_ok(cudaGetLastError(), "b2rf_predict") ? B2RF_OK : B2RF_ERR_CUDA;
}

}  // extern "C"
