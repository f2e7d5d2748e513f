// b2rf (sm_100a) — K4 histogram build (DTStatsAggregator.update via binSeqOp): shared-memory histograms over per-node row segments (default), per-group counting sort, column scan + global RED; K6 routing fused into the scan kernels.
// Compiled with -fmad=false: MLlib's arithmetic is unfused JVM double arithmetic.
#include "b2rf_internal.h"
#include "b2rf_device.cuh"
#include "../../include/b2rf.h"

#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

namespace b2rf {

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
  count_k4_variant(B2RF_K4_SCAN);
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
// Per-warp staging table for the row broadcast: lane k publishes row k's (packed entry, w*y, w*y*y) as two 16-byte words and
// every lane reads row r back with two 128-bit loads.  All lanes of a row slot read the same address, so each load is one
// broadcast wavefront; the five 32-bit shuffles this replaces go through the same shared-memory pipe as the histogram atomics
// and cost five (measured on C3: K4 0.657 s -> 0.616 s per fit).
constexpr int STAGE_BYTES_PER_WARP = 32 * 32;
__host__ __device__ __forceinline__ size_t bucket_stage_offset(int maxCells) { return ((size_t)maxCells * 20 + 15) & ~(size_t)15; }
__device__ __forceinline__ void stage_row(uint4* st, int lane, uint32_t entry, double wy, double wyy) {
  __syncwarp();                                                                // the previous block's readers are done
  st[lane * 2] = make_uint4(entry, 0u, (uint32_t)__double2loint(wy), (uint32_t)__double2hiint(wy));
  st[lane * 2 + 1] = make_uint4((uint32_t)__double2loint(wyy), (uint32_t)__double2hiint(wyy), 0u, 0u);
  __syncwarp();
}
__device__ __forceinline__ void staged_row(const uint4* st, int r, uint32_t& entry, double& wy, double& wyy) {
  const uint4 v0 = st[r * 2], v1 = st[r * 2 + 1];
  entry = v0.x;
  wy = __hiloint2double((int)v0.w, (int)v0.z);
  wyy = __hiloint2double((int)v1.y, (int)v1.x);
}

// Rows [0, n) of one (slot, chunk) task, one warp row-block at a time.  Lane k of a warp owns row k of the current 32-row
// block: it fetches the packed entry and the label once and forms (w, w*y, w*y*y); the values are then broadcast row by row
// while every lane (== feature) gathers its own bin code and updates its own histogram column.
template <bool DIRECT>
__device__ __forceinline__ void bucket_rows32(const BucketLaunch& a, const uint32_t* rl, const double* yl, int n, int warp, int lane, bool active, uint32_t featCol,
                                            double* hb, uint32_t gbase, double* ssum, double* ssq, unsigned* scnt, uint32_t sbase, uint32_t sstep,
                                            uint4* st) {
  constexpr int W = BUCKET_THREADS / 32;
  constexpr uint32_t NONE = 0xffffffffu;
  for (int base = warp * 32; base < n; base += W * 32) {
    const uint32_t mine = (base + lane < n) ? rl[base + lane] : 0u;          // w == 0 marks padding
    const uint32_t myw = mine >> 27;
    const double myy = myw ? (yl ? yl[base + lane] : a.y[mine & 0x7ffffffu]) : 0.0;   // label from the list when it is carried there
    const double mywy = (double)myw * myy;                                     // MLlib: instanceWeight * label * label == (w*y)*y
    const double mywyy = mywy * myy;
    stage_row(st, lane, mine, mywy, mywyy);
    // All 32 bin-code gathers of the block are issued before the first atomic: the loop is bound by the latency of these
    // loads and of the atomics' read-modify-write chains, so the more loads in flight per warp the better (microbenchmark
    // scratch/ubench_private.cu: 4 / 8 / 16 / 32 rows in flight -> 212 / 229 / 239 / 243 G updates/s).
    uint32_t bin[32];
#pragma unroll
    for (int u = 0; u < 32; u++) {
      const uint32_t p = st[u * 2].x;                                          // rows past the end carry w == 0
      bin[u] = ((p >> 27) && active) ? a.bins_rm[(size_t)(p & 0x7ffffffu) * a.ldr + featCol] : NONE;
    }
#pragma unroll
    for (int u = 0; u < 32; u++) {
      if (bin[u] != NONE) {
        uint32_t p; double wy, wyy;
        staged_row(st, u, p, wy, wyy);
        const uint32_t wv = p >> 27;
        if (DIRECT) {
          double* c = hb + (size_t)(gbase + bin[u]) * 4;
          atomicAdd(c, (double)wv); atomicAdd(c + 1, wy); atomicAdd(c + 2, wyy);
        } else {
          const uint32_t cell = sbase + bin[u] * sstep;
          atomicAdd(&scnt[cell], wv); atomicAdd(&ssum[cell], wy); atomicAdd(&ssq[cell], wyy);
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
  uint4* st = reinterpret_cast<uint4*>(smraw + bucket_stage_offset(a.maxCells)) + warp * (STAGE_BYTES_PER_WARP / 16);
  const uint32_t total = a.chunk_prefix[a.nSlots];
  for (uint32_t task = blockIdx.x; task < total; task += gridDim.x) {
    int lo = 0, hi = a.nSlots;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (a.chunk_prefix[mid] <= task) lo = mid; else hi = mid; }
    const int slot = lo;
    const uint32_t chunk = task - a.chunk_prefix[slot];
    uint32_t cnt; const uint32_t* rl; const double* yl = nullptr;
    const uint32_t begin = chunk * (uint32_t)a.CH;
    if (a.dseg != nullptr) {
      const DSeg sg = a.dseg[a.slot_pos[slot]];
      cnt = sg.count;
      rl = (sg.buf ? a.idx1 : a.idx0) + sg.start + begin;
      if (a.y0) yl = (sg.buf ? a.y1 : a.y0) + sg.start + begin;
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
      if (direct) bucket_rows32<true>(a, rl, yl, n, warp, lane, active, e.x, hb, e.y, ssum, ssq, scnt, sbase, sstep, st);
      else bucket_rows32<false>(a, rl, yl, n, warp, lane, active, e.x, hb, e.y, ssum, ssq, scnt, sbase, sstep, st);
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
__device__ __forceinline__ void bucket_rows_impl(const BucketLaunch& a, const uint32_t* rl, const double* yl, int n, int warp, int lane, bool active, uint32_t featCol,
                                                 double* hb, uint32_t gbase, double* ssum, double* ssq, unsigned* scnt, uint32_t sbase, uint32_t sstep,
                                                 int Rrt, int rrt, uint4* st, uint32_t& z_c, double& z_s, double& z_q) {
  const int R = ONE_ROW ? 1 : Rrt, r = ONE_ROW ? 0 : rrt;
  constexpr int W = THREADS / 32;
  constexpr int U = ONE_ROW ? 16 : 8;                                          // row groups in flight (gathers issued before the first atomic)
  constexpr uint32_t NONE = 0xffffffffu;
  for (int base = warp * 32; base < n; base += W * 32) {
    const uint32_t mine = (base + lane < n) ? rl[base + lane] : 0u;          // w == 0 marks padding
    const uint32_t myw = mine >> 27;
    const double myy = myw ? (yl ? yl[base + lane] : a.y[mine & 0x7ffffffu]) : 0.0;   // label from the list when it is carried there
    const double mywy = (double)myw * myy;                                     // MLlib: instanceWeight * label * label == (w*y)*y
    const double mywyy = mywy * myy;
    stage_row(st, lane, mine, mywy, mywyy);
    const int nb = min(32, n - base);
    for (int q = 0; q < nb; q += U * R) {
      uint32_t bin[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        const int ri = q + u * R + r;
        const uint32_t p = ri < 32 ? st[ri * 2].x : 0u;                        // rows past the end of the list carry w == 0
        bin[u] = ((p >> 27) && active) ? load_code(a.bins_rm, (size_t)(p & 0x7ffffffu) * a.ldr + featCol, CB ? CB : a.codeBytes) : NONE;
      }
#pragma unroll
      for (int u = 0; u < U; u++) {
        if (bin[u] != NONE) {
          uint32_t p; double wy, wyy;
          staged_row(st, q + u * R + r, p, wy, wyy);
          const uint32_t wv = p >> 27;
          if (DIRECT) {
            double* c = hb + (size_t)(gbase + bin[u]) * 4;
            atomicAdd(c, (double)wv); atomicAdd(c + 1, wy); atomicAdd(c + 2, wyy);
          } else if (a.zeroReg && bin[u] == 0u) {
            // bin 0 of every feature is kept in registers and added once per task: for categorical features indexed by
            // frequency (StringIndexer) and for zero-heavy counts it is the hottest cell of the column, the one every warp of
            // the CTA would otherwise fight over with compare-and-swap loops
            z_c += wv; z_s += wy; z_q += wyy;
          } else {
            const uint32_t cell = sbase + bin[u] * sstep;
            atomicAdd(&scnt[cell], wv); atomicAdd(&ssum[cell], wy); atomicAdd(&ssq[cell], wyy);
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
  uint4* st = reinterpret_cast<uint4*>(smraw + bucket_stage_offset(a.maxCells)) + warp * (STAGE_BYTES_PER_WARP / 16);
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
    uint32_t cnt; const uint32_t* rl; const double* yl = nullptr;
    const uint32_t begin = chunk * (uint32_t)a.CH;
    if (a.dseg != nullptr) {
      const DSeg sg = a.dseg[a.slot_pos[slot]];
      cnt = sg.count;
      rl = (sg.buf ? a.idx1 : a.idx0) + sg.start + begin;
      if (a.y0) yl = (sg.buf ? a.y1 : a.y0) + sg.start + begin;
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
      uint32_t z_c = 0u; double z_s = 0.0, z_q = 0.0;
      if (direct) bucket_rows_impl<true, THREADS, ONE_ROW, CB>(a, rl, yl, n, warp, lane, active, e.x, hb, e.y, ssum, ssq, scnt, 0u, 0u, 32 / G, r, st, z_c, z_s, z_q);
      else {
        bucket_rows_impl<false, THREADS, ONE_ROW, CB>(a, rl, yl, n, warp, lane, active, e.x, hb, e.y, ssum, ssq, scnt, (uint32_t)fl, (uint32_t)G, 32 / G, r, st, z_c, z_s, z_q);
        if (z_c) { atomicAdd(&scnt[fl], z_c); atomicAdd(&ssum[fl], z_s); atomicAdd(&ssq[fl], z_q); }
      }
    } else {
      for (int j0 = 0; j0 < a.Fsub; j0 += 32) {
        const int j = j0 + lane;
        const bool active = j < a.Fsub;
        const uint2 e = active ? fo[j] : make_uint2(0u, 0u);
        uint32_t z_c = 0u; double z_s = 0.0, z_q = 0.0;
        if (direct) bucket_rows_impl<true, THREADS, true, CB>(a, rl, yl, n, warp, lane, active, e.x, hb, e.y, ssum, ssq, scnt, 0u, 0u, 1, 0, st, z_c, z_s, z_q);
        else {
          bucket_rows_impl<false, THREADS, true, CB>(a, rl, yl, n, warp, lane, active, e.x, hb, e.y, ssum, ssq, scnt, e.y, 1u, 1, 0, st, z_c, z_s, z_q);
          if (z_c) { atomicAdd(&scnt[e.y], z_c); atomicAdd(&ssum[e.y], z_s); atomicAdd(&ssq[e.y], z_q); }
        }
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
    count_k4_variant(B2RF_K4_FP64_32);
    bucket_hist32_kernel<<<grid, BUCKET_THREADS, smem, s>>>(b);
    return cudaGetLastError();
  }
  count_k4_variant(B2RF_K4_FP64_GENERAL);
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

// CTAs per SM and block size from the histogram size: the task histogram (20 bytes per cell) plus the warps' staging tables
// must fit 228 KB per SM together with the 1 KB the runtime reserves per CTA.  Small histograms: 256 threads, up to four CTAs
// (64 registers x 256 threads x 4 fills the register file); big ones: 16 warps per CTA, two CTAs or one.
static cudaError_t launch_bucket_sized(const BucketLaunch& b, cudaStream_t s) {
  const size_t hist = bucket_stage_offset(b.maxCells);
  const size_t budget = (size_t)228 * 1024;
  int perSM = (int)(budget / (hist + (BUCKET_THREADS / 32) * STAGE_BYTES_PER_WARP + 1024));
  if (perSM > 4) perSM = 4;
  int threads = BUCKET_THREADS;
  if (perSM <= 2) {
    threads = 512;
    perSM = (int)(budget / (hist + (512 / 32) * STAGE_BYTES_PER_WARP + 1024));
    if (perSM < 1) perSM = 1;
    if (perSM > 2) perSM = 2;
  }
  const size_t smem = hist + (size_t)(threads / 32) * STAGE_BYTES_PER_WARP;
  return launch_bucket_kernel(b, device_sm_count() * perSM, threads, smem, s);
}

// ---- segment mode: task table from the nodes' row segments, then the same shared-memory histogram kernel ----
__global__ void __launch_bounds__(1024) seg_tasks_kernel(const DSeg* dseg, const int64_t* pos, int n, int CH, uint32_t* chunkPrefix, int ticket,
                                                         unsigned long long* rows_acc) {
  __shared__ uint32_t s_chunks[1024];
  const int tid = threadIdx.x;
  const int per = (n + 1023) / 1024;
  const int b = tid * per, e = min(n, b + per);
  uint32_t c = 0;
  unsigned long long rows = 0;
  for (int i = b; i < e; i++) { const uint32_t k = dseg[pos[i]].count; c += (k + CH - 1) / CH; rows += k; }
  if (rows_acc && rows) atomicAdd(rows_acc, rows);
  s_chunks[tid] = c;
  __syncthreads();
  if (tid == 0) {
    uint32_t ac = 0;
    for (int i = 0; i < 1024; i++) { const uint32_t tc = s_chunks[i]; s_chunks[i] = ac; ac += tc; }
    chunkPrefix[n] = ac;
    if (ticket) chunkPrefix[n + 1] = 0u;
  }
  __syncthreads();
  c = s_chunks[tid];
  for (int i = b; i < e; i++) { chunkPrefix[i] = c; c += (dseg[pos[i]].count + CH - 1) / CH; }
}

cudaError_t launch_seg_tasks(const DSeg* dseg, const int64_t* pos, int n, int CH, uint32_t* chunkPrefix, cudaStream_t s, bool ticket,
                             unsigned long long* rows_acc) {
  seg_tasks_kernel<<<1, 1024, 0, s>>>(dseg, pos, n, CH, chunkPrefix, ticket ? 1 : 0, rows_acc);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(1024) seg_tasks_slots_kernel(const SlotSrc* slots, int n, int CH, uint32_t* chunkPrefix) {
  __shared__ uint32_t s_chunks[1024];
  const int tid = threadIdx.x;
  const int per = (n + 1023) / 1024;
  const int b = tid * per, e = min(n, b + per);
  uint32_t c = 0;
  for (int i = b; i < e; i++) c += (slots[i].seg->count + CH - 1) / CH;
  s_chunks[tid] = c;
  __syncthreads();
  if (tid == 0) {
    uint32_t ac = 0;
    for (int i = 0; i < 1024; i++) { const uint32_t tc = s_chunks[i]; s_chunks[i] = ac; ac += tc; }
    chunkPrefix[n] = ac;
    chunkPrefix[n + 1] = 0u;   // work ticket
  }
  __syncthreads();
  c = s_chunks[tid];
  for (int i = b; i < e; i++) { chunkPrefix[i] = c; c += (slots[i].seg->count + CH - 1) / CH; }
}
cudaError_t launch_seg_tasks_slots(const SlotSrc* slots, int n, int CH, uint32_t* chunkPrefix, cudaStream_t s) {
  seg_tasks_slots_kernel<<<1, 1024, 0, s>>>(slots, n, CH, chunkPrefix);
  return cudaGetLastError();
}

cudaError_t launch_segment_hist(const BucketLaunch& b, cudaStream_t s) {
  cudaError_t e = launch_seg_tasks(b.dseg, b.slot_pos, b.nSlots, b.CH, const_cast<uint32_t*>(b.chunk_prefix), s, false, b.rows_acc);
  if (e != cudaSuccess) return e;
  return launch_bucket_sized(b, s);
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
  return launch_bucket_sized(b, s);
}


}  // namespace b2rf

using namespace b2rf;

extern "C" {

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

}  // extern "C"
