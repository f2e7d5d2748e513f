// b2rf (sm_100a) — K4, limb form (DTStatsAggregator.update via binSeqOp): the default histogram kernel of segment mode.
//
// Why.  Shared memory has one native read-modify-write wide enough to be cheap: the 32-bit integer add (SASS ATOMS.ADD, one
// wavefront, nothing returned).  An fp64 add is a load + compare-and-swap loop (ATOMS.CAST.SPIN.64), about seven wavefronts and
// a dozen instructions, and the (count, sum w*y, sum w*y*y) update of the fp64 kernel (k4_hist.cu) spent 19 shared-memory
// wavefronts and 51 warp instructions per row.  Here a task (node slot, <= 4096-row chunk, feature group) first takes the
// largest label magnitude of its rows, which fixes one binary scale for the task, and every row's w*y and w*y*y become
// fixed-point integers cut into three 20-bit limbs each.  4096 rows x 20 bits cannot overflow a 32-bit counter, so an update is
// seven independent native adds (count + 3 + 3) with no carries, and the per-task sums are EXACT integers (order-independent);
// they are converted to fp64 once per task and added to the global histogram with RED.ADD.F64, as before.
//   * resolution: |y| <= Ymax_task maps to |q| < 2^54, so labels at the task's top binade keep all 53 mantissa bits and the
//     absolute error of a task sum is below 2^-56 * Ymax_task per row — at or below what sequential fp64 summation commits.
//   * integer-valued labels (and class indicators) give bit-exact sums.
//   * tasks whose labels are not finite, or so large / small that y*y leaves the normal fp64 range (|y| beyond 2^+-480), take
//     the plain fp64 path (global atomics, MLlib's own overflow / underflow behaviour), as do tasks with fewer updates than cells.
// Data path.  Labels travel WITH the per-tree row lists (y list parallel to the packed row | weight list, moved by the K6
// partitions), so a row visit costs one 64 B code-row gather instead of that plus a 32 B label sector; the code rows of the
// next 32-row block are copied global -> shared asynchronously (cp.async, SASS LDGSTS, two stages per warp) while the warp adds
// the current block, and a lane reads its feature's code with one LDS.U8 at an immediate offset: no 64-bit address arithmetic
// and no load-to-use stall in the add loop.  Per row: 3 shared loads + 1 multiply-add + 7 adds.
// Compiled with -fmad=false like the rest of the library.
#include "b2rf_internal.h"
#include "b2rf_device.cuh"
#include "../../include/b2rf.h"

#include <atomic>
#include <cstdlib>
#include <cstring>

namespace b2rf {

constexpr int LIMB_PLANES = 7;             // count, 3 limbs of sum w*y, 3 limbs of sum w*y*y (both on the task's binary grid)
constexpr uint32_t LIMB_MASK = (1u << 20) - 1u;
constexpr int LIMB_EXP_MAX = 480;          // |exponent| of the task scale kept inside the range where 2^(2e-53) is a normal double

__device__ __forceinline__ void cp_async16(uint32_t dst_smem, const void* src, uint32_t nbytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst_smem), "l"(src), "r"(nbytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ double pow2d(int k) { return __hiloint2double((1023 + k) << 20, 0); }   // 2^k, k in the normal range

// One row's fixed-point limbs: v0 = (w, a0, a1, a2), v1 = (b0, b1, b2, packed entry).  The two products are MLlib's own
// (instanceWeight * label, then * label, each rounded to fp64); what is cut into limbs is their exact value on the task's
// binary grid: sY = 2^(53 - e), sP = 2^(53 - 2e) with 2^e <= max |y| < 2^(e+1), so |v| < 2^59 and u < 2^60, and every row in the
// task's top binade converts without any rounding (a one-row node then gets sum = fl(w*y) and sumsq = fl(fl(w*y)*y) exactly as
// MLlib computes them, hence impurity exactly 0).
__device__ __forceinline__ void limb_convert(uint32_t ent, double yv, double sY, double sP, uint4& v0, uint4& v1) {
  const uint32_t w = ent >> 27;
  const double wy = (double)w * yv;
  const double wyy = wy * yv;
  const long long v = __double2ll_rn(wy * sY);
  const unsigned long long u = __double2ull_rn(wyy * sP);
  v0 = make_uint4(w, (uint32_t)v & LIMB_MASK, (uint32_t)(v >> 20) & LIMB_MASK, (uint32_t)(int32_t)(v >> 40));
  v1 = make_uint4((uint32_t)u & LIMB_MASK, (uint32_t)(u >> 20) & LIMB_MASK, (uint32_t)(u >> 40), ent);
}

// G    features per task group = lanes per row (32, 16 or 8); a warp instruction covers 32/G rows
// LDR  bytes per row of the row-major code matrix (0: read a.ldr)
// BR   rows per staged block (32, or 16 when rows are wide and shared memory is short)
template <int G, int LDR, int BR, int THREADS>
__global__ void __launch_bounds__(THREADS, THREADS == 256 ? 2 : 1) limb_hist_kernel(const BucketLaunch a, uint32_t* ticket) {
  extern __shared__ __align__(16) unsigned char smraw[];
  __shared__ uint32_t s_red[THREADS / 32];
  __shared__ uint32_t s_task;
  constexpr int W = THREADS / 32;
  constexpr int R = 32 / G;                  // rows per warp instruction
  constexpr int NU = BR / R;                 // warp instructions per staged block
  constexpr unsigned FULL = 0xffffffffu;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ldr = LDR ? LDR : (int)a.ldr;
  const int cpr = ldr >> 4;                  // 16-byte pieces per code row
  const int padB = a.padB;
  uint32_t* hist = reinterpret_cast<uint32_t*>(smraw);                 // [padB][7][G]
  const size_t histBytes = (size_t)padB * LIMB_PLANES * G * 4;
  const int stageBytes = BR * ldr + BR * 32;
  unsigned char* ring = smraw + histBytes + (size_t)warp * 2 * stageBytes;
  const int fl = lane & (G - 1), r = lane / G;
  const int nG = (a.Fsub + G - 1) / G;
  const uint32_t total = a.chunk_prefix[a.nSlots] * (uint32_t)nG;
  const uint8_t* bins_rm = a.bins_rm;

  for (;;) {
    if (tid == 0) s_task = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t task = s_task;
    if (task >= total) break;
    const uint32_t bt = task / (uint32_t)nG;
    const int g = (int)(task - bt * (uint32_t)nG);
    int lo = 0, hi = a.nSlots;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (a.chunk_prefix[mid] <= bt) lo = mid; else hi = mid; }
    const int slot = lo;
    const uint32_t begin = (bt - a.chunk_prefix[slot]) * (uint32_t)a.CH;
    const DSeg sg = a.dseg[a.slot_pos[slot]];
    const uint32_t* rl = (sg.buf ? a.idx1 : a.idx0) + sg.start + begin;
    const double* yl = (sg.buf ? a.y1 : a.y0) + sg.start + begin;
    const int n = (int)min((uint32_t)a.CH, sg.count - begin);
    const uint2* fo = a.slot_fo + (size_t)slot * a.Fsub;
    double* hb = a.hist + (size_t)a.slot_base[slot] * 4;
    const int j = g * G + fl;
    const bool active = j < a.Fsub;
    const uint2 e = active ? fo[j] : make_uint2(0u, 0u);
    // ---- task scale: largest label magnitude of the task's rows (the high word is monotone in |y|) ----
    uint32_t ex = 0;
    for (int i = tid; i < n; i += THREADS) ex = max(ex, (uint32_t)__double2hiint(yl[i]) & 0x7fffffffu);
    ex = __reduce_max_sync(FULL, ex);
    if (lane == 0) s_red[warp] = ex;
    bool direct = (int64_t)n * a.Fsub < (int64_t)a.slot_cells[slot];   // fewer updates than cells: skip the shared-memory round trip
    if (!direct) {
      uint4* h4 = reinterpret_cast<uint4*>(hist);
      for (int c = tid; c < (int)(histBytes >> 4); c += THREADS) h4[c] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < W; i++) ex = max(ex, s_red[i]);
    int ey = (int)(ex >> 20) - 1023;
    if (ex == 0u) ey = 0;                                             // all labels zero (e.g. a pure class-0 node)
    else if (ey > LIMB_EXP_MAX || ey < -LIMB_EXP_MAX) direct = true;   // non-finite labels, or y*y outside the normal fp64 range: plain fp64 path

    if (direct) {
      for (int base = warp * 32; base < n; base += W * 32) {
        const bool in = base + lane < n;
        const uint32_t ent = in ? rl[base + lane] : 0u;
        const double yv = in ? yl[base + lane] : 0.0;
        const double wy = (double)(ent >> 27) * yv;                    // MLlib: instanceWeight * label * label == (w*y)*y
        const double wyy = wy * yv;
#pragma unroll 4
        for (int u = 0; u < G; u++) {
          const int src = u * R + r;
          const uint32_t e2 = __shfl_sync(FULL, ent, src);
          const double wy2 = __shfl_sync(FULL, wy, src), wyy2 = __shfl_sync(FULL, wyy, src);
          if (active && (e2 >> 27)) {
            const uint32_t code = bins_rm[(size_t)(e2 & 0x7ffffffu) * ldr + e.x];
            double* c = hb + (size_t)(e.y + code) * 4;
            atomicAdd(c, (double)(e2 >> 27)); atomicAdd(c + 1, wy2); atomicAdd(c + 2, wyy2);
          }
        }
      }
      continue;   // the next iteration's __syncthreads (ticket) separates the tasks
    }

    const double sY = pow2d(53 - ey), sP = pow2d(53 - 2 * ey);
    const int nblk = (n + BR - 1) / BR;
    uint32_t ent = 0u; double yv = 0.0;
    auto load = [&](int blk) {
      const int i = blk * BR + lane;
      const bool in = blk < nblk && lane < BR && i < n;
      ent = in ? rl[i] : 0u;
      yv = in ? yl[i] : 0.0;
    };
    auto stage = [&](int s) {
      unsigned char* tile = ring + s * stageBytes;
      uint4* vals = reinterpret_cast<uint4*>(tile + BR * ldr);
      if (lane < BR) {
        uint4 v0, v1;
        limb_convert(ent, yv, sY, sP, v0, v1);
        vals[lane * 2] = v0; vals[lane * 2 + 1] = v1;
      }
      __syncwarp();
      const uint32_t tile_s = smem_u32(tile);
      for (int c = lane; c < BR * cpr; c += 32) {
        const int row = c / cpr, seg = c - row * cpr;
        const uint32_t e2 = reinterpret_cast<const uint32_t*>(vals)[row * 8 + 7];
        const bool valid = (e2 >> 27) != 0u;                           // rows past the end of the chunk: zero-filled code row, zero limbs
        const uint8_t* src = valid ? bins_rm + (size_t)(e2 & 0x7ffffffu) * ldr + seg * 16 : bins_rm;
        cp_async16(tile_s + c * 16, src, valid ? 16u : 0u);
      }
      cp_async_commit();
    };
    int b = warp;
    if (b < nblk) { load(b); stage(0); load(b + W); }
    uint32_t* hcell = hist + fl;
    for (int k = 0; b < nblk; k++, b += W) {
      const int s = k & 1;
      if (b + W < nblk) { stage(s ^ 1); load(b + 2 * W); cp_async_wait<1>(); }
      else cp_async_wait<0>();
      __syncwarp();
      const unsigned char* tl = ring + s * stageBytes + r * ldr + e.x;
      const uint4* vv = reinterpret_cast<const uint4*>(ring + s * stageBytes + BR * ldr) + r * 2;
      if (active) {
        // one row ahead: the next row's code and limbs are loaded before this row's adds are issued (the compiler keeps shared
        // loads behind earlier shared atomics, so the overlap has to be written out)
        uint32_t code = tl[0];
        uint4 v0 = vv[0], v1 = vv[1];
#pragma unroll
        for (int u = 0; u < NU; u++) {
          const uint32_t cc = code; const uint4 w0 = v0, w1 = v1;
          if (u + 1 < NU) { code = tl[(u + 1) * R * ldr]; v0 = vv[(u + 1) * R * 2]; v1 = vv[(u + 1) * R * 2 + 1]; }
          uint32_t* c = hcell + cc * (LIMB_PLANES * G);
          atomicAdd(c, w0.x); atomicAdd(c + G, w0.y); atomicAdd(c + 2 * G, w0.z); atomicAdd(c + 3 * G, w0.w);
          atomicAdd(c + 4 * G, w1.x); atomicAdd(c + 5 * G, w1.y); atomicAdd(c + 6 * G, w1.z);
        }
      }
      __syncwarp();
    }
    __syncthreads();
    // ---- flush: exact limb sums -> fp64, added to the global histogram ----
    const double sA = pow2d(ey - 53), sB = pow2d(2 * ey - 53);
    for (int c = tid; c < padB * G; c += THREADS) {
      const int bin = c / G, f2 = c & (G - 1);
      const uint32_t* cell = hist + (size_t)bin * (LIMB_PLANES * G) + f2;
      const uint32_t cnt = cell[0];
      if (!cnt) continue;                                                // untouched cell (also every cell of an idle lane)
      const double A = ((double)(int32_t)cell[3 * G] * 0x1p40 + (double)cell[2 * G] * 0x1p20) + (double)cell[G];
      const double B = ((double)cell[6 * G] * 0x1p40 + (double)cell[5 * G] * 0x1p20) + (double)cell[4 * G];
      double* gp = hb + ((size_t)fo[g * G + f2].y + (size_t)bin) * 4;
      atomicAdd(gp, (double)cnt); atomicAdd(gp + 1, A * sA); atomicAdd(gp + 2, B * sB);
    }
    // the ticket __syncthreads at the top of the next iteration orders this flush before the next task's zeroing
  }
}

static std::atomic<long long> g_variant_count[B2RF_K4_VARIANTS];
void count_k4_variant(int v) { if (v >= 0 && v < B2RF_K4_VARIANTS) g_variant_count[v].fetch_add(1, std::memory_order_relaxed); }
long long k4_variant_count(int v) { return (v >= 0 && v < B2RF_K4_VARIANTS) ? g_variant_count[v].load(std::memory_order_relaxed) : 0; }

// Shared memory of one CTA: the limb histogram plus two staged blocks per warp.
static size_t limb_smem(int padB, int G, int ldr, int BR, int threads) {
  return (size_t)padB * LIMB_PLANES * G * 4 + (size_t)(threads / 32) * 2 * ((size_t)BR * ldr + (size_t)BR * 32);
}

// Picks (G, block rows, threads) for a launch, or returns false when the limb kernel does not apply (the caller then runs the
// fp64 compare-and-swap kernel of k4_hist.cu): uint8 codes, labels carried in the row lists, 16-byte aligned code rows, and a
// task histogram that leaves room for the staging ring.
bool limb_plan(const BucketLaunch& b, int maxBinsAll, LimbPlan* p) {
  if (b.codeBytes != 1 || b.dseg == nullptr || b.y0 == nullptr || b.y1 == nullptr) return false;
  if ((b.ldr & 15) != 0 || (reinterpret_cast<uintptr_t>(b.bins_rm) & 15) != 0 || b.ldr > 1024) return false;
  const size_t cap2 = (size_t)113 * 1024, cap1 = (size_t)226 * 1024;   // two CTAs per SM / one
  int G = 32;
  while (G > 8 && G / 2 >= b.Fsub) G /= 2;
  for (; G >= 8; G /= 2) {
    if (limb_smem(maxBinsAll, G, (int)b.ldr, 32, 256) <= cap2) { *p = {G, 32, 256, maxBinsAll}; return true; }
    if (limb_smem(maxBinsAll, G, (int)b.ldr, 32, 512) <= cap1) { *p = {G, 32, 512, maxBinsAll}; return true; }
    if (limb_smem(maxBinsAll, G, (int)b.ldr, 16, 512) <= cap1) { *p = {G, 16, 512, maxBinsAll}; return true; }
  }
  return false;
}

template <int G, int LDR, int BR, int THREADS>
static cudaError_t launch_limb_inst(const BucketLaunch& b, uint32_t* ticket, int grid, size_t smem, cudaStream_t s) {
  cudaError_t e = cudaFuncSetAttribute(limb_hist_kernel<G, LDR, BR, THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  limb_hist_kernel<G, LDR, BR, THREADS><<<grid, THREADS, smem, s>>>(b, ticket);
  return cudaGetLastError();
}

template <int G>
static cudaError_t launch_limb_g(const BucketLaunch& b, const LimbPlan& p, uint32_t* ticket, int grid, size_t smem, cudaStream_t s) {
  if (p.threads == 256) {
    if (b.ldr == 64) return launch_limb_inst<G, 64, 32, 256>(b, ticket, grid, smem, s);
    if (b.ldr == 128) return launch_limb_inst<G, 128, 32, 256>(b, ticket, grid, smem, s);
    return launch_limb_inst<G, 0, 32, 256>(b, ticket, grid, smem, s);
  }
  if (p.BR == 32) {
    if (b.ldr == 128) return launch_limb_inst<G, 128, 32, 512>(b, ticket, grid, smem, s);
    return launch_limb_inst<G, 0, 32, 512>(b, ticket, grid, smem, s);
  }
  if (b.ldr == 128) return launch_limb_inst<G, 128, 16, 512>(b, ticket, grid, smem, s);
  return launch_limb_inst<G, 0, 16, 512>(b, ticket, grid, smem, s);
}

// Task table (seg_tasks, which also zeroes the ticket behind the prefix array) + the limb kernel.
cudaError_t launch_limb_hist(const BucketLaunch& b0, const LimbPlan& p, cudaStream_t s) {
  BucketLaunch b = b0;
  b.padB = p.padB; b.G = p.G;
  cudaError_t e = launch_seg_tasks(b.dseg, b.slot_pos, b.nSlots, b.CH, const_cast<uint32_t*>(b.chunk_prefix), s, /*ticket=*/true, b.rows_acc);
  if (e != cudaSuccess) return e;
  uint32_t* ticket = const_cast<uint32_t*>(b.chunk_prefix) + b.nSlots + 1;
  const size_t smem = limb_smem(p.padB, p.G, (int)b.ldr, p.BR, p.threads);
  const int grid = device_sm_count() * (p.threads == 256 ? 2 : 1);
  count_k4_variant(B2RF_K4_LIMB);
  if (p.G == 32) return launch_limb_g<32>(b, p, ticket, grid, smem, s);
  if (p.G == 16) return launch_limb_g<16>(b, p, ticket, grid, smem, s);
  return launch_limb_g<8>(b, p, ticket, grid, smem, s);
}

}  // namespace b2rf

extern "C" {

/* How many times each K4 implementation has been launched by this process (test / evidence hook):
 * 0 = limb kernel (k4_limb.cu), 1 = fp64 kernel for <= 32 features (bucket_hist32_kernel), 2 = general fp64 kernel,
 * 3 = multiclass class-count kernel, 4 = column-scan kernels (hist_kernel<0|1|2>). */
int b2rf_k4_launch_counts(int64_t* h_counts, int32_t n) {
  if (!h_counts || n <= 0) return B2RF_ERR_ARG;
  for (int i = 0; i < n; i++) h_counts[i] = (int64_t)b2rf::k4_variant_count(i);
  return B2RF_OK;
}

}  // extern "C"
