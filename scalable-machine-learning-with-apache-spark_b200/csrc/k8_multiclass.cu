// b2rf (sm_100a) — multiclass classification (numClasses > 2): the per-class forms of K4, K5 and K7.
//   K4c  bucket_hist_cls_kernel   per-(node, feature, bin) CLASS COUNTS over the node's row segment (GiniAggregator.update /
//                                 EntropyAggregator.update: allStats(offset + label) += numSamples)
//   K5c  best_split_cls_kernel    binsToBestSplit for the three feature kinds of a multiclass tree: continuous, ordered categorical
//                                 (categories sorted by their impurity) and UNORDERED categorical (every subset split)
//   K7c  predict_raw_mc_kernel    rawPrediction over numClasses
// Everything accumulated here is an integer (bag counts), so sums are exact in u32 / fp64 whatever the order and the Gini arithmetic
// below reproduces the oracle bit for bit (compiled with -fmad=false).
#include "b2rf_internal.h"
#include "b2rf_device.cuh"
#include "../../include/b2rf.h"

#include <cfloat>
#include <cmath>
#include <cstring>

namespace b2rf {

// =================================================================================================
// K4c: class-count histograms.  Same task structure as the regression kernel (one CTA per (node slot, 4096-row chunk,
// group of 32 features), lane == feature, bank == lane), but a cell holds numClasses u32 counters laid out
// [class][bin][lane] and an update is ONE native shared-memory add.  Unordered categorical features are aggregated per category
// like ordered ones; K5c forms the subset sums (MLlib keeps one left-statistics bin per subset split instead: same integers).
// =================================================================================================
constexpr int CLS_THREADS = 256;

__global__ void __launch_bounds__(CLS_THREADS) bucket_hist_cls_kernel(const ClsHistLaunch a) {
  extern __shared__ __align__(16) unsigned char smraw[];
  unsigned* scnt = reinterpret_cast<unsigned*>(smraw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  constexpr int W = CLS_THREADS / 32;
  const unsigned FULL = 0xffffffffu;
  const int nG = (a.Fsub + 31) / 32;
  const uint32_t total = a.chunk_prefix[a.nSlots] * (uint32_t)nG;
  const int plane = a.padB * 32;
  const int scells = a.C * plane;
  const size_t stride = (size_t)4 * a.M;                    // doubles per global histogram cell
  for (uint32_t task = blockIdx.x; task < total; task += gridDim.x) {
    const uint32_t bt = task / (uint32_t)nG;
    const int g = (int)(task % (uint32_t)nG);
    int lo = 0, hi = a.nSlots;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (a.chunk_prefix[mid] <= bt) lo = mid; else hi = mid; }
    const int slot = lo;
    const uint32_t chunk = bt - a.chunk_prefix[slot];
    const DSeg sg = a.dseg[a.slot_pos[slot]];
    const uint32_t begin = chunk * (uint32_t)a.CH;
    const uint32_t* rl = (sg.buf ? a.idx1 : a.idx0) + sg.start + begin;
    const int n = (int)min((uint32_t)a.CH, sg.count - begin);
    const uint2* fo = a.slot_fo + (size_t)slot * a.Fsub;
    double* hb = a.hist + (size_t)a.slot_base[slot] * stride;
    const int j = g * 32 + lane;
    const bool active = j < a.Fsub;
    const uint2 e = active ? fo[j] : make_uint2(0u, 0u);
    const bool direct = a.direct != 0 || (int64_t)n * 32 < (int64_t)scells;   // few rows (or no room): straight to global memory
    if (!direct) {
      for (int c = tid; c < scells; c += CLS_THREADS) scnt[c] = 0u;
      __syncthreads();
    }
    for (int base = warp * 32; base < n; base += W * 32) {
      const uint32_t mine = (base + lane < n) ? rl[base + lane] : 0u;          // w == 0 marks padding
      const int mylab = (mine >> 27) ? (int)a.y[mine & 0x7ffffffu] : 0;
      uint32_t bin[32];
#pragma unroll
      for (int u = 0; u < 32; u++) {                                           // all gathers of the block before the first atomic
        const uint32_t p = __shfl_sync(FULL, mine, u);
        bin[u] = ((p >> 27) && active) ? load_code(a.bins_rm, (size_t)(p & 0x7ffffffu) * a.ldr + e.x, a.codeBytes) : 0xffffffffu;
      }
#pragma unroll
      for (int u = 0; u < 32; u++) {
        const uint32_t p = __shfl_sync(FULL, mine, u);
        const int lab = __shfl_sync(FULL, mylab, u);
        if (bin[u] != 0xffffffffu) {
          if (direct) atomicAdd(hb + (size_t)(e.y + bin[u]) * stride + lab, (double)(p >> 27));
          else atomicAdd(&scnt[(lab * a.padB + (int)bin[u]) * 32 + lane], p >> 27);
        }
      }
    }
    if (!direct) {
      __syncthreads();
      for (int c = tid; c < scells; c += CLS_THREADS) {
        const unsigned k = scnt[c];
        if (!k) continue;
        const int cls = c / plane, rem = c - cls * plane;
        const int jj = g * 32 + (rem & 31);                                     // untouched cells of idle lanes stay 0
        atomicAdd(hb + (size_t)(fo[jj].y + (uint32_t)(rem >> 5)) * stride + cls, (double)k);
      }
      __syncthreads();
    }
  }
}

// ---- staged form (default when labels travel with the row lists and code rows are 64 or 128 bytes of uint8 codes) ----
// The skeleton of the regression limb kernel (k4_limb.cu): dynamic ticket scheduling over (slot, chunk, feature group) tasks, 12
// warps x 3 CTAs per SM (2 when the counters are large; 56 registers per thread), each warp register-stages the code rows of its next 32-row block (LDG.128 -> STS.128, in flight while the
// current block is added) and writes one 8-byte (weight, class offset) pair per row next to them.  A lane then reads its feature's
// code with one LDS.U8 and the row's pair with one broadcast LDS.64 and issues ONE native 32-bit add (ATOMS.ADD): three
// shared-memory wavefronts per 32 updates against the gather kernel's dependent global loads.  Counter layout [bin][class][lane]
// (bank == lane); bag counts are exact in u32 (<= 4096 rows x weight <= 31 per task).
template <int G, int LDR>
__global__ void __launch_bounds__(384, 3) staged_hist_cls_kernel(const ClsHistLaunch a, uint32_t* ticket) {
  extern __shared__ __align__(16) unsigned char smraw[];
  __shared__ uint32_t s_task;
  constexpr int THREADS = 384, W = THREADS / 32, BR = 32;
  constexpr int R = 32 / G, NU = BR / R;
  constexpr int NPC = (BR * LDR) / 512, CPR = LDR / 16;
  constexpr unsigned FULL = 0xffffffffu;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int C = a.C, padB = a.padB;
  uint32_t* hist = reinterpret_cast<uint32_t*>(smraw);                 // [padB][C][G]
  const int cellsPerBin = C * G;
  const size_t histBytes = (size_t)padB * cellsPerBin * 4;
  constexpr int tileBytes = BR * LDR, valBytes = BR * 8, warpBytes = tileBytes + 2 * valBytes;
  unsigned char* ring = smraw + histBytes + (size_t)warp * warpBytes;
  unsigned char* valsBase = ring + tileBytes;
  const int fl = lane & (G - 1), r = lane / G;
  const int nG = (a.Fsub + G - 1) / G;
  const uint32_t total = a.chunk_prefix[a.nSlots] * (uint32_t)nG;
  const uint8_t* bins_rm = reinterpret_cast<const uint8_t*>(a.bins_rm);
  const size_t stride = (size_t)4 * a.M;
  // rows past the end of a chunk carry weight 0; their staged code only has to be a valid bin index (see k4_limb.cu)
  for (int c = lane; c < warpBytes >> 4; c += 32) reinterpret_cast<uint4*>(ring)[c] = make_uint4(0u, 0u, 0u, 0u);
  __syncthreads();

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
    const uint2* fo = a.slot_fo + (size_t)slot * a.Fsub;
    double* hb = a.hist + (size_t)a.slot_base[slot] * stride;
    const int n = (int)min((uint32_t)a.CH, sg.count - begin);
    const int j = g * G + fl;
    const bool active = j < a.Fsub;
    const uint2 e = active ? fo[j] : make_uint2(0u, 0u);
    const bool direct = (int64_t)n * a.Fsub < (int64_t)a.slot_cells[slot] * C;   // fewer updates than counters: straight to global memory
    uint32_t ent = 0u; uint32_t cls = 0u;
    auto load = [&](int blk, int nblk) {
      const int i = blk * BR + lane;
      const bool in = blk < nblk && i < n;
      ent = in ? rl[i] : 0u;
      cls = in ? min((uint32_t)(int)yl[i], (uint32_t)(C - 1)) : 0u;
    };
    if (direct) {
      for (int base = warp * 32; base < n; base += W * 32) {
        const bool in = base + lane < n;
        const uint32_t en = in ? rl[base + lane] : 0u;
        const uint32_t cl = in ? min((uint32_t)(int)yl[base + lane], (uint32_t)(C - 1)) : 0u;
#pragma unroll 4
        for (int u = 0; u < G; u++) {
          const int src = u * R + r;
          const uint32_t e2 = __shfl_sync(FULL, en, src), c2 = __shfl_sync(FULL, cl, src);
          if (active && (e2 >> 27)) {
            const uint32_t code = bins_rm[(size_t)(e2 & 0x7ffffffu) * LDR + e.x];
            red_add_f64(hb + (size_t)(e.y + code) * stride + c2, (double)(e2 >> 27));
          }
        }
      }
      continue;   // the next iteration's __syncthreads (ticket) separates the tasks
    }
    {
      uint4* h4 = reinterpret_cast<uint4*>(hist);
      for (int c = tid; c < (int)(histBytes >> 4); c += THREADS) h4[c] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
    const int nblk = (n + BR - 1) / BR;
    uint32_t* hcell = hist + fl;
    auto put_vals = [&](int s) {
      *reinterpret_cast<uint2*>(valsBase + s * valBytes + lane * 8) = make_uint2(ent >> 27, cls * (uint32_t)G);
    };
    uint4 creg[NPC];
    auto fetch = [&](int blk) {                                    // (ent, cls) hold block blk
      const int nvalid = min(BR, n - blk * BR);
#pragma unroll
      for (int q = 0; q < NPC; q++) {
        const int c = lane + 32 * q, row = c / CPR, seg = c - row * CPR;
        const uint32_t e2 = __shfl_sync(FULL, ent, row);
        creg[q] = row < nvalid ? ldg128_sparse(bins_rm + (size_t)(e2 & 0x7ffffffu) * LDR + seg * 16) : make_uint4(0u, 0u, 0u, 0u);
      }
    };
    int b = warp;
    if (b < nblk) { load(b, nblk); fetch(b); put_vals(0); load(b + W, nblk); }
    for (int k = 0; b < nblk; k++, b += W) {
      const int s = k & 1;
#pragma unroll
      for (int q = 0; q < NPC; q++) reinterpret_cast<uint4*>(ring)[lane + 32 * q] = creg[q];
      __syncwarp();
      if (b + W < nblk) { fetch(b + W); put_vals(s ^ 1); load(b + 2 * W, nblk); }
      if (active) {
        const unsigned char* tl = ring + r * LDR + e.x;
        const unsigned char* vv = valsBase + s * valBytes + r * 8;
        uint32_t code = tl[0];
        uint2 q = *reinterpret_cast<const uint2*>(vv);
#pragma unroll
        for (int u = 0; u < NU; u++) {
          const uint32_t cc = code; const uint2 p = q;
          if (u + 1 < NU) { code = tl[(u + 1) * R * LDR]; q = *reinterpret_cast<const uint2*>(vv + (u + 1) * R * 8); }
          atomicAdd(hcell + cc * (uint32_t)cellsPerBin + p.y, p.x);       // padding rows add 0
        }
      }
      __syncwarp();
    }
    __syncthreads();
    for (int c = tid; c < padB * cellsPerBin; c += THREADS) {
      const uint32_t k = hist[c];
      if (!k) continue;                                                    // untouched counter (also every counter of an idle lane)
      const int bin = c / cellsPerBin, rem = c - bin * cellsPerBin;
      const int cl = rem / G, f2 = rem & (G - 1);
      red_add_f64(hb + ((size_t)fo[g * G + f2].y + (size_t)bin) * stride + cl, (double)k);
    }
    // the ticket __syncthreads at the top of the next iteration orders this flush before the next task's zeroing
  }
}

static size_t staged_cls_smem(int padB, int C, int G, int ldr) { return (size_t)padB * C * G * 4 + (size_t)12 * (32 * ldr + 2 * 32 * 8); }

template <int G, int LDR>
static cudaError_t launch_staged_cls(const ClsHistLaunch& a, uint32_t* ticket, size_t smem, cudaStream_t s) {
  cudaError_t e = cudaFuncSetAttribute(staged_hist_cls_kernel<G, LDR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int perSM = (smem + 1024) * 3 <= (size_t)227 * 1024 ? 3 : 2;   // the kernel is bound by the latency of its gathers: as many warps as fit
  staged_hist_cls_kernel<G, LDR><<<device_sm_count() * perSM, 384, smem, s>>>(a, ticket);
  return cudaGetLastError();
}

cudaError_t launch_cls_hist(const ClsHistLaunch& a0, cudaStream_t s) {
  ClsHistLaunch a = a0;
  // staged form: uint8 codes in 64- or 128-byte rows, labels in the row lists, and a task histogram that leaves two CTAs per SM
  int G = 0;
  if (a.staged && a.codeBytes == 1 && (a.ldr == 64 || a.ldr == 128) && a.y0 && a.y1 && a.slot_cells &&
      (reinterpret_cast<uintptr_t>(a.bins_rm) & 15) == 0) {
    G = 32;
    while (G > 8 && G / 2 >= a.Fsub) G /= 2;
    while (G >= 8 && staged_cls_smem(a.padB, a.C, G, (int)a.ldr) > (size_t)113 * 1024) G /= 2;
    if (G < 8) G = 0;
  }
  cudaError_t e = launch_seg_tasks(a.dseg, a.slot_pos, a.nSlots, a.CH, const_cast<uint32_t*>(a.chunk_prefix), s, /*ticket=*/G != 0, a.rows_acc);
  if (e != cudaSuccess) return e;
  count_k4_variant(B2RF_K4_CLASS_COUNTS);
  if (G) {
    uint32_t* ticket = const_cast<uint32_t*>(a.chunk_prefix) + a.nSlots + 1;
    const size_t smem = staged_cls_smem(a.padB, a.C, G, (int)a.ldr);
    count_k4_variant(B2RF_K4_CLASS_STAGED);
    if (a.ldr == 64) {
      if (G == 32) return launch_staged_cls<32, 64>(a, ticket, smem, s);
      if (G == 16) return launch_staged_cls<16, 64>(a, ticket, smem, s);
      return launch_staged_cls<8, 64>(a, ticket, smem, s);
    }
    if (G == 32) return launch_staged_cls<32, 128>(a, ticket, smem, s);
    if (G == 16) return launch_staged_cls<16, 128>(a, ticket, smem, s);
    return launch_staged_cls<8, 128>(a, ticket, smem, s);
  }
  size_t smem = (size_t)a.C * a.padB * 32 * 4;
  a.direct = smem > (size_t)200 * 1024;
  if (a.direct) smem = 16;
  int perSM = (int)((size_t)224 * 1024 / (smem + 1024));
  if (perSM < 1) perSM = 1;
  if (perSM > 4) perSM = 4;
  e = cudaFuncSetAttribute(bucket_hist_cls_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  bucket_hist_cls_kernel<<<device_sm_count() * perSM, CLS_THREADS, smem, s>>>(a);
  return cudaGetLastError();
}

// =================================================================================================
// K5c: binsToBestSplit over class counts — one warp per node slot
// =================================================================================================
// Gini.calculate / Entropy.calculate over `C` counts at p[0], p[st], p[2 st], ...; total = their sum in class order
__device__ __forceinline__ double cls_impurity(int kind, const double* p, int st, int C, double total) {
  if (total == 0) return 0.0;
  if (kind == 1) {
    double imp = 1.0;
    for (int k = 0; k < C; k++) { const double f = p[(size_t)k * st] / total; imp -= f * f; }
    return imp;
  }
  double imp = 0.0;
  for (int k = 0; k < C; k++) {
    const double c = p[(size_t)k * st];
    if (c != 0) { const double f = c / total; imp -= f * (log(f) / log(2.0)); }
  }
  return imp;
}

__host__ __device__ inline size_t cls_split_warp_bytes(int maxB, int C, int maskWords) {
  size_t b = (size_t)maxB * C * 8 * 2        // raw[maxB][C], cum[maxB][C]
             + (size_t)maxB * 8              // key[maxB]
             + (size_t)C * 8 * 4             // tot, par, bestL, bestR
             + (size_t)maskWords * 4 + (size_t)maxB * 2;
  return (b + 15) & ~(size_t)15;
}

__global__ void __launch_bounds__(128) best_split_cls_kernel(const ClsSplitLaunch a) {
  extern __shared__ __align__(16) unsigned char smraw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slot = blockIdx.x * (blockDim.x >> 5) + warp;
  if (slot >= a.nSlots) return;
  const int C = a.C, maxB = a.maxB;
  unsigned char* base = smraw + (size_t)warp * cls_split_warp_bytes(maxB, C, a.maskWords);
  double* raw = reinterpret_cast<double*>(base);                      // [bin][class]
  double* cum = raw + (size_t)maxB * C;                               // [position in order][class]
  double* key = cum + (size_t)maxB * C;
  double* tot = key + maxB;
  double* par = tot + C;
  double* bestL = par + C;
  double* bestR = bestL + C;
  uint32_t* mask = reinterpret_cast<uint32_t*>(bestR + C);
  uint16_t* order = reinterpret_cast<uint16_t*>(mask + a.maskWords);
  const unsigned FULL = 0xffffffffu;
  const size_t stride = (size_t)4 * a.M;

  bool haveP = a.hasParent[slot] != 0;
  double Pimp = 0, Ptot = 0;
  if (haveP) {
    for (int k = lane; k < C; k += 32) par[k] = a.parentCls[(size_t)slot * a.Cp + k];
    __syncwarp();
    for (int k = 0; k < C; k++) Ptot += par[k];
    Pimp = cls_impurity(a.impurity, par, 1, C, Ptot);
  }
  bool any = false;
  double bestGain = 0; int bestF = -1, bestS = -1;
  for (int q = lane; q < a.maskWords; q += 32) mask[q] = 0u;
  const uint2* fo = a.slot_fo + (size_t)slot * a.Fsub;
  const double* hb = a.hist + (size_t)a.slot_base[slot] * stride;

  for (int j = 0; j < a.Fsub; j++) {
    const uint2 e = fo[j];
    const int f = (int)e.x;
    const int nb = a.cellBins[f];
    const int kind = a.kind[f];                    // 0 continuous, 1 ordered categorical, 2 unordered categorical
    const int ns = kind == 2 ? (1 << (nb - 1)) - 1 : nb - 1;
    if (ns <= 0) continue;
    const double* cells = hb + (size_t)e.y * stride;
    __syncwarp();
    for (int i = lane; i < nb * C; i += 32) raw[i] = cells[(size_t)(i / C) * stride + (i % C)];
    __syncwarp();
    for (int k = lane; k < C; k += 32) { double t = 0; for (int b = 0; b < nb; b++) t += raw[(size_t)b * C + k]; tot[k] = t; }
    __syncwarp();
    double T = 0;
    for (int k = 0; k < C; k++) T += tot[k];
    if (kind == 1) {        // ordered categorical, multiclass: categories sorted by the impurity of their own class counts
      for (int b = lane; b < nb; b += 32) {
        double cnt = 0;
        for (int k = 0; k < C; k++) cnt += raw[(size_t)b * C + k];
        key[b] = cnt != 0 ? cls_impurity(a.impurity, raw + (size_t)b * C, 1, C, cnt) : DBL_MAX;
      }
      __syncwarp();
      for (int b = lane; b < nb; b += 32) {        // stable rank by counting == List.sortBy(_._2)
        const double kb = key[b];
        int rank = 0;
        for (int c = 0; c < nb; c++) { const double kc = key[c]; rank += (kc < kb) || (kc == kb && c < b); }
        order[rank] = (uint16_t)b;
      }
    } else {
      for (int b = lane; b < nb; b += 32) order[b] = (uint16_t)b;
    }
    __syncwarp();
    if (kind != 2) {
      for (int k = lane; k < C; k += 32) {
        double run = 0;
        for (int q = 0; q < nb; q++) { run = raw[(size_t)order[q] * C + k] + run; cum[(size_t)q * C + k] = run; }
      }
      __syncwarp();
    }
    if (!haveP) {           // root: parent = left + right of the first evaluated candidate == the node totals (exact integers)
      for (int k = lane; k < C; k += 32) par[k] = tot[k];
      __syncwarp();
      Ptot = T;
      Pimp = cls_impurity(a.impurity, par, 1, C, Ptot);
      haveP = true;
    }
    double lg = 0; int ls = -1;
    for (int s = lane; s < ns; s += 32) {
      double lc = 0, rc = 0;
      double gain;
      if (kind != 2) {
        for (int k = 0; k < C; k++) { const double L = cum[(size_t)s * C + k]; lc += L; rc += tot[k] - L; }
      } else {
        for (int k = 0; k < C; k++) {
          double L = 0;
          for (int c = 0; c < nb; c++) if (((s + 1) >> c) & 1) L += raw[(size_t)c * C + k];
          lc += L; rc += tot[k] - L;
        }
      }
      if (lc < a.minInstances || rc < a.minInstances || lc < a.minWeight || rc < a.minWeight) gain = -DBL_MAX;
      else {
        // left / right impurities streamed over the classes (same operation order as Gini.calculate / Entropy.calculate)
        double li, ri;
        if (a.impurity == 1) { li = 1.0; ri = 1.0; } else { li = 0.0; ri = 0.0; }
        for (int k = 0; k < C; k++) {
          double L;
          if (kind != 2) L = cum[(size_t)s * C + k];
          else { L = 0; for (int c = 0; c < nb; c++) if (((s + 1) >> c) & 1) L += raw[(size_t)c * C + k]; }
          const double R = tot[k] - L;
          if (a.impurity == 1) {
            const double fl = L / lc; li -= fl * fl;
            const double fr = R / rc; ri -= fr * fr;
          } else {
            if (L != 0) { const double fl = L / lc; li -= fl * (log(fl) / log(2.0)); }
            if (R != 0) { const double fr = R / rc; ri -= fr * (log(fr) / log(2.0)); }
          }
        }
        const double total = lc + rc;
        const double lw = lc / total, rw = rc / total;
        gain = Pimp - lw * li - rw * ri;
        if (gain < a.minInfoGain) gain = -DBL_MAX;
      }
      if (ls < 0 || gain > lg) { lg = gain; ls = s; }
    }
    for (int o = 16; o > 0; o >>= 1) {            // warp argmax: larger gain wins, ties -> lower split index
      const double og = __shfl_xor_sync(FULL, lg, o);
      const int os = __shfl_xor_sync(FULL, ls, o);
      if (os >= 0 && (ls < 0 || og > lg || (og == lg && os < ls))) { lg = og; ls = os; }
    }
    if (!any || lg > bestGain) {                  // strict >: the earliest feature in subset order wins ties
      any = true; bestGain = lg; bestF = f; bestS = ls;
      __syncwarp();
      for (int k = lane; k < C; k += 32) {
        double L;
        if (kind != 2) L = cum[(size_t)ls * C + k];
        else { L = 0; for (int c = 0; c < nb; c++) if (((ls + 1) >> c) & 1) L += raw[(size_t)c * C + k]; }
        bestL[k] = L; bestR[k] = tot[k] - L;
      }
      for (int q = lane; q < a.maskWords; q += 32) mask[q] = 0u;
      __syncwarp();
      if (kind == 1) { for (int q = lane; q <= ls; q += 32) { const int c = order[q]; atomicOr(&mask[c >> 5], 1u << (c & 31)); } }
      else if (kind == 2 && lane == 0) mask[0] = (uint32_t)(ls + 1);      // extractMultiClassCategories(splitIndex + 1, arity): the set bits
      __syncwarp();
    }
  }
  if (!any && !haveP) {       // no feature with candidate splits at a root: statistics = totals of the first feature's cells
    const int f0 = (int)fo[0].x;
    const int nb = a.cellBins[f0];
    const double* cells = hb + (size_t)fo[0].y * stride;
    for (int k = lane; k < C; k += 32) { double t = 0; for (int b = 0; b < nb; b++) t += cells[(size_t)b * stride + k]; par[k] = t; }
    __syncwarp();
    Ptot = 0;
    for (int k = 0; k < C; k++) Ptot += par[k];
    Pimp = cls_impurity(a.impurity, par, 1, C, Ptot);
  }
  __syncwarp();
  double* out3 = a.clsOut + (size_t)slot * 3 * a.Cp;
  for (int k = lane; k < C; k += 32) { out3[k] = par[k]; out3[a.Cp + k] = any ? bestL[k] : 0.0; out3[2 * a.Cp + k] = any ? bestR[k] : 0.0; }
  if (lane == 0) {
    SplitRec r;
    memset(&r, 0, sizeof r);
    if (!any) { r.gain = -DBL_MAX; r.valid = 0; r.feature = (int)fo[0].x; r.splitIdx = 0; }
    else {
      r.gain = bestGain; r.valid = bestGain != -DBL_MAX; r.feature = bestF; r.splitIdx = bestS;
      double lc = 0, rc = 0;
      for (int k = 0; k < C; k++) { lc += bestL[k]; rc += bestR[k]; }
      r.leftImp = cls_impurity(a.impurity, bestL, 1, C, lc);
      r.rightImp = cls_impurity(a.impurity, bestR, 1, C, rc);
      r.left[0] = lc; r.right[0] = rc;
    }
    r.impurity = Pimp;
    r.parent[0] = Ptot;
    a.out[slot] = r;
  }
  __syncwarp();
  for (int q = lane; q < a.maskWords; q += 32) a.catmask_out[(size_t)slot * a.maskWords + q] = any ? mask[q] : 0u;
}

cudaError_t launch_best_split_cls(const ClsSplitLaunch& a, cudaStream_t s) {
  if (a.nSlots <= 0) return cudaSuccess;
  const size_t per = cls_split_warp_bytes(a.maxB, a.C, a.maskWords);
  int warps = (int)(((size_t)200 * 1024) / per);
  if (warps > 4) warps = 4;
  if (warps < 1) return cudaErrorInvalidValue;
  const size_t smem = per * warps;
  cudaError_t e = cudaFuncSetAttribute(best_split_cls_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  best_split_cls_kernel<<<(a.nSlots + warps - 1) / warps, warps * 32, smem, s>>>(a);
  return cudaGetLastError();
}

// =================================================================================================
// K7c: rawPrediction over numClasses (RandomForestClassificationModel.predictRaw / DecisionTreeClassificationModel.predictRaw)
// =================================================================================================
__global__ void predict_raw_mc_kernel(const double* X, int64_t ldx, int64_t n_rows, int T, const int64_t* tree_off, const int32_t* feature,
                                      const double* threshold, const int32_t* left, const int32_t* right, const double* counts, int C,
                                      const int32_t* ncat, const int32_t* cat_off, const int32_t* cat_len, const double* cats, int mode, double* out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows) return;
  const double* x = X + (size_t)i * ldx;
  double* o = out + (size_t)i * C;
  for (int k = 0; k < C; k++) o[k] = 0.0;
  const int nt = mode == 1 ? 1 : T;
  for (int t = 0; t < nt; t++) {
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
    const double* cc = counts + (size_t)(base + n) * C;
    if (mode == 1) { for (int k = 0; k < C; k++) o[k] = cc[k]; }
    else {
      double total = 0;
      for (int k = 0; k < C; k++) total += cc[k];
      if (total != 0) for (int k = 0; k < C; k++) o[k] += cc[k] / total;
    }
  }
}

}  // namespace b2rf

using namespace b2rf;

extern "C" {

int b2rf_predict_raw_mc(const double* d_X, int64_t ldx, int64_t n_rows, int32_t T, const int64_t* d_tree_off, const int32_t* d_feature,
                        const double* d_threshold, const int32_t* d_left, const int32_t* d_right, const double* d_class_counts, int32_t numClasses,
                        const int32_t* d_ncat, const int32_t* d_cat_off, const int32_t* d_cat_len, const double* d_cats, int32_t mode,
                        double* d_out, void* stream) {
  if (mode != 0 && mode != 1) { set_error("b2rf_predict_raw_mc: mode must be 0 (forest votes) or 1 (single-tree class counts)"); return B2RF_ERR_ARG; }
  if (numClasses < 2) { set_error("b2rf_predict_raw_mc: numClasses must be >= 2"); return B2RF_ERR_ARG; }
  if (n_rows <= 0) return B2RF_OK;
  predict_raw_mc_kernel<<<(unsigned)((n_rows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_X, ldx, n_rows, T, d_tree_off, d_feature, d_threshold, d_left,
                                                                                            d_right, d_class_counts, numClasses, d_ncat, d_cat_off,
                                                                                            d_cat_len, d_cats, mode, d_out);
  return cuda_ok(cudaGetLastError(), "b2rf_predict_raw_mc") ? B2RF_OK : B2RF_ERR_CUDA;
}

}  // extern "C"
