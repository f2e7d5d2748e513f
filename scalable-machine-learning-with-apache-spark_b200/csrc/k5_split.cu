// b2rf (sm_100a) — K5 best split (RandomForest.binsToBestSplit, calculateImpurityStats, Variance.calculate).
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
// K5: binsToBestSplit — one warp per node slot
// =================================================================================================
constexpr int SPLIT_WARPS = 4;   // at most; fewer when a warp's shared arrays (58 B per bin) would not fit
// per-warp shared arrays, carved from dynamic shared memory: raw[maxB][3], cum[maxB][3], key[maxB] (f64), mask[maskWords] (u32),
// order[maxB] (u16)
__host__ __device__ inline size_t split_warp_bytes(int maxB, int maskWords) {
  size_t b = (size_t)maxB * (3 + 3 + 1) * 8 + (size_t)maskWords * 4 + (size_t)maxB * 2;
  return (b + 15) & ~(size_t)15;
}

// Variance.calculate / Gini.calculate / Entropy.calculate on one calculator's statistics (cnt = sum w, sum = sum w*y, sq = sum w*y*y).
// Classification (kind 1, 2) is binary: the class counts are (cnt - sum, sum), exact integers, and stats.sum == cnt.
__device__ __forceinline__ double impurity_calc(int kind, double cnt, double sum, double sq) {
  if (cnt == 0) return 0.0;
  if (kind == 0) {
    const double sl = sq - (sum * sum) / cnt;
    return sl / cnt;
  }
  const double c0 = cnt - sum, c1 = sum;
  const double total = c0 + c1;
  if (kind == 1) {                                // impurity = 1 - sum_k (count_k / total)^2, classes in index order
    double imp = 1.0;
    const double f0 = c0 / total; imp -= f0 * f0;
    const double f1 = c1 / total; imp -= f1 * f1;
    return imp;
  }
  double imp = 0.0;                               // impurity = - sum_k freq_k * log2(freq_k), log2(x) = log(x) / log(2), empty classes skipped
  if (c0 != 0) { const double f0 = c0 / total; imp -= f0 * (log(f0) / log(2.0)); }
  if (c1 != 0) { const double f1 = c1 / total; imp -= f1 * (log(f1) / log(2.0)); }
  return imp;
}

struct SplitView {   // one warp's shared arrays
  double (*raw)[3];
  double (*cum)[3];
  double* key;
  uint32_t* mask;
  uint16_t* order;
};

// WPS = warps per node slot.  1: a warp walks the slot's features in subset order.  > 1 (every slot of the launch has its parent's
// statistics, i.e. no root node): the block's warps share ONE slot, warp w takes features w, w + WPS, ... of the subset, and the
// block keeps the candidate with the largest gain, ties to the earliest subset position — the same winner as the sequential
// strict-'>' scan (maxBy keeps the first maximum).  Groups of a few hundred wide slots (C4: 43 features x 256 bins) otherwise leave
// most SMs idle behind one warp's serial prefix sums.
template <int WPS>
__global__ void __launch_bounds__(SPLIT_WARPS * 32) best_split_kernel(const SplitLaunch a) {
  extern __shared__ __align__(16) unsigned char smraw[];
  __shared__ double s_gain[SPLIT_WARPS];
  __shared__ int s_pos[SPLIT_WARPS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int warpsPerBlock = blockDim.x >> 5;
  const int slot = WPS > 1 ? blockIdx.x : blockIdx.x * warpsPerBlock + warp;
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
    Pimp = impurity_calc(a.impurity, P[0], P[1], P[2]);
  }
  bool any = false;
  double bestGain = 0; int bestF = -1, bestS = -1, bestJ = -1;
  double bL[3] = {0, 0, 0}, bR[3] = {0, 0, 0};
  for (int q = lane; q < a.maskWords; q += 32) sm.mask[q] = 0u;
  int firstF = -1;
  const uint2* fo = a.slot_fo + (size_t)slot * a.Fsub;
  const double* hb = a.hist + (size_t)a.slot_base[slot] * 4;

  if (WPS > 1) firstF = (int)fo[0].x;
  for (int j = (WPS > 1 ? warp : 0); j < a.Fsub; j += WPS) {
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
      // ordering key of a category: regression -> its mean label; binary classification -> its class-1 count (categoryStats.stats(1))
      for (int b = lane; b < nb; b += 32) sm.key[b] = sm.raw[b][0] != 0 ? (a.impurity == 0 ? sm.raw[b][1] / sm.raw[b][0] : sm.raw[b][1]) : DBL_MAX;
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
    if (lane < 3) {  // sequential prefix in MLlib's association order (h[next] += h[cur]); lanes 0..2 each run one statistic's chain
      double c = sm.raw[cat ? sm.order[0] : 0][lane];
      sm.cum[0][lane] = c;
      if (cat) for (int q = 1; q < nb; q++) { c = sm.raw[sm.order[q]][lane] + c; sm.cum[q][lane] = c; }
      else for (int q = 1; q < nb; q++) { c = sm.raw[q][lane] + c; sm.cum[q][lane] = c; }
    }
    __syncwarp();
    const double T0 = sm.cum[nb - 1][0], T1 = sm.cum[nb - 1][1], T2 = sm.cum[nb - 1][2];
    if (!haveP) {  // root: parent = left + right of the first evaluated candidate (calculateImpurityStats, stats == null)
      const double l0 = sm.cum[0][0], l1 = sm.cum[0][1], l2 = sm.cum[0][2];
      P[0] = l0 + (T0 - l0); P[1] = l1 + (T1 - l1); P[2] = l2 + (T2 - l2);
      Pimp = impurity_calc(a.impurity, P[0], P[1], P[2]);
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
        const double li = impurity_calc(a.impurity, L0, L1, L2), ri = impurity_calc(a.impurity, R0, R1, R2);
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
  bool writer = true;
  if (WPS > 1) {   // the block's winner: largest gain, then earliest subset position; a warp without candidates never wins
    if (lane == 0) { s_gain[warp] = bestGain; s_pos[warp] = any ? bestJ : -1; }
    __syncthreads();
    int win = -1;
    for (int w = 0; w < WPS; w++) {
      if (s_pos[w] < 0) continue;
      if (win < 0 || s_gain[w] > s_gain[win] || (s_gain[w] == s_gain[win] && s_pos[w] < s_pos[win])) win = w;
    }
    writer = (win < 0) ? (warp == 0) : (warp == win);   // no candidate anywhere: warp 0 reports the dummy split
    if (!writer) return;
  }
  if (lane == 0) {
    SplitRec r;
    memset(&r, 0, sizeof r);
    if (!any) {  // no feature with candidate splits: dummy split, invalid stats on the node totals
      if (!haveP && firstF >= 0) {
        const double* cells = hb + (size_t)fo[0].y * 4;
        const int nb = a.numBins[firstF];
        for (int b = 0; b < nb; b++) { P[0] += cells[(size_t)b * 4]; P[1] += cells[(size_t)b * 4 + 1]; P[2] += cells[(size_t)b * 4 + 2]; }
        Pimp = impurity_calc(a.impurity, P[0], P[1], P[2]);
      }
      r.gain = -DBL_MAX; r.valid = 0; r.feature = firstF < 0 ? 0 : firstF; r.splitIdx = 0;
    } else {
      r.gain = bestGain; r.valid = bestGain != -DBL_MAX; r.feature = bestF; r.splitIdx = bestS;
      for (int q = 0; q < 3; q++) { r.left[q] = bL[q]; r.right[q] = bR[q]; }
      r.left[3] = bL[0]; r.right[3] = bR[0];
      r.leftImp = impurity_calc(a.impurity, bL[0], bL[1], bL[2]);
      r.rightImp = impurity_calc(a.impurity, bR[0], bR[1], bR[2]);
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
  cudaError_t e = cudaFuncSetAttribute(best_split_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  // several warps per slot when no slot is a root (their parent statistics come with the launch) and the per-warp arrays fit
  if (a.allHaveParent && warps == SPLIT_WARPS && a.Fsub >= 2 * SPLIT_WARPS) {
    e = cudaFuncSetAttribute(best_split_kernel<SPLIT_WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    best_split_kernel<SPLIT_WARPS><<<a.nSlots, SPLIT_WARPS * 32, smem, s>>>(a);
    return cudaGetLastError();
  }
  best_split_kernel<1><<<(a.nSlots + warps - 1) / warps, warps * 32, smem, s>>>(a);
  return cudaGetLastError();
}


}  // namespace b2rf

using namespace b2rf;

extern "C" {


}  // extern "C"
