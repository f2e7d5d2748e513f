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

cudaError_t launch_cls_hist(const ClsHistLaunch& a0, cudaStream_t s) {
  cudaError_t e = launch_seg_tasks(a0.dseg, a0.slot_pos, a0.nSlots, a0.CH, const_cast<uint32_t*>(a0.chunk_prefix), s, false, a0.rows_acc);
  if (e != cudaSuccess) return e;
  ClsHistLaunch a = a0;
  size_t smem = (size_t)a.C * a.padB * 32 * 4;
  a.direct = smem > (size_t)200 * 1024;
  if (a.direct) smem = 16;
  int perSM = (int)((size_t)224 * 1024 / (smem + 1024));
  if (perSM < 1) perSM = 1;
  if (perSM > 4) perSM = 4;
  e = cudaFuncSetAttribute(bucket_hist_cls_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  count_k4_variant(B2RF_K4_CLASS_COUNTS);
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
