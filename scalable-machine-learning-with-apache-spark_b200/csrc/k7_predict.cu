// b2rf (sm_100a) — K7 forest prediction on raw features (TreeEnsembleModel.predict).
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

// ---- packed-node traversal -------------------------------------------------------------------------------------------
// One 16-byte record per node, read with a single 128-bit load per visit (the plain kernel above reads five arrays):
//   leaf:                feat = -1,      v = prediction
//   continuous split:    feat = f >= 0,  v = threshold,                     left child = next record (pre-order), right = `right`
//   categorical split:   feat = -2 - f,  v = bits of (cat_off, cat_len)
// TREES_IN_FLIGHT trees are walked at once per row so several dependent load chains overlap; their leaf values are then
// added in tree order, which keeps MLlib's summation order.
struct __align__(16) PackedNode { double v; int32_t feat; int32_t right; };

__global__ void pack_forest_kernel(int64_t n_nodes, const int32_t* feature, const double* threshold, const int32_t* right, const double* prediction,
                                   const int32_t* ncat, const int32_t* cat_off, const int32_t* cat_len, const int32_t* mask_idx, PackedNode* out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  PackedNode p;
  const int f = feature[i];
  if (f < 0) { p.v = prediction[i]; p.feat = -1; p.right = -1; }
  else if (ncat[i] == 0) { p.v = threshold[i]; p.feat = f; p.right = right[i]; }
  else if (mask_idx != nullptr) { p.v = __hiloint2double(ncat[i], mask_idx[i]); p.feat = -2 - f; p.right = right[i]; }   // (arity, bit-mask index)
  else { p.v = __hiloint2double(cat_len[i], cat_off[i]); p.feat = -2 - f; p.right = right[i]; }
  out[i] = p;
}

constexpr int TREES_IN_FLIGHT = 4;
template <bool LEAVES>
__global__ void __launch_bounds__(256) predict_packed_kernel(const double* __restrict__ X, int64_t ldx, int64_t n_rows, int T, const int64_t* __restrict__ tree_off,
                                                             const PackedNode* __restrict__ nodes, const double* __restrict__ cats, double* out, int32_t* leaf_out,
                                                             const uint32_t* __restrict__ masks, int maskWords) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows) return;
  const double* x = X + (size_t)i * ldx;
  double acc = 0.0;
  for (int t0 = 0; t0 < T; t0 += TREES_IN_FLIGHT) {
    const PackedNode* base[TREES_IN_FLIGHT];
    int n[TREES_IN_FLIGHT];
    PackedNode cur[TREES_IN_FLIGHT];
    bool live[TREES_IN_FLIGHT];
    bool any = false;
#pragma unroll
    for (int k = 0; k < TREES_IN_FLIGHT; k++) {
      live[k] = t0 + k < T;
      base[k] = nodes + (live[k] ? tree_off[t0 + k] : 0);
      n[k] = 0;
      if (live[k]) { cur[k] = base[k][0]; live[k] = cur[k].feat != -1; }
      else { cur[k].v = 0.0; cur[k].feat = -1; cur[k].right = -1; }
      any |= live[k];
    }
    while (any) {
      double xv[TREES_IN_FLIGHT];
#pragma unroll
      for (int k = 0; k < TREES_IN_FLIGHT; k++) {
        const int f = cur[k].feat >= 0 ? cur[k].feat : -2 - cur[k].feat;
        xv[k] = live[k] ? x[f] : 0.0;
      }
      any = false;
#pragma unroll
      for (int k = 0; k < TREES_IN_FLIGHT; k++) {
        if (!live[k]) continue;
        bool l;
        if (cur[k].feat >= 0) l = xv[k] <= cur[k].v;
        else {
          const int co = __double2loint(cur[k].v), cl = __double2hiint(cur[k].v);
          l = false;
          if (masks != nullptr) {      // (bit-mask index, arity): categories.contains(x) as one bit test; non-integral or out-of-range x -> right
            const int xi = (int)xv[k];
            if (xv[k] == (double)xi && xi >= 0 && xi < cl) l = (masks[(size_t)co * maskWords + (xi >> 5)] >> (xi & 31)) & 1u;
          } else {
            for (int c = 0; c < cl; c++) if (cats[co + c] == xv[k]) { l = true; break; }
          }
        }
        n[k] = l ? n[k] + 1 : cur[k].right;
        cur[k] = base[k][n[k]];
        live[k] = cur[k].feat != -1;
        any |= live[k];
      }
    }
#pragma unroll
    for (int k = 0; k < TREES_IN_FLIGHT; k++) {
      if (t0 + k < T) {
        acc += cur[k].v;
        if (LEAVES) leaf_out[(size_t)i * T + t0 + k] = n[k];
      }
    }
  }
  if (out != nullptr) out[i] = acc / (double)T;
}

// ---- traversal on bin codes ---------------------------------------------------------------------------------------------
// For a forest whose splits are all continuous, `x <= threshold` only asks how many of the forest's own thresholds of that feature
// lie strictly below x: with tab[f] the sorted distinct thresholds the forest uses on feature f and code(x) = #{t in tab[f] : t < x}
// (what K2 computes), x <= tab[f][c] <=> code(x) <= c.  Rows shrink from 8 bytes per feature to one code, so a block's rows stay in
// L1 while every tree is walked, and a node is 8 bytes: (feature | split code << 16, right child); leaves are feature 0xffff and take
// their value from the prediction array.  Exactly the same leaves as the f64 traversal, so the same sums.
struct __align__(8) CodedNode { uint32_t fc; int32_t right; };

template <typename CodeT, bool LEAVES>
__global__ void __launch_bounds__(256) predict_coded_kernel(const CodeT* __restrict__ codes, int64_t ldr, int64_t n_rows, int T, const int64_t* __restrict__ tree_off,
                                                            const CodedNode* __restrict__ nodes, const double* __restrict__ prediction, double* out,
                                                            int32_t* leaf_out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows) return;
  const CodeT* x = codes + (size_t)i * ldr;
  double acc = 0.0;
  for (int t0 = 0; t0 < T; t0 += TREES_IN_FLIGHT) {
    int64_t base[TREES_IN_FLIGHT];
    int n[TREES_IN_FLIGHT];
    CodedNode cur[TREES_IN_FLIGHT];
    bool live[TREES_IN_FLIGHT];
    bool any = false;
#pragma unroll
    for (int k = 0; k < TREES_IN_FLIGHT; k++) {
      live[k] = t0 + k < T;
      base[k] = live[k] ? tree_off[t0 + k] : 0;
      n[k] = 0;
      if (live[k]) { cur[k] = nodes[base[k]]; live[k] = (cur[k].fc & 0xffffu) != 0xffffu; }
      else { cur[k].fc = 0xffffu; cur[k].right = -1; }
      any |= live[k];
    }
    while (any) {
      uint32_t c[TREES_IN_FLIGHT];
#pragma unroll
      for (int k = 0; k < TREES_IN_FLIGHT; k++) c[k] = live[k] ? (uint32_t)x[cur[k].fc & 0xffffu] : 0u;
      any = false;
#pragma unroll
      for (int k = 0; k < TREES_IN_FLIGHT; k++) {
        if (!live[k]) continue;
        n[k] = c[k] <= (cur[k].fc >> 16) ? n[k] + 1 : cur[k].right;
        cur[k] = nodes[base[k] + n[k]];
        live[k] = (cur[k].fc & 0xffffu) != 0xffffu;
        any |= live[k];
      }
    }
#pragma unroll
    for (int k = 0; k < TREES_IN_FLIGHT; k++) {
      if (t0 + k < T) {
        acc += prediction[base[k] + n[k]];
        if (LEAVES) leaf_out[(size_t)i * T + t0 + k] = n[k];
      }
    }
  }
  if (out != nullptr) out[i] = acc / (double)T;
}

// Classification rawPrediction (numClasses == 2).  stats4[node] = (sum w, sum w*y, ., .): class counts (c0, c1) = (stats[0] - stats[1], stats[1]).
// mode 0: RandomForestClassificationModel.predictRaw — votes(k) += classCounts(k) / total for every tree, in tree order;
// mode 1: DecisionTreeClassificationModel.predictRaw — the class counts of tree 0's leaf.
__global__ void predict_raw_kernel(const double* X, int64_t ldx, int64_t n_rows, int T, const int64_t* tree_off, const int32_t* feature,
                                   const double* threshold, const int32_t* left, const int32_t* right, const double* stats4,
                                   const int32_t* ncat, const int32_t* cat_off, const int32_t* cat_len, const double* cats, int mode, double* out2) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows) return;
  const double* x = X + (size_t)i * ldx;
  double v0 = 0.0, v1 = 0.0;
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
    const double c1 = stats4[(size_t)(base + n) * 4 + 1];
    const double c0 = stats4[(size_t)(base + n) * 4] - c1;
    if (mode == 1) { v0 = c0; v1 = c1; }
    else {
      const double total = c0 + c1;
      if (total != 0) { v0 += c0 / total; v1 += c1 / total; }
    }
  }
  out2[(size_t)i * 2] = v0; out2[(size_t)i * 2 + 1] = v1;
}


}  // namespace b2rf

using namespace b2rf;

extern "C" {

int b2rf_predict(const double* d_X, int64_t ldx, int64_t n_rows, int32_t T, const int64_t* d_tree_off, const int32_t* d_feature,
                 const double* d_threshold, const int32_t* d_left, const int32_t* d_right, const double* d_prediction, const int32_t* d_ncat,
                 const int32_t* d_cat_off, const int32_t* d_cat_len, const double* d_cats, double* d_out, void* stream) {
  if (n_rows <= 0) return B2RF_OK;
  predict_kernel<<<(unsigned)((n_rows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_X, ldx, n_rows, T, d_tree_off, d_feature, d_threshold, d_left, d_right,
                                                                                     d_prediction, d_ncat, d_cat_off, d_cat_len, d_cats, d_out);
  return cuda_ok(cudaGetLastError(), "b2rf_predict") ? B2RF_OK : B2RF_ERR_CUDA;
}

size_t b2rf_forest_pack_bytes(int64_t n_nodes) { return (size_t)(n_nodes > 0 ? n_nodes : 0) * sizeof(PackedNode); }

int b2rf_forest_pack(int64_t n_nodes, const int32_t* d_feature, const double* d_threshold, const int32_t* d_right, const double* d_prediction,
                     const int32_t* d_ncat, const int32_t* d_cat_off, const int32_t* d_cat_len, const int32_t* d_mask_idx, void* d_packed, void* stream) {
  if (n_nodes <= 0) return B2RF_OK;
  if (!d_packed) { set_error("b2rf_forest_pack: no output buffer"); return B2RF_ERR_ARG; }
  pack_forest_kernel<<<(unsigned)((n_nodes + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n_nodes, d_feature, d_threshold, d_right, d_prediction, d_ncat,
                                                                                         d_cat_off, d_cat_len, d_mask_idx, (PackedNode*)d_packed);
  return cuda_ok(cudaGetLastError(), "b2rf_forest_pack") ? B2RF_OK : B2RF_ERR_CUDA;
}

int b2rf_predict_packed(const double* d_X, int64_t ldx, int64_t n_rows, int32_t T, const int64_t* d_tree_off, const void* d_packed,
                        const double* d_cats, const uint32_t* d_masks, int32_t maskWords, double* d_out, int32_t* d_leaf, void* stream) {
  if (n_rows <= 0) return B2RF_OK;
  if (!d_out && !d_leaf) { set_error("b2rf_predict_packed: neither d_out nor d_leaf given"); return B2RF_ERR_ARG; }
  if (d_masks && maskWords <= 0) { set_error("b2rf_predict_packed: maskWords must be positive with d_masks"); return B2RF_ERR_ARG; }
  const unsigned grid = (unsigned)((n_rows + 255) / 256);
  if (d_leaf) predict_packed_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(d_X, ldx, n_rows, T, d_tree_off, (const PackedNode*)d_packed, d_cats, d_out, d_leaf, d_masks, maskWords);
  else predict_packed_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(d_X, ldx, n_rows, T, d_tree_off, (const PackedNode*)d_packed, d_cats, d_out, d_leaf, d_masks, maskWords);
  return cuda_ok(cudaGetLastError(), "b2rf_predict_packed") ? B2RF_OK : B2RF_ERR_CUDA;
}

int b2rf_predict_coded(const void* d_codes, int64_t ldr, int32_t code_bytes, int64_t n_rows, int32_t T, const int64_t* d_tree_off, const void* d_nodes8,
                       const double* d_prediction, double* d_out, int32_t* d_leaf, void* stream) {
  if (code_bytes != 1 && code_bytes != 2) { set_error("b2rf_predict_coded: code_bytes must be 1 or 2"); return B2RF_ERR_ARG; }
  if (!d_out && !d_leaf) { set_error("b2rf_predict_coded: neither d_out nor d_leaf given"); return B2RF_ERR_ARG; }
  if (n_rows <= 0) return B2RF_OK;
  const unsigned grid = (unsigned)((n_rows + 255) / 256);
  cudaStream_t s = (cudaStream_t)stream;
  const CodedNode* nd = (const CodedNode*)d_nodes8;
  if (code_bytes == 1) {
    if (d_leaf) predict_coded_kernel<uint8_t, true><<<grid, 256, 0, s>>>((const uint8_t*)d_codes, ldr, n_rows, T, d_tree_off, nd, d_prediction, d_out, d_leaf);
    else predict_coded_kernel<uint8_t, false><<<grid, 256, 0, s>>>((const uint8_t*)d_codes, ldr, n_rows, T, d_tree_off, nd, d_prediction, d_out, d_leaf);
  } else {
    if (d_leaf) predict_coded_kernel<uint16_t, true><<<grid, 256, 0, s>>>((const uint16_t*)d_codes, ldr, n_rows, T, d_tree_off, nd, d_prediction, d_out, d_leaf);
    else predict_coded_kernel<uint16_t, false><<<grid, 256, 0, s>>>((const uint16_t*)d_codes, ldr, n_rows, T, d_tree_off, nd, d_prediction, d_out, d_leaf);
  }
  return cuda_ok(cudaGetLastError(), "b2rf_predict_coded") ? B2RF_OK : B2RF_ERR_CUDA;
}

int b2rf_predict_raw(const double* d_X, int64_t ldx, int64_t n_rows, int32_t T, const int64_t* d_tree_off, const int32_t* d_feature,
                     const double* d_threshold, const int32_t* d_left, const int32_t* d_right, const double* d_stats4, const int32_t* d_ncat,
                     const int32_t* d_cat_off, const int32_t* d_cat_len, const double* d_cats, int32_t mode, double* d_out2, void* stream) {
  if (mode != 0 && mode != 1) { set_error("b2rf_predict_raw: mode must be 0 (forest votes) or 1 (single-tree class counts)"); return B2RF_ERR_ARG; }
  if (n_rows <= 0) return B2RF_OK;
  predict_raw_kernel<<<(unsigned)((n_rows + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_X, ldx, n_rows, T, d_tree_off, d_feature, d_threshold, d_left,
                                                                                         d_right, d_stats4, d_ncat, d_cat_off, d_cat_len, d_cats, mode, d_out2);
  return cuda_ok(cudaGetLastError(), "b2rf_predict_raw") ? B2RF_OK : B2RF_ERR_CUDA;
}

}  // extern "C"
