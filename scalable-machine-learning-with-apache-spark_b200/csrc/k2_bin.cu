// b2rf (sm_100a) — K2 binning (TreePoint.findBin): row-major f64 -> column-major + row-major bin codes (uint8 / uint16).
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
        if (x != x) lo = nsplits[f];     // Arrays.binarySearch orders NaN above every double: last bin (and `NaN <= t` is false at predict time)
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
  if (bins == nullptr) return;     // row-major codes only (K7 on bin codes)
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


}  // namespace b2rf

using namespace b2rf;

extern "C" {

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

}  // extern "C"
