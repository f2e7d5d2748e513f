// b2rf (sm_100a) — K1 split finding (RandomForest.findSplits / findSplitsForContinuousFeature) and the findSplits row sample.
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
#include <thread>
#include <vector>
#include <algorithm>

namespace b2rf {

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

// The same decision for callers whose rows are still on the host: only the local ids of the kept rows (any order), warp-aggregated append.
__global__ void sample_row_ids_kernel(int64_t n_rows, int64_t row0, uint64_t seedWord, double fraction, int64_t* out, int64_t cap,
                                      unsigned long long* count) {
  const int lane = threadIdx.x & 31;
  for (int64_t base = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) - lane; base < n_rows; base += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = base + lane;
    bool keep = false;
    if (row < n_rows) {
      const uint64_t h = mix64(seedWord + (uint64_t)(row0 + row) * 0x9E3779B97F4A7C15ULL);
      keep = ((double)(h >> 11) * 0x1.0p-53) < fraction;
    }
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (!m) continue;
    unsigned long long pos = 0;
    if (lane == 0) pos = atomicAdd(count, (unsigned long long)__popc(m));
    pos = __shfl_sync(0xffffffffu, pos, 0) + __popc(m & ((1u << lane) - 1u));
    if (keep && (int64_t)pos < cap) out[pos] = row;
  }
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


}  // namespace b2rf

using namespace b2rf;

extern "C" {

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

int b2rf_sample_row_ids(int64_t n_rows, int64_t row0_global, int64_t seed, double fraction, int64_t* d_ids_out, int64_t cap, int64_t* d_count,
                        void* stream) {
  if (n_rows < 0 || !d_ids_out || !d_count || cap < 0) { set_error("b2rf_sample_row_ids: bad argument"); return B2RF_ERR_ARG; }
  if (n_rows == 0) return B2RF_OK;
  const uint64_t seedWord = ((uint64_t)(uint32_t)xorshift_first_int(seed) << 32) | 0x9E3779B9u;
  const int64_t blocks = std::min<int64_t>((n_rows + 255) / 256, (int64_t)device_sm_count() * 16);
  sample_row_ids_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(n_rows, row0_global, seedWord, fraction, d_ids_out, cap,
                                                                               reinterpret_cast<unsigned long long*>(d_count));
  return cuda_ok(cudaGetLastError(), "b2rf_sample_row_ids") ? B2RF_OK : B2RF_ERR_CUDA;
}

int64_t b2rf_sample_indices(int64_t seed, int64_t row0_global, int64_t n_rows, double fraction, int64_t* h_idx_out, int64_t cap) {
  const uint64_t seedWord = ((uint64_t)(uint32_t)xorshift_first_int(seed) << 32) | 0x9E3779B9u;
  // hash of every row index: a few host threads over contiguous row ranges, kept rows concatenated in range order (ascending)
  const int nt = (int)std::max<int64_t>(1, std::min<int64_t>({(int64_t)8, (int64_t)std::thread::hardware_concurrency(), n_rows / 262144 + 1}));
  std::vector<std::vector<int64_t>> part(nt);
  auto work = [&](int t) {
    const int64_t lo = n_rows * t / nt, hi = n_rows * (t + 1) / nt;
    std::vector<int64_t>& out = part[t];
    out.reserve((size_t)((double)(hi - lo) * fraction * 1.2) + 64);
    for (int64_t i = lo; i < hi; i++) {
      const uint64_t h = mix64(seedWord + (uint64_t)(row0_global + i) * 0x9E3779B97F4A7C15ULL);
      if (((double)(h >> 11) * 0x1.0p-53) < fraction) out.push_back(i);
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < nt; t++) pool.emplace_back(work, t);
  work(0);
  for (auto& th : pool) th.join();
  int64_t k = 0;
  for (int t = 0; t < nt; t++)
    for (int64_t i : part[t]) { if (k < cap) h_idx_out[k] = i; k++; }
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

}  // extern "C"
