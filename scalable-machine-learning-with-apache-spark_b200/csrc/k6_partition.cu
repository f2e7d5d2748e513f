// b2rf (sm_100a) — K6 row routing in segment mode (LearningNode.predictImpl on binned features): initial per-tree row lists, stable partition of split nodes' row segments, node-table maintenance.
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

// ---- initial row lists: per tree, the in-bag rows in ascending order, packed row | weight << 27 ----
// A block covers COMPACT_ROWS = 2048 consecutive rows of one tree, a thread 8 of them (one 8-byte load of bag weights; the weight rows
// are padded to multiples of 128, so whole words can be read and rows past n_rows are masked).
__device__ __forceinline__ uint64_t load_w8(const uint8_t* w, int64_t ldw, int t, int64_t r0, int64_t n_rows) {
  if (r0 >= n_rows) return 0ull;
  uint64_t v = *reinterpret_cast<const uint64_t*>(w + (size_t)t * ldw + r0);
  const int64_t left = n_rows - r0;
  if (left < 8) v &= (1ull << (8 * left)) - 1ull;
  return v;
}
__device__ __forceinline__ uint32_t nonzero_bytes(uint64_t v) {   // bit k set <=> byte k != 0
  uint64_t m = v | (v >> 4); m |= m >> 2; m |= m >> 1; m &= 0x0101010101010101ull;     // bit 0 of every byte = OR of its bits
  return (uint32_t)((m * 0x0102040810204080ull) >> 56);                                  // gather the eight flag bits
}
__global__ void __launch_bounds__(256) compact_count_kernel(const uint8_t* w, int64_t ldw, int64_t n_rows, uint32_t* blockCnt, int nb) {
  const int t = blockIdx.y, b = blockIdx.x;
  const int64_t r0 = (int64_t)b * COMPACT_ROWS + (int64_t)threadIdx.x * 8;
  uint32_t c = __popc(nonzero_bytes(load_w8(w, ldw, t, r0, n_rows)));
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
                                                             const uint64_t* treeBase, uint32_t* idx, const double* y, double* ylist) {
  const int t = blockIdx.y, b = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ uint32_t sw[8];
  const int64_t r0 = (int64_t)b * COMPACT_ROWS + (int64_t)threadIdx.x * 8;
  const uint64_t wv = load_w8(w, ldw, t, r0, n_rows);
  const uint32_t flags = nonzero_bytes(wv);
  const uint32_t mine = __popc(flags);
  uint32_t incl = mine;                                      // exclusive rank of this thread's first kept row inside the block
  for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += u; }
  if (lane == 31) sw[warp] = incl;
  __syncthreads();
  uint32_t before = incl - mine;
  for (int i = 0; i < warp; i++) before += sw[i];
  const size_t base = treeBase[t] + blockOff[(size_t)t * nb + b] + before;
  uint32_t k = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    if ((flags >> j) & 1u) {
      uint32_t c = (uint32_t)(wv >> (8 * j)) & 0xffu;
      if (c > 31u) c = 31u;   // packed in 5 bits
      idx[base + k] = (uint32_t)(r0 + j) | (c << 27);
      if (ylist) ylist[base + k] = y[r0 + j];
      k++;
    }
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
                                 const double* y, double* ylist, cudaStream_t s) {
  const int nb = (int)((n_rows + COMPACT_ROWS - 1) / COMPACT_ROWS);
  if (nb <= 0 || T <= 0) return cudaSuccess;
  compact_write_kernel<<<dim3((unsigned)nb, (unsigned)T), 256, 0, s>>>(w, ldw, n_rows, blockOff, nb, treeBase, idx, y, ylist);
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
        const uint32_t bin = load_code_sparse(col, src[i] & 0x7ffffffu, p.codeBytes);
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
    const double* ysrc = p.y0 ? (sg.buf ? p.y1 : p.y0) + sg.start + (size_t)c * SEG_CH : nullptr;
    double* ydst = p.y0 ? (sg.buf ? p.y0 : p.y1) + sg.start : nullptr;
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
      const uint32_t at = ((bits >> (i & 31)) & 1u) ? lbase + lrank : rbase + ((uint32_t)i - lrank);
      dst[at] = e;
      if (ysrc) ydst[at] = ysrc[i];
    }
    __syncthreads();
  }
}

// p.split_pos holds 2 * nSplits entries: parents' flat positions, then the left children's flat positions.
cudaError_t launch_partition(const PartLaunch& p, cudaStream_t s) {
  if (p.nSplits <= 0) return cudaSuccess;
  { const cudaError_t e = launch_seg_tasks(p.dseg, p.split_pos, p.nSplits, SEG_CH, p.chunk_prefix, s); if (e != cudaSuccess) return e; }
  const int grid = device_sm_count() * 8;
  part_count_kernel<<<grid, 256, 0, s>>>(p);
  part_scan_kernel<<<(p.nSplits + 127) / 128, 128, 0, s>>>(p);
  part_scatter_kernel<<<grid, 256, 0, s>>>(p);
  return cudaGetLastError();
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


}  // namespace b2rf

using namespace b2rf;

extern "C" {


}  // extern "C"
