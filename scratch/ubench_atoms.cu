// Cost of one shared-memory ATOMS.ADD warp instruction as a function of the lane -> bank pattern (B200).
// Patterns (word address = bank + 32 * row-in-table; the table row changes every iteration and differs between lane groups):
//  0 identity          lane i -> bank i                                   (the G = 32 limb kernel)
//  1 halves, straight  lanes 0-15 -> banks 0-15, lanes 16-31 -> banks 16-31, different table rows   (G = 16, bins of different parity)
//  2 halves, crossed   lanes 0-15 -> banks 16-31, lanes 16-31 -> banks 0-15
//  3 halves, same      both half-warps -> banks 0-15, different table rows  (G = 16, bins of equal parity: 2-way conflict)
//  4 halves, random    each half-warp picks its bank half from a hash      (what the G = 16 kernel sees on average)
//  5 quarters, distinct  lane group q -> banks 8q..8q+7 rotated by the iteration (plane-split mapping for G = 8)
//  6 quarters, random    each 8-lane group picks one of 4 bank quarters from a hash (the G = 8 kernel today)
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scratch/ubench_atoms scratch/ubench_atoms.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t hash32(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

template <int PAT>
__global__ void __launch_bounds__(384) atoms_kernel(int iters, long long* cyc, uint32_t* sink) {
  extern __shared__ uint32_t tab[];             // 64 table rows x 32 banks x (blockDim / 32) warps? no: one shared table of 256 rows
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 256 * 32; i += blockDim.x) tab[i] = 0;
  __syncthreads();
  const int half = lane >> 4, quarter = lane >> 3;
  // eight addresses per lane, computed once (the loop below is nothing but the adds)
  uint32_t* ad[8];
#pragma unroll
  for (int u = 0; u < 8; u++) {
    const uint32_t k = (uint32_t)u;
    uint32_t bank, row;
    if (PAT == 0) { bank = lane; row = hash32(k + warp * 977u) & 255u; }
    else if (PAT <= 4) {
      const uint32_t h = hash32(k * 2u + half + warp * 977u);
      row = h & 255u;
      uint32_t sel;
      if (PAT == 1) sel = half; else if (PAT == 2) sel = half ^ 1u; else if (PAT == 3) sel = 0u; else sel = (h >> 8) & 1u;
      bank = (lane & 15) + 16u * sel;
    } else {
      const uint32_t h = hash32(k * 4u + quarter + warp * 977u);
      row = h & 255u;
      const uint32_t sel = PAT == 5 ? ((quarter + k) & 3u) : ((h >> 8) & 3u);
      bank = (lane & 7) + 8u * sel;
    }
    ad[u] = &tab[row * 32u + bank];
  }
  const long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 8; u++) atomicAdd(ad[u], 1u);
  }
  const long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  if (tab[threadIdx.x] == 0xdeadbeefu) *sink = 1;
}

template <int PAT> static void run(const char* name, int sms) {
  long long* d; uint32_t* s; cudaMalloc(&d, 8 * sms * 2); cudaMalloc(&s, 4);
  const int iters = 20000;
  for (int threads : {32, 384}) {
    for (int ctas : {1, 2}) {
      atoms_kernel<PAT><<<sms * ctas, threads, 256 * 32 * 4>>>(iters, d, s);
      cudaDeviceSynchronize();
      long long h[1024]; cudaMemcpy(h, d, 8 * sms * ctas, cudaMemcpyDeviceToHost);
      long long mx = 0; for (int i = 0; i < sms * ctas; i++) mx = h[i] > mx ? h[i] : mx;
      const double instr = (double)iters * 8 * (threads / 32) * ctas;      // warp instructions per SM
      printf("%-22s %3d threads x %d CTA/SM: %.2f clk per ATOMS warp instruction per SM\n", name, threads, ctas, (double)mx / instr);
    }
  }
}

int main() {
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  run<0>("identity", sms); run<1>("halves straight", sms); run<2>("halves crossed", sms); run<3>("halves same", sms); run<4>("halves random", sms);
  run<5>("quarters distinct", sms); run<6>("quarters random", sms);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
