// How many DRAM bytes does a gather of aligned ROW-byte rows cost on B200?  rows = 10M x ROW bytes (640 MB at 64: > L2), the index list
// is ascending with random gaps (like a node's row list) or fully random; 4 (ROW=64) lanes fetch one row with LDG.128.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -o scratch/ubench_gather scratch/ubench_gather.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

template <int HINT> __device__ __forceinline__ uint4 ldg16(const void* p) {
  uint4 v;
  if (HINT == 64) asm volatile("ld.global.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  else if (HINT == 128) asm volatile("ld.global.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  else if (HINT == 256) asm volatile("ld.global.L2::256B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  else if (HINT == 1) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  else if (HINT == 2) asm volatile("ld.global.cs.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  else v = *reinterpret_cast<const uint4*>(p);
  return v;
}

template <int ROW, int HINT>
__global__ void gather_kernel(const uint8_t* __restrict__ m, const uint32_t* __restrict__ idx, int64_t n, unsigned long long* out) {
  constexpr int LPR = ROW / 16;                 // lanes per row
  constexpr int RPW = 32 / LPR;                 // rows per warp instruction
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  unsigned long long acc = 0;
  for (int64_t base = warp * 32; base < n; base += nw * 32) {
    const uint32_t mine = base + lane < n ? idx[base + lane] : 0u;
#pragma unroll 4
    for (int u = 0; u < 32 / RPW; u++) {
      const uint32_t r = __shfl_sync(0xffffffffu, mine, u * RPW + lane / LPR);
      const uint4 v = ldg16<HINT>(m + (size_t)r * ROW + (lane % LPR) * 16);
      acc += v.x ^ v.y ^ v.z ^ v.w;
    }
  }
  if (acc == 0x1234567812345678ull) *out = acc;
}

template <int ROW, int HINT = 0>
static void run(const char* name, const uint8_t* dm, const uint32_t* didx, int64_t n, unsigned long long* dout) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  gather_kernel<ROW, HINT><<<148 * 8, 256>>>(dm, didx, n, dout);
  cudaEventRecord(a);
  gather_kernel<ROW, HINT><<<148 * 8, 256>>>(dm, didx, n, dout);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  printf("%s ROW=%d hint=%d: %.3f ms, %.1f GB/s of rows\n", name, ROW, HINT, ms, (double)n * ROW / ms / 1e6);
}

int main() {
  const int64_t R = 10000000, n = 63200000;    // 10 list passes of 6.32M in-bag rows
  uint8_t* dm; cudaMalloc(&dm, (size_t)R * 128); cudaMemset(dm, 1, (size_t)R * 128);
  std::vector<uint32_t> asc(n), rnd(n);
  uint64_t s = 88172645463325252ull;
  auto nxt = [&] { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
  for (int p = 0; p < 10; p++) {                // each pass: an ascending 63 % sample of the rows
    int64_t k = 0;
    for (int64_t r = 0; r < R && k < n / 10; r++) if (nxt() % 1000 < 640) asc[p * (n / 10) + k++] = (uint32_t)r;
    for (; k < n / 10; k++) asc[p * (n / 10) + k] = (uint32_t)(R - 1);
  }
  for (int64_t i = 0; i < n; i++) rnd[i] = (uint32_t)(nxt() % R);
  // node lists at level 4: 16 interleaved ascending subsequences
  std::vector<uint32_t> node(n);
  for (int p = 0; p < 10; p++) {
    std::vector<std::vector<uint32_t>> b(16);
    for (int64_t i = 0; i < n / 10; i++) { const uint32_t r = asc[p * (n / 10) + i]; b[(r * 2654435761u) >> 28].push_back(r); }
    int64_t k = p * (n / 10);
    for (auto& v : b) for (uint32_t r : v) node[k++] = r;
  }
  uint32_t* didx; cudaMalloc(&didx, n * 4);
  unsigned long long* dout; cudaMalloc(&dout, 8);
  struct { const char* name; std::vector<uint32_t>* v; } sets[] = {{"ascending", &asc}, {"16 node lists", &node}, {"random", &rnd}};
  for (auto& st : sets) {
    cudaMemcpy(didx, st.v->data(), n * 4, cudaMemcpyHostToDevice);
    run<32>(st.name, dm, didx, n, dout);
    run<64>(st.name, dm, didx, n, dout);
    run<128>(st.name, dm, didx, n, dout);
    run<64, 64>(st.name, dm, didx, n, dout);
    run<64, 128>(st.name, dm, didx, n, dout);
    run<64, 256>(st.name, dm, didx, n, dout);
    run<64, 1>(st.name, dm, didx, n, dout);
    run<64, 2>(st.name, dm, didx, n, dout);
    run<32, 64>(st.name, dm, didx, n, dout);
  }
  cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
