// Time-boxed go / no-go: the K4 histogram as an exact integer contraction on the tensor cores.
//   hist[bin][limb] = sum_r onehot[bin][r] * limb[r]      tcgen05.mma kind::i8, M = 128 (two features x 64 bins), N = 32 byte-limb
//   columns (count + 2 x 80-bit windows = 21 used), K = 32 rows per instruction, u8 operands from shared memory, s32 accumulators in TMEM.
// One CTA per SM.  Per 32-row block a node with 22 features per node needs 11 MMAs (one per feature pair, 4 KB one-hot tile each) and
// 2 x 704 byte stores (set the block's one-hot bytes, clear them again before the tile is reused).  This harness measures the UPPER BOUND
// of that scheme: the MMAs and the byte stores run concurrently but UNSYNCHRONISED (no full / empty barriers, no code-row staging, no
// limb tile construction, no epilogue), so a real kernel can only be slower.
//   mode 0: MMAs only          mode 1: MMAs + one-hot byte stores by 4 / 8 producer warps          mode 2: byte stores only
// Result check (mode 0): D[bin][n] must equal iters * sum_{r : code(r) == bin} B[n][r]  — proves the descriptors / layouts are the real thing.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -o scratch/ubench_tc_hist scratch/ubench_tc_hist.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr int PAIRS = 11;                 // feature pairs per row block (22 features per node)
constexpr int TILE_A = 4096;              // 128 bins-rows x 32 K-bytes
constexpr int TILE_B = 1024;              // 32 limb columns x 32 K-bytes
constexpr int STAGES = 2;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// canonical K-major no-swizzle layout: 8-row x 16-byte core matrices; K chunk stride (LBO) 128 B, 8-row group stride (SBO) 256 B
__device__ __forceinline__ uint32_t a_off(uint32_t row, uint32_t k) { return (row >> 3) * 256u + (k >> 4) * 128u + (row & 7u) * 16u + (k & 15u); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);          // start address
  d |= (uint64_t)(128u >> 4) << 16;                  // leading byte offset (between the two 16-byte K chunks)
  d |= (uint64_t)(256u >> 4) << 32;                  // stride byte offset (between 8-row groups)
  d |= (uint64_t)1 << 46;                            // descriptor version (Blackwell)
  return d;                                          // layout type 0 = no swizzle
}
__device__ __forceinline__ uint32_t code_of(int pair, int r, int it) { return (uint32_t)((pair * 37 + r * 11 + (r >> 2) * 5 + it * 3) & 127); }

__global__ void __launch_bounds__(288) tc_hist_kernel(int iters, int mode, int producerWarps, long long* cycles, int32_t* dout) {
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  uint8_t* A = sm;                                   // [STAGES][PAIRS][4 KB]
  uint8_t* B = sm + STAGES * PAIRS * TILE_A;         // [1 KB]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int mmaWarp = 8;
  for (int i = tid; i < (STAGES * PAIRS * TILE_A + TILE_B) / 16; i += blockDim.x) reinterpret_cast<uint4*>(sm)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  // stage 0: a real one-hot block (row r of pair p has code code_of(p, r, 0)); B[n][r] = (n * 7 + r * 3) & 0xff
  for (int i = tid; i < PAIRS * 32; i += blockDim.x) { const int p = i >> 5, r = i & 31; A[p * TILE_A + a_off(code_of(p, r, 0), r)] = 1; }
  for (int i = tid; i < 32 * 32; i += blockDim.x) { const int n = i >> 5, r = i & 31; B[a_off(n, r)] = (uint8_t)((n * 7 + r * 3) & 0xff); }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == mmaWarp) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes of the tiles -> visible to the MMA's async proxy
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  // instruction descriptor: D = s32, A = B = u8, both K-major, N = 32, M = 128
  const uint32_t idesc = (2u << 4) | (0u << 7) | (0u << 10) | ((32u >> 3) << 17) | ((128u >> 4) << 24);
  const long long t0 = clock64();
  if (warp == mmaWarp && mode != 2) {
    if (lane == 0) {
      const uint64_t bdesc = make_desc(smem_u32(B));
      for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int p = 0; p < PAIRS; p++) {
          const uint64_t adesc = make_desc(smem_u32(A + p * TILE_A));          // stage 0 tiles (the producers scribble on stage 1)
          const uint32_t acc = it > 0 ? 1u : 0u;
          asm volatile(
              "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
              "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n" ::"r"(tmem + (uint32_t)p * 32u),
              "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc), "r"(0u)
              : "memory");
        }
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    __syncwarp();
    uint32_t ok = 0;
    while (!ok) {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(ok) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    }
  } else if (warp < producerWarps && mode != 0) {
    // one-hot byte traffic on stage 1: per iteration (32-row block) 704 sets + 704 clears, lane = row, warp w takes features w, w + PW, ...
    uint8_t* A1 = A + PAIRS * TILE_A;
    for (int it = 0; it < iters; it++) {
      for (int f = warp; f < 2 * PAIRS; f += producerWarps) {
        const uint32_t codeNew = (code_of(f >> 1, lane, it + 1) & 63u) + 64u * (f & 1), codeOld = (code_of(f >> 1, lane, it) & 63u) + 64u * (f & 1);
        uint8_t* t = A1 + (f >> 1) * TILE_A;
        t[a_off(codeOld, lane)] = 0;
        t[a_off(codeNew, lane)] = 1;
      }
    }
  }
  const long long t1 = clock64();
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (tid == 0) cycles[blockIdx.x] = t1 - t0;       // thread 0 is a producer (or idle) — the MMA warp's own time:
  if (warp == mmaWarp && lane == 0) cycles[gridDim.x + blockIdx.x] = t1 - t0;
  // epilogue (correctness only): warps 0..3 read their 32 TMEM lanes of pair 0's accumulator (32 columns)
  if (warp < 4 && blockIdx.x == 0 && mode == 0) {
    uint32_t v[32];
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]),
          "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]),
          "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int j = 0; j < 32; j++) dout[(warp * 32 + lane) * 32 + j] = (int32_t)v[j];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == mmaWarp) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u));
}

int main(int argc, char** argv) {
  const int iters = argc > 1 ? atoi(argv[1]) : 20000;
  int dev = 0, sms = 0, khz = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  const size_t smem = 120 * 1024;   // tiles need 91 KB; 120 KB keeps one CTA per SM (each CTA allocates all 512 TMEM columns)
  cudaFuncSetAttribute(tc_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  long long* dcyc; int32_t* dout;
  cudaMalloc(&dcyc, sizeof(long long) * 2 * sms); cudaMalloc(&dout, 128 * 32 * 4);
  std::vector<long long> cyc(2 * sms); std::vector<int32_t> out(128 * 32);
  struct { int mode, pw; const char* name; } runs[] = {{0, 0, "MMAs only"}, {2, 4, "byte stores only, 4 producer warps"}, {2, 8, "byte stores only, 8 producer warps"},
                                                       {1, 4, "MMAs + byte stores, 4 producer warps"}, {1, 8, "MMAs + byte stores, 8 producer warps"}};
  for (auto& r : runs) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    tc_hist_kernel<<<sms, 288, smem>>>(iters / 10 + 1, r.mode, r.pw, dcyc, dout);    // warm-up
    cudaMemset(dout, 0, 128 * 32 * 4);
    cudaEventRecord(a);
    tc_hist_kernel<<<sms, 288, smem>>>(iters, r.mode, r.pw, dcyc, dout);
    cudaEventRecord(b);
    cudaError_t e = cudaEventSynchronize(b);
    if (e != cudaSuccess) { printf("%s: %s\n", r.name, cudaGetErrorString(e)); return 1; }
    float ms; cudaEventElapsedTime(&ms, a, b);
    cudaMemcpy(cyc.data(), dcyc, sizeof(long long) * 2 * sms, cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < 2 * sms; i++) mx = cyc[i] > mx ? cyc[i] : mx;
    const double updates = (double)iters * 704.0 * sms;   // (row, feature) updates represented by the work
    printf("%-40s: %.3f ms, %.1f clk per 32-row block per SM (%.2f clk per row, %.2f clk per MMA), %.0f G updates/s\n", r.name, ms, (double)mx / iters,
           (double)mx / iters / 32.0, (double)mx / iters / PAIRS, updates / ms / 1e6);
    if (r.mode == 0) {
      cudaMemcpy(out.data(), dout, 128 * 32 * 4, cudaMemcpyDeviceToHost);
      int bad = 0;
      for (int bin = 0; bin < 128; bin++)
        for (int n = 0; n < 32; n++) {
          long long want = 0;
          for (int rr = 0; rr < 32; rr++) if ((unsigned)((0 * 37 + rr * 11 + (rr >> 2) * 5) & 127) == (unsigned)bin) want += (n * 7 + rr * 3) & 0xff;
          want *= iters;
          if ((long long)out[bin * 32 + n] != (int32_t)want) { if (bad < 5) printf("  mismatch bin %d col %d: got %d want %lld\n", bin, n, out[bin * 32 + n], want); bad++; }
        }
      printf("  accumulator check (pair 0, 128 x 32): %s\n", bad ? "FAILED" : "exact");
    }
  }
  printf("clock rate attr %.0f MHz; production limb kernel: 18.5 clk per row per SM at 22 features (40.3 ms per 6.32e8-visit level), 347 G updates/s\n", khz / 1e3);
  return 0;
}
