// Small device helpers shared by the kernel files: mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP), bin-code loads.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2rf {

// =================================================================================================
// small PTX helpers: mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP)
// =================================================================================================
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// Bin codes are uint8 when maxBins <= 256 and uint16 otherwise (codeBytes = 1 / 2, uniform per launch).
__device__ __forceinline__ uint32_t load_code(const void* base, size_t idx, int codeBytes) {
  return codeBytes == 2 ? (uint32_t) reinterpret_cast<const uint16_t*>(base)[idx] : (uint32_t) reinterpret_cast<const uint8_t*>(base)[idx];
}
__device__ __forceinline__ bool in_cat_mask(const uint32_t* masks, uint32_t aux, int maskWords, uint32_t bin) {
  return (masks[(size_t)(aux & 0x7fffffffu) * maskWords + (bin >> 5)] >> (bin & 31)) & 1u;
}

// fp64 add to the GLOBAL histogram, no return value (SASS REDG.E.ADD.F64).  Written as PTX because the histogram pointer comes out of a
// per-slot table: through atomicAdd() the compiler cannot prove the address space and emits a generic atomic — an address-space query
// plus a shared-memory compare-and-swap path and a returning global ATOM per call site.
__device__ __forceinline__ void red_add_f64(double* p, double v) { asm volatile("red.global.add.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory"); }

// Gathers from the code matrices.  A plain LDG that misses L2 makes B200 fetch the whole 128-byte line from DRAM (measured,
// scratch/ubench_gather.cu / profiles/r02_ubench_gather.md: a scattered 64-byte row costs 127 B of DRAM traffic, a single byte ~100 B);
// the L2::64B prefetch-size qualifier (SASS LDG.E.LTC64B) halves that: 68 B per 64-byte row.  Read-only data (ld.global.nc).
__device__ __forceinline__ uint4 ldg128_sparse(const void* p) {
  uint4 v;
  asm("ld.global.nc.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t load_code_sparse(const void* base, size_t idx, int codeBytes) {
  uint32_t v;
  if (codeBytes == 2) { uint16_t h; asm("ld.global.nc.L2::64B.u16 %0, [%1];" : "=h"(h) : "l"(reinterpret_cast<const uint16_t*>(base) + idx)); v = h; }
  else asm("ld.global.nc.L2::64B.u8 %0, [%1];" : "=r"(v) : "l"(reinterpret_cast<const uint8_t*>(base) + idx));
  return v;
}

}  // namespace b2rf
