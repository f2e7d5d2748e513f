// b2rf (sm_100a) — K3 bagging (BaggedPoint.convertToBaggedRDD): commons-math3 PoissonDistribution over Well19937c, XORShift Bernoulli.
// Compiled with -fmad=false: MLlib's arithmetic is unfused JVM double arithmetic.
#include "b2rf_internal.h"
#include "b2rf_device.cuh"
#include "../../include/b2rf.h"

#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <mutex>

namespace b2rf {

// =================================================================================================
// K3: bagging.  One thread per emulated Spark partition; generator state in local memory.
// =================================================================================================
constexpr int BAG_THREADS = 64;
// Well19937c with its 624-word state in shared memory, laid out [word][thread] so that a thread's word w lives in bank
// (thread % 32) whatever w is: the per-thread indices diverge (variable draws per sample) without bank conflicts.
struct Well19937cSm {
  uint32_t* v;   // &state[0][threadIdx.x]; word w is v[w * BAG_THREADS]
  int index;
  // The recurrence's loop-carried values stay in registers: c0 = v[index] (the z4 of the previous step) and cTop = the word
  // whose top bit feeds z0.  Of the three words a step writes, two are overwritten before anything reads them, so only z3
  // is stored; the four old words a step needs are fetched one step ahead (p*), off the dependent chain.
  uint32_t c0, cTop, pM1, pM2, pM3, pR2;
  __device__ __forceinline__ uint32_t& at(int w) { return v[w * BAG_THREADS]; }
  __device__ __forceinline__ void prefetch() {
    int i1 = index + 70; if (i1 >= 624) i1 -= 624;
    int i2 = index + 179; if (i2 >= 624) i2 -= 624;
    int i3 = index + 449; if (i3 >= 624) i3 -= 624;
    int r2 = index - 2; if (r2 < 0) r2 += 624;
    pM1 = at(i1); pM2 = at(i2); pM3 = at(i3); pR2 = at(r2);
  }
  __device__ void seed(int64_t sd) {
    uint32_t a = (uint32_t)((uint64_t)sd >> 32), b = (uint32_t)((uint64_t)sd & 0xffffffffULL);
    at(0) = a; at(1) = b;
    for (int i = 2; i < 624; i++) {
      const int64_t l = (int64_t)(int32_t)a;
      const uint32_t c = (uint32_t)((1812433253LL * (l ^ (l >> 30)) + i) & 0xffffffffLL);
      at(i) = c;
      a = b; b = c;
    }
    index = 0;
    c0 = at(0); cTop = at(623);
    prefetch();
  }
  __device__ __forceinline__ uint32_t next(int bits) {
    const uint32_t v0 = c0, vM1 = pM1, vM2 = pM2, vM3 = pM3, vR2 = pR2;
    const uint32_t z0 = (0x80000000u & cTop) ^ (0x7FFFFFFFu & vR2);
    const uint32_t z1 = (v0 ^ (v0 << 25)) ^ (vM1 ^ (vM1 >> 27));
    const uint32_t z2 = (vM2 >> 9) ^ (vM3 ^ (vM3 >> 1));
    const uint32_t z3 = z1 ^ z2;
    uint32_t z4 = z0 ^ (z1 ^ (z1 << 9)) ^ (z2 ^ (z2 << 21)) ^ (z3 ^ (z3 >> 21));
    at(index) = z3;               // v[index-1] = z4 and v[index-2] &= 0x80000000 are dead stores: both words are rewritten
    c0 = z4;                      // (as z3 of the next two steps) before they are read from memory
    cTop = vR2;
    index = index == 0 ? 623 : index - 1;
    prefetch();
    z4 ^= (z4 << 7) & 0xe46e1700u;
    z4 ^= (z4 << 15) & 0x9b868000u;
    return z4 >> (32 - bits);
  }
  __device__ __forceinline__ double nextDouble() {
    const int64_t high = (int64_t)next(26) << 26;
    const int64_t low = (int64_t)next(26);
    return (double)(high | low) * 0x1.0p-52;
  }
};

__device__ inline uint32_t rotl32d(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
__device__ uint32_t murmur3_8bytes(const uint8_t* data, uint32_t seed) {
  uint32_t h = seed;
  for (int i = 0; i < 8; i += 4) {
    uint32_t k = (uint32_t)data[i] | ((uint32_t)data[i + 1] << 8) | ((uint32_t)data[i + 2] << 16) | ((uint32_t)data[i + 3] << 24);
    k *= 0xcc9e2d51u; k = rotl32d(k, 15); k *= 0x1b873593u;
    h ^= k; h = rotl32d(h, 13); h = h * 5 + 0xe6546b64u;
  }
  h ^= 8u;
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
struct XorShiftDev {
  uint64_t s;
  __device__ void seed(int64_t init) {
    uint8_t b[8];
    for (int i = 0; i < 8; i++) b[i] = (uint8_t)((uint64_t)init >> (56 - 8 * i));
    const uint32_t lo = murmur3_8bytes(b, 0x3c074a61u);
    const uint32_t hi = murmur3_8bytes(b, lo);
    s = ((uint64_t)hi << 32) | (uint64_t)lo;
  }
  __device__ __forceinline__ int32_t next(int bits) {
    uint64_t n = s ^ (s << 21);
    n ^= n >> 35;
    n ^= n << 4;
    s = n;
    return (int32_t)(n & ((1ULL << bits) - 1));
  }
  __device__ __forceinline__ double nextDouble() { return (double)(((int64_t)next(26) << 27) + next(27)) * 0x1.0p-53; }
};

// mode 0: Poisson (with replacement); 1: Bernoulli (without replacement)
__global__ void __launch_bounds__(BAG_THREADS) bag_kernel(int64_t seed, const int64_t* part_offsets, int P, int part0, int T, double rate, double expNegRate,
                                                          int mode, uint8_t* w, int64_t ldw, int32_t* node, int64_t ldn) {
  extern __shared__ uint32_t bag_state[];   // [624][BAG_THREADS]
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const int64_t r0 = part_offsets[p], r1 = part_offsets[p + 1];
  if (mode == 0) {
    Well19937cSm rng;
    rng.v = bag_state + threadIdx.x;
    rng.seed(seed + (int64_t)(part0 + p) + 1);
    const double lim = 1000 * rate;
    // One uniform per loop trip for every lane: PoissonDistribution.sample() consumes a data-dependent number of uniforms, and
    // a per-sample inner loop would make each warp pay for its slowest lane on every sample.  Flattened, lanes only diverge
    // on the short "sample finished" branch.
    int64_t i = r0;
    int t = 0;
    long n = 0;
    double r = 1.0;
    while (i < r1) {
      bool finished = true;
      if ((double)n < lim) {
        r *= rng.nextDouble();
        if (r >= expNegRate) { n++; finished = false; }
      }
      if (finished) {
        const int c = n > 255 ? 255 : (int)n;
        w[(size_t)t * ldw + i] = (uint8_t)c;
        if (node != nullptr) node[(size_t)t * ldn + i] = c > 0 ? 0 : -1;
        n = 0; r = 1.0;
        if (++t == T) { t = 0; i++; }
      }
    }
  } else if (mode == 1) {
    XorShiftDev rng;
    rng.seed(seed + (int64_t)(part0 + p) + 1);
    for (int64_t i = r0; i < r1; i++)
      for (int t = 0; t < T; t++) {
        const int c = rng.nextDouble() < rate ? 1 : 0;
        w[(size_t)t * ldw + i] = (uint8_t)c;
        if (node != nullptr) node[(size_t)t * ldn + i] = c > 0 ? 0 : -1;
      }
  }
}
// ---------------------------------------------------------------------------------------------------------------------
// K3, warp-parallel form: ONE WARP per emulated partition instead of one thread.
//
// Well19937c is serial only through one 32-bit word: with x_k the word the step at ring index i-k reads from v[i-k],
//     x_{k+1} = M(x_k) ^ d_k,   M(x) = L(x ^ (x << 25)),  L(t) = (t << 9) ^ (t >>> 21),
//     d_k = L(a_k) ^ z0_k ^ (z2_k << 21) ^ (z2_k >>> 21)        (the z1 and z2 terms of z3 ^ (z3 >>> 21) cancel in z4)
// where a_k, z0_k, z2_k only involve ring words written at least 70 steps earlier.  M is linear over GF(2), so 32 consecutive
// steps are a prefix scan of affine maps: five shuffle stages, stage s applying M^(2^s) through four 256-entry byte tables.
// The Poisson sampler consumes a data-dependent number of uniforms per sample; it is parallelised exactly (same products in
// the same order as PoissonDistribution.sample): every lane computes the sample that WOULD start at its uniform, lane 0
// finishes the sample carried in from the previous window and walks the start positions, marked lanes emit.
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t well_M(uint32_t x) { const uint32_t t = x ^ (x << 25); return (t << 9) ^ (t >> 21); }

struct BagWarpSmem {
  uint32_t v[624];
  double u[32];
  uint32_t len[32];
};

__device__ __forceinline__ uint32_t well_apply(const uint32_t* tab /* [4][256] */, uint32_t y) {
  return tab[y & 255u] ^ tab[256 + ((y >> 8) & 255u)] ^ tab[512 + ((y >> 16) & 255u)] ^ tab[768 + (y >> 24)];
}

// d_k of the step that reads ring index i (see the header comment), and its a_k / z2_k for the ring write-back
__device__ __forceinline__ uint32_t well_d(const uint32_t* v, int i, uint32_t& a, uint32_t& z2) {
  int i1 = i + 70;  if (i1 >= 624) i1 -= 624;
  int i2 = i + 179; if (i2 >= 624) i2 -= 624;
  int i3 = i + 449; if (i3 >= 624) i3 -= 624;
  int r1 = i - 1;   if (r1 < 0) r1 += 624;
  int r2 = i - 2;   if (r2 < 0) r2 += 624;
  const uint32_t vM1 = v[i1], vM2 = v[i2], vM3 = v[i3];
  const uint32_t z0 = (0x80000000u & v[r1]) ^ (0x7FFFFFFFu & v[r2]);
  a = vM1 ^ (vM1 >> 27);
  z2 = (vM2 >> 9) ^ (vM3 ^ (vM3 >> 1));
  const uint32_t la = (a << 9) ^ (a >> 21);
  return la ^ z0 ^ (z2 << 21) ^ (z2 >> 21);
}
__device__ __forceinline__ uint32_t well_temper(uint32_t z4) {
  z4 ^= (z4 << 7) & 0xe46e1700u;
  z4 ^= (z4 << 15) & 0x9b868000u;
  return z4;
}

// 64 generator steps at once: lane l owns steps 2l and 2l+1.  With X_k the word step k reads (X_0 = carried word) and
// x_{k+1} = M X_k ^ d_k, the pair is the affine map X_{2l+2} = M^2 X_{2l} ^ (M d_{2l} ^ d_{2l+1}); one 32-lane scan over the pairs
// (powers M^2, M^4, ..., M^32) gives every X_{2l+2}, and X_{2l+1} = M X_{2l} ^ d_{2l} follows locally.  None of the ring words a step
// reads is written inside the same 64 steps (offsets 70 / 179 / 449 ahead; i-1 and i-2 are still the old revolution).
// Returns the tempered outputs of steps 2l (hi) and 2l+1 (lo) — exactly the two next(26) calls of one nextDouble().
__device__ __forceinline__ void well_batch64(uint32_t* v, int& index, uint32_t& c0, const uint32_t* tabs, int lane, uint32_t& out_a, uint32_t& out_b) {
  int ia = index - 2 * lane; if (ia < 0) ia += 624;
  int ib = ia - 1;           if (ib < 0) ib += 624;
  uint32_t aa, z2a, ab, z2b;
  const uint32_t da = well_d(v, ia, aa, z2a);
  const uint32_t db = well_d(v, ib, ab, z2b);
  uint32_t e = well_M(da) ^ db;
  if (lane == 0) e ^= well_M(well_M(c0));
#pragma unroll
  for (int st = 0; st < 5; st++) {
    const int o = 1 << st;
    const uint32_t y = __shfl_up_sync(0xffffffffu, e, o);
    // (M^2)^(2^st) = M^(2^(st+1)): M^2 and M^4 by shifts and xors, M^8 / M^16 / M^32 by byte tables
    const uint32_t my = st == 0 ? well_M(well_M(y)) : (st == 1 ? well_M(well_M(well_M(well_M(y)))) : well_apply(tabs + (st + 1) * 1024, y));
    if (lane >= o) e ^= my;
  }
  uint32_t xa = __shfl_up_sync(0xffffffffu, e, 1);         // X_{2l}
  if (lane == 0) xa = c0;
  const uint32_t xb = well_M(xa) ^ da;                     // X_{2l+1} == z4 of step 2l
  __syncwarp();
  v[ia] = ((xa ^ (xa << 25)) ^ aa) ^ z2a;                  // z3 of each step; the other two words a step writes are dead
  v[ib] = ((xb ^ (xb << 25)) ^ ab) ^ z2b;
  __syncwarp();
  c0 = __shfl_sync(0xffffffffu, e, 31);
  index -= 64; if (index < 0) index += 624;
  out_a = well_temper(xb);
  out_b = well_temper(e);
}

// One block per SM, blockDim / 32 warps each; warp w of block b takes partitions b + gridDim * (w + warps * i).  The kernel is bound by
// instruction issue, so what matters is that every SM carries the same number of partitions: with the earlier 4-warp blocks handed out by
// the block scheduler an SM ended up with 16 or 20 warps (2442 partitions of C3 over 148 SMs) and the kernel took as long as the 20.
__global__ void __launch_bounds__(1024) bag_warp_kernel(int64_t seed, const int64_t* part_offsets, int P, int part0, int T, double rate,
                                                        double expNegRate, uint8_t* w, int64_t ldw, int32_t* node, int64_t ldn) {
  __shared__ uint32_t tabs[6 * 1024];              // M^(2^s), s = 0..5, as 4 x 256 byte tables
  extern __shared__ __align__(16) unsigned char bag_dyn_smem[];
  BagWarpSmem* sm_all = reinterpret_cast<BagWarpSmem*>(bag_dyn_smem);
  // tables: entry (s, byte, val) = M^(2^s)(val << 8*byte)
  for (int e = threadIdx.x; e < 6 * 1024; e += blockDim.x) {
    const int st = e >> 10, byte = (e >> 8) & 3, val = e & 255;
    uint32_t x = (uint32_t)val << (8 * byte);
    for (int r = 0; r < (1 << st); r++) x = well_M(x);
    tabs[e] = x;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  BagWarpSmem& sm = sm_all[warp];
  const unsigned FULL = 0xffffffffu;
  for (int p = blockIdx.x + gridDim.x * warp; p < P; p += gridDim.x * (blockDim.x >> 5)) {
  // AbstractWell.setSeed: v[0..1] = (hi, lo); v[i] = 1812433253 * (l ^ (l >> 30)) + i with l = (long) v[i - 2]: two serial chains
  {
    const int64_t sd = seed + (int64_t)(part0 + p) + 1;
    if (lane < 2) {
      uint32_t cur = lane == 0 ? (uint32_t)((uint64_t)sd >> 32) : (uint32_t)((uint64_t)sd & 0xffffffffULL);
      sm.v[lane] = cur;
      for (int i = lane + 2; i < 624; i += 2) {
        const int64_t l = (int64_t)(int32_t)cur;
        cur = (uint32_t)((1812433253LL * (l ^ (l >> 30)) + i) & 0xffffffffLL);
        sm.v[i] = cur;
      }
    }
  }
  __syncwarp();
  int index = 0;
  uint32_t c0 = sm.v[0];
  const int64_t r0 = part_offsets[p], r1 = part_offsets[p + 1];
  const int64_t total = (r1 - r0) * (int64_t)T;     // samples of this partition, row-major over trees
  const long limI = (long)ceil(1000 * rate);
  int64_t ord = 0;                                   // samples emitted so far == ordRow * T + ordT (kept split: no 64-bit divisions)
  int64_t ordRow = 0; uint32_t ordT = 0;
  double rc = 1.0; long nc = 0;                      // the sample that is in progress across windows
  const bool wideT = T > 32;                         // then ordT + k < 2 T: one conditional subtraction replaces the division
  auto emit = [&](uint32_t k, long n) {              // sample number ord + k (k <= 32)
    const uint32_t tt = ordT + k;
    uint32_t q, t;
    if (wideT) { q = tt >= (uint32_t)T ? 1u : 0u; t = tt - q * (uint32_t)T; }
    else { q = tt / (uint32_t)T; t = tt - q * (uint32_t)T; }
    const int64_t row = r0 + ordRow + (int64_t)q;
    const int c = n > 255 ? 255 : (int)n;
    w[(size_t)t * ldw + row] = (uint8_t)c;
    if (node != nullptr) node[(size_t)t * ldn + row] = c > 0 ? 0 : -1;
  };
  while (ord < total) {
    // ---- 64 generator steps -> 32 uniforms (BitsStreamGenerator.nextDouble: two 26-bit halves) ----
    {
      uint32_t oa, ob;
      well_batch64(sm.v, index, c0, tabs, lane, oa, ob);
      sm.u[lane] = (double)(((int64_t)(oa >> 6) << 26) | (int64_t)(ob >> 6)) * 0x1.0p-52;
    }
    __syncwarp();
    // ---- every lane: the sample that would START at uniform `lane` (r = 1; r *= u; ends when r < p) ----
    uint32_t mylen = 0; double myr = 1.0;            // mylen == 0: runs past the window (myr = product so far)
    {   // the first four uniforms of the chain are loaded together (most samples end within them); same products, same order
      double uu[4];
#pragma unroll
      for (int j = 0; j < 4; j++) uu[j] = sm.u[min(lane + j, 31)];
#pragma unroll
      for (int j = 0; j < 4; j++)
        if (mylen == 0u && lane + j < 32) { myr *= uu[j]; if (!(myr >= expNegRate)) mylen = (uint32_t)(j + 1); }
      if (mylen == 0u)
        for (int k = lane + 4; k < 32; k++) { myr *= sm.u[k]; if (!(myr >= expNegRate)) { mylen = (uint32_t)(k - lane + 1); break; } }
    }
    sm.len[lane] = mylen;
    __syncwarp();
    // ---- lane 0 finishes the carried-in sample; the start positions after it are found by pointer doubling ----
    int first_done = 0; long first_n = 0; int s0 = 32;
    if (lane == 0) {
      int pos = 0; bool open = true;
      while (pos < 32) {
        if (nc >= limI) { open = false; break; }      // iteration cap of sample(): ends without consuming (practically never)
        rc *= sm.u[pos]; pos++;
        if (rc >= expNegRate) nc++; else { open = false; break; }
      }
      if (!open) { first_done = 1; first_n = nc; rc = 1.0; nc = 0; s0 = pos; }
    }
    first_done = __shfl_sync(FULL, first_done, 0);
    s0 = __shfl_sync(FULL, s0, 0);
    // samples start at s0, s0 + len[s0], ... until a chain runs past the window.  J = position after this lane's chain
    // (>= 32: leaves the window, 64: chain incomplete); S = set of reached starts, doubled five times (up to 31 hops).
    uint32_t J = mylen == 0 ? 64u : (uint32_t)lane + mylen;
    uint32_t S = (first_done && s0 < 32) ? (1u << s0) : 0u;
#pragma unroll
    for (int rnd = 0; rnd < 5; rnd++) {
      const uint32_t contrib = (((S >> lane) & 1u) && J < 32u) ? (1u << J) : 0u;
      S |= __reduce_or_sync(FULL, contrib);
      const uint32_t Jn = __shfl_sync(FULL, J, J & 31u);
      if (J < 32u) J = Jn;
    }
    const uint32_t openMask = __ballot_sync(FULL, ((S >> lane) & 1u) && mylen == 0);   // at most one reached chain is incomplete
    const int carry_pos = openMask ? (__ffs(openMask) - 1) : -1;
    const uint32_t marks = S & ~openMask;
    // ---- emit ----
    if (first_done) {
      if (lane == 0 && ord < total) emit(0u, first_n);
      if ((marks >> lane) & 1u) {
        const uint32_t k = 1u + (uint32_t)__popc(marks & ((1u << lane) - 1u));
        if (ord + k < total) emit(k, (long)mylen - 1);
      }
      const uint32_t done = 1u + (uint32_t)__popc(marks);
      ord += done;
      ordT += done;
      if (wideT) { if (ordT >= (uint32_t)T) { ordT -= (uint32_t)T; ordRow++; } }     // done <= 33 <= T
      else { ordRow += ordT / (uint32_t)T; ordT %= (uint32_t)T; }
      // the sample left open at the window's end (if any) continues in the next window
      const double cr = __shfl_sync(FULL, myr, carry_pos < 0 ? 0 : carry_pos);
      if (carry_pos >= 0) { rc = cr; nc = 32 - carry_pos; } else { rc = 1.0; nc = 0; }
    }
    // lane 0 holds the authoritative in-progress sample in both cases (it alone advanced it when the carried-in sample did
    // not finish inside the window)
    rc = __shfl_sync(FULL, rc, 0);
    nc = (long)__shfl_sync(FULL, (long long)nc, 0);
    __syncwarp();
  }
  __syncwarp();
  }   // next partition of this warp
}

__global__ void bag_ones_kernel(int64_t n, int T, uint8_t* w, int64_t ldw, int32_t* node, int64_t ldn) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int t = 0; t < T; t++) { w[(size_t)t * ldw + i] = 1; if (node != nullptr) node[(size_t)t * ldn + i] = 0; }
}


}  // namespace b2rf

using namespace b2rf;

extern "C" {

int b2rf_bag(int64_t seed, const int64_t* h_part_offsets, int32_t P, int32_t part0, int32_t T, double subsamplingRate, int32_t bootstrap,
             uint8_t* d_w, int64_t ldw, int32_t* d_node, int64_t ldn, void* stream) {
  if (P <= 0 || T <= 0 || !h_part_offsets) { set_error("b2rf_bag: bad argument"); return B2RF_ERR_ARG; }
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t n = h_part_offsets[P];
  if (!bootstrap && T == 1 && subsamplingRate == 1.0) {
    if (n > 0) bag_ones_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(n, T, d_w, ldw, d_node, ldn);
    return cuda_ok(cudaGetLastError(), "b2rf_bag") ? B2RF_OK : B2RF_ERR_CUDA;
  }
  int64_t* d_po = nullptr;
  {   // keep the stream-ordered pool's memory across synchronisations: with the default release threshold (0) the driver returns
      // the pool's backing memory to the OS at the next synchronize and maps it again on the next call
    static std::mutex mu;
    static bool tuned[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> g(mu);
    if (dev >= 0 && dev < 64 && !tuned[dev]) {
      cudaMemPool_t pool;
      if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        uint64_t keep = (uint64_t)64 << 20;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
      }
      tuned[dev] = true;
    }
  }
  if (!cuda_ok(cudaMallocAsync((void**)&d_po, (size_t)(P + 1) * 8, s), "b2rf_bag alloc")) return B2RF_ERR_CUDA;
  if (!cuda_ok(cudaMemcpyAsync(d_po, h_part_offsets, (size_t)(P + 1) * 8, cudaMemcpyHostToDevice, s), "b2rf_bag copy")) return B2RF_ERR_CUDA;
  // exp(-rate) is evaluated once on the host so that every consumer of the stream compares against the same constant
  static const bool serialOnly = [] { const char* m = std::getenv("B2RF_BAG_SERIAL"); return m && *m == '1'; }();
  bool ok = true;
  if (bootstrap && 1000 * subsamplingRate >= 64 && !serialOnly) {   // warp per partition (window of 32 uniforms < iteration cap)
    const int sms = device_sm_count();
    int warps = (P + sms - 1) / sms;                       // partitions per SM, one warp each (more than 32: warps loop)
    warps = warps < 1 ? 1 : (warps > 32 ? 32 : warps);
    const int grid = std::min(sms, (P + warps - 1) / warps);
    const size_t smem = (size_t)warps * sizeof(BagWarpSmem);
    ok = cuda_ok(cudaFuncSetAttribute(bag_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "b2rf_bag attr");
    if (ok) {
      bag_warp_kernel<<<grid, warps * 32, smem, s>>>(seed, d_po, P, part0, T, subsamplingRate, std::exp(-subsamplingRate), d_w, ldw, d_node, ldn);
      ok = cuda_ok(cudaGetLastError(), "b2rf_bag");
    }
  } else {
    const size_t smem = (size_t)624 * BAG_THREADS * 4;
    ok = cuda_ok(cudaFuncSetAttribute(bag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "b2rf_bag attr");
    if (ok) {
      bag_kernel<<<(P + BAG_THREADS - 1) / BAG_THREADS, BAG_THREADS, smem, s>>>(seed, d_po, P, part0, T, subsamplingRate, std::exp(-subsamplingRate),
                                                                               bootstrap ? 0 : 1, d_w, ldw, d_node, ldn);
      ok = cuda_ok(cudaGetLastError(), "b2rf_bag");
    }
  }
  cudaFreeAsync(d_po, s);   // stream-ordered.  No host sync: cudaMemcpyAsync from pageable memory stages h_part_offsets before returning
  return ok ? B2RF_OK : B2RF_ERR_CUDA;
}

}  // extern "C"
