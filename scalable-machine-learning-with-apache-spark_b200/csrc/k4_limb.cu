// b2rf (sm_100a) — K4, limb form (DTStatsAggregator.update via binSeqOp): the default histogram kernel of segment mode.
//
// Why.  Shared memory has one native read-modify-write that is cheap: the 32-bit integer add (SASS ATOMS.ADD, one wavefront,
// nothing returned).  An fp64 add is a load + compare-and-swap loop (ATOMS.CAST.SPIN.64): the (count, sum w*y, sum w*y*y) update of
// the fp64 kernels (k4_hist.cu) costs 19 shared-memory wavefronts and 51 warp instructions per row.  Here every row's two
// products — MLlib's own fl(w*y) and fl(fl(w*y)*y) — are placed on a binary grid fixed per task (node slot, <= 4096-row chunk,
// feature group) by the largest label magnitude of the task's rows, and cut into 20-bit limbs; 4096 rows x 20 bits cannot
// overflow a 32-bit counter, so an update is a handful of independent native adds with no carries.
//   * The grid is wide enough that a row converts WITHOUT rounding unless its label is more than ~2^-11 below the task's largest
//     (80-bit windows, four limbs per sum: "wide" form, 9 adds per update), so a cell holds the EXACT sum of MLlib's per-row
//     terms, rounded to fp64 once when the task's histogram is added to the global one (RED.ADD.F64).  Whatever MLlib computes
//     exactly (one-row cells, equal labels, integer labels, class indicators) comes out bit-identical — which matters
//     structurally: a child whose impurity is exactly 0 is a leaf and never draws a feature subset.
//   * Rows that would need rounding (labels tiny next to the task's largest) are not put on the grid at all: their three fp64
//     terms go straight to the global histogram, so a cell that holds only such a row is still exact.
//   * When all labels of the fit are integers of modest size (classification; the labs' price labels) the 60-bit windows of
//     the "narrow" form (three limbs per sum, 7 adds per update) are exact for every row, and that form runs.
//   * Tasks whose labels are not finite or lie outside 2^+-460 take the plain fp64 path (global atomics: MLlib's own overflow /
//     underflow behaviour), as do tasks with fewer updates than histogram cells.
// Data path.  Labels travel WITH the per-tree row lists (a y list parallel to the packed row | weight list, moved by the K6
// partitions): a row visit costs one code-row gather instead of that plus a 32 B label sector.  The code rows of a warp's next block
// are fetched one block ahead — by default with 16-byte loads held in registers (ld.global.nc.L2::64B: a 64-byte DRAM fetch per
// miss instead of the 128-byte line) and stored to the warp's shared-memory tile just before use; for row widths other than 64 /
// 128 bytes, or on request, by the TMA engine (one cp.async.bulk per row completing on an mbarrier, two stages per warp).  A lane
// reads its feature's code with one LDS.U8 and the row's packed limbs with one or two broadcast loads at immediate offsets: no
// 64-bit address arithmetic and no load-to-use stall in the loop.  Feature groups narrower than 32 split the limb PLANES, not the
// rows, over the lane groups, so every add is conflict-free (see the kernel's parameter list).
// Compiled with -fmad=false like the rest of the library.
#include "b2rf_internal.h"
#include "b2rf_device.cuh"
#include "../../include/b2rf.h"

#include <atomic>
#include <cstdlib>
#include <cstring>

namespace b2rf {

constexpr uint32_t LIMB_MASK = (1u << 20) - 1u;
constexpr int LIMB_EXP_MAX = 460;          // |exponent| of the task scale kept inside the range where every scale factor is a normal double
constexpr int LIMB_WBITS = 5;              // bag weights are packed in 5 bits (<= 31)

__device__ __forceinline__ double pow2d(int k) { return __hiloint2double((1023 + k) << 20, 0); }   // 2^k, k in the normal range
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}

// Per-task grid.  2^ey <= max |y| < 2^(ey+1); |w*y| < 2^Ea and w*y*y < 2^Eb with Ea = ey + 1 + 5, Eb = 2 ey + 2 + 5.
//   narrow: v = w*y * 2^(59 - Ea) (|v| < 2^59), u = w*y*y * 2^(60 - Eb) (u < 2^60): three 20-bit limbs each
//   wide  : w*y * 2^(39 - Ea) = hi + lo / 2^40 with |hi| < 2^39, 0 <= lo < 2^40 (four limbs), likewise w*y*y * 2^(40 - Eb)
struct LimbScale { double sa, sb; };

// Packed values of one row; `slow` = the row is not on the task's grid (its limbs are then all zero: the add loop adds nothing for it,
// and the block's slow pass sends its fp64 terms to the global histogram).  Narrow form, one 16-byte word:
//   x = (v mod 2^59) | w << 59 ;  y = u
__device__ __forceinline__ uint4 limb_pack_narrow(uint32_t w, double yv, LimbScale sc, bool& slow) {
  const double wy = (double)w * yv, wyy = wy * yv;                     // MLlib: instanceWeight * label, then * label
  const double ta = wy * sc.sa, tb = wyy * sc.sb;
  const long long v = __double2ll_rn(ta);
  const unsigned long long u = __double2ull_rn(tb);
  slow = !((double)v == ta && (double)u == tb);                        // on the grid <=> exact
  if (slow) return make_uint4(0u, 0u, 0u, 0u);
  const unsigned long long x = ((unsigned long long)v & ((1ull << 59) - 1ull)) | ((unsigned long long)w << 59);
  return make_uint4((uint32_t)x, (uint32_t)(x >> 32), (uint32_t)u, (uint32_t)(u >> 32));
}
// Wide form, 24 bytes: x0 = loA | (hiB >> 24) << 40 | w << 56 ;  x1 = hiA (two's complement) ;  x2 = loB | (hiB & 0xffffff) << 40
__device__ __forceinline__ void limb_pack_wide(uint32_t w, double yv, LimbScale sc, uint4& p0, uint2& p1, bool& slow) {
  const double wy = (double)w * yv, wyy = wy * yv;
  const double ta = wy * sc.sa, tb = wyy * sc.sb;                       // |ta| < 2^39, 0 <= tb < 2^40 (power-of-two scales: exact)
  const double ta40 = ta * 0x1p40, tb40 = tb * 0x1p40;
  slow = !(ta40 == rint(ta40) && tb40 == rint(tb40));                  // bits below the window
  unsigned long long x0 = 0ull, x1 = 0ull, x2 = 0ull;
  if (!slow) {
    const double fa = floor(ta), fb = floor(tb);
    const long long hiA = __double2ll_rz(fa);                           // fractional parts: exact (their exact values are representable)
    const unsigned long long loA = __double2ull_rz((ta - fa) * 0x1p40), hiB = __double2ull_rz(fb), loB = __double2ull_rz((tb - fb) * 0x1p40);
    x0 = loA | ((hiB >> 24) << 40) | ((unsigned long long)w << 56);
    x1 = (unsigned long long)hiA;
    x2 = loB | ((hiB & 0xffffffull) << 40);
  }
  p0 = make_uint4((uint32_t)x0, (uint32_t)(x0 >> 32), (uint32_t)x1, (uint32_t)(x1 >> 32));
  p1 = make_uint2((uint32_t)x2, (uint32_t)(x2 >> 32));
}

// The per-plane add values of one row in plane order (what the packed forms above encode), for the plane-split lane mapping:
//   wide  : L[0] = w, L[1..4] = the four 20-bit limbs of w*y (top one signed), L[5..8] = those of w*y*y
//   narrow: L[0] = w, L[1..3] / L[4..6] likewise with three limbs
template <int NP>
__device__ __forceinline__ void limb_values(uint32_t w, double yv, LimbScale sc, uint32_t (&L)[NP], bool& slow) {
#pragma unroll
  for (int i = 0; i < NP; i++) L[i] = 0u;
  if constexpr (NP == 2) { slow = false; L[0] = w; L[1] = yv != 0.0 ? w : 0u; return; }
  const double wy = (double)w * yv, wyy = wy * yv;
  const double ta = wy * sc.sa, tb = wyy * sc.sb;
  if constexpr (NP == 9) {
    const double ta40 = ta * 0x1p40, tb40 = tb * 0x1p40;
    slow = !(ta40 == rint(ta40) && tb40 == rint(tb40));
    if (slow) return;
    const double fa = floor(ta), fb = floor(tb);
    const long long hiA = __double2ll_rz(fa);
    const unsigned long long loA = __double2ull_rz((ta - fa) * 0x1p40), hiB = __double2ull_rz(fb), loB = __double2ull_rz((tb - fb) * 0x1p40);
    L[0] = w;
    L[1] = (uint32_t)loA & LIMB_MASK; L[2] = (uint32_t)(loA >> 20); L[3] = (uint32_t)hiA & LIMB_MASK; L[4] = (uint32_t)(hiA >> 20);
    L[5] = (uint32_t)loB & LIMB_MASK; L[6] = (uint32_t)(loB >> 20); L[7] = (uint32_t)hiB & LIMB_MASK; L[8] = (uint32_t)(hiB >> 20);
  } else if constexpr (NP == 7) {
    const long long v = __double2ll_rn(ta);
    const unsigned long long u = __double2ull_rn(tb);
    slow = !((double)v == ta && (double)u == tb);
    if (slow) return;
    L[0] = w;
    L[1] = (uint32_t)v & LIMB_MASK; L[2] = (uint32_t)(v >> 20) & LIMB_MASK; L[3] = (uint32_t)(v >> 40);
    L[4] = (uint32_t)u & LIMB_MASK; L[5] = (uint32_t)(u >> 20) & LIMB_MASK; L[6] = (uint32_t)(u >> 40);
  }
}

// G    features per task group = lanes per row (32, 16 or 8).  G = 32: a warp instruction is one row, lane = feature.  G < 32
//      ("plane split"): a warp instruction is still ONE row; the H = 32 / G lane groups take the planes p == q (mod H) of the row's
//      cells.  Cells are [bin][plane][G], so the lanes of one feature (same code) land in different bank groups whatever the code is:
//      always one wavefront.  (Two or four ROWS per instruction collide whenever any feature's codes agree mod H: measured 2.00
//      wavefronts per add for G = 16 on C4.)  Packed values: H records per row, record q = the add values of planes q, q + H, ...
// LDR  bytes per row of the row-major code matrix (0: read a.ldr)
// BR   rows per staged block (32 or 16)
// NP   planes per cell: 7 = count + 3 + 3 limbs (narrow), 9 = count + 4 + 4 (wide), 2 = count + sum of w*y (labels all 0 or 1:
//      binary classification; w*y*y == w*y, and both sums are small integers)
// TMA  how a block's code rows reach shared memory.  true: one bulk async copy per row on the TMA engine (UBLKCP + mbarrier, two
//      tile slots per warp); costs no LSU wavefronts, but UBLKCP is a uniform-datapath instruction — one copy per warp instruction —
//      so a block is BR serialized R2UR / UBLKCP sequences, ~12 issue slots per row (ncu: profiles/r02_ncu_limb.md).  false (needs a
//      compile-time LDR): every lane fetches 16-byte pieces with LDG.128 one block ahead, holds them in registers while the warp adds
//      the current block, and stores them with STS.128 (0.5 wavefronts and ~1 issue slot per row; one tile slot per warp).
template <int G, int LDR, int BR, int THREADS, int NP, bool TMA>
__global__ void __launch_bounds__(THREADS, THREADS <= 384 ? (NP == 2 ? 3 : 2) : 1) limb_hist_kernel(const BucketLaunch a, uint32_t* ticket) {
  extern __shared__ __align__(16) unsigned char smraw[];
  __shared__ uint32_t s_red[THREADS / 32];
  __shared__ uint32_t s_task;
  constexpr int W = THREADS / 32;
  constexpr int R = 32 / G;                  // rows per warp instruction
  constexpr int NU = BR / R;                 // warp instructions per staged block
  constexpr bool WIDE = NP == 9, BIN = NP == 2;
  constexpr int H = 32 / G;                    // lane groups
  constexpr bool PS = G < 32;                  // plane-split mapping
  constexpr bool SQ = NP == 9 && H == 2;       // five planes for lane group 0: the count rides in bits 20.. of plane 2's word
  constexpr int KW = SQ ? 4 : (NP + H - 1) / H;                     // words per record
  constexpr int RB = KW <= 1 ? 4 : (KW == 2 ? 8 : 16);              // bytes per record
  constexpr int VB = PS ? H * RB : (WIDE ? 32 : (BIN ? 8 : 16));    // bytes of packed values per row (wide, G = 32: 24 used)
  constexpr unsigned FULL = 0xffffffffu;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ldr = LDR ? LDR : (int)a.ldr;
  const int padB = a.padB;
  uint32_t* hist = reinterpret_cast<uint32_t*>(smraw);                 // [padB][NP][G]
  const size_t histBytes = (size_t)padB * NP * G * 4;
  static_assert(TMA || LDR > 0, "register staging needs a compile-time row width");
  constexpr int NPC = TMA ? 1 : (BR * (LDR > 0 ? LDR : 16)) / 512;   // 16-byte pieces of a block's code rows per lane
  static_assert(TMA || (BR * LDR) % 512 == 0, "a block's code rows must be a whole number of 512-byte warp stores");
  // per-warp ring: [tile slots: 2 (TMA) or 1][packed values: 2 slots]
  const int tileBytes = BR * ldr, valBytes = BR * VB;
  const int warpBytes = (TMA ? 2 : 1) * tileBytes + 2 * valBytes;
  unsigned char* ring = smraw + histBytes + (size_t)warp * warpBytes;
  unsigned char* valsBase = ring + (TMA ? 2 : 1) * tileBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smraw + histBytes + (size_t)W * warpBytes) + warp * 2;
  const int fl = lane & (G - 1), r = lane / G;
  const int nG = (a.Fsub + G - 1) / G;
  const uint32_t total = a.chunk_prefix[a.nSlots] * (uint32_t)nG;
  const uint8_t* bins_rm = a.bins_rm;

  // staged code rows of blocks that are not full keep whatever the slot held before; rows past the end of a chunk carry zero limbs,
  // so such a code only has to be a valid bin index: start from zeros (and every later content is a real code row)
  for (int c = lane; c < warpBytes >> 4; c += 32) reinterpret_cast<uint4*>(ring)[c] = make_uint4(0u, 0u, 0u, 0u);
  if (TMA) {
    if (lane == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");        // generic-proxy zeros before the async proxy writes the same bytes
  }
  __syncthreads();
  uint32_t phase = 0u;                                                   // bit s: parity the next wait on stage s expects

  for (;;) {
    if (tid == 0) s_task = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t task = s_task;
    if (task >= total) break;
    const uint32_t bt = task / (uint32_t)nG;
    const int g = (int)(task - bt * (uint32_t)nG);
    int lo = 0, hi = a.nSlots;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (a.chunk_prefix[mid] <= bt) lo = mid; else hi = mid; }
    const int slot = lo;
    const uint32_t begin = (bt - a.chunk_prefix[slot]) * (uint32_t)a.CH;
    DSeg sg; const uint32_t* rl; const double* yl; const uint2* fo; double* hb; int slotCells;
    if (a.slots) {   // slots of several trainers in one launch (fitMultiple shared passes)
      const SlotSrc ss = a.slots[slot];
      sg = *ss.seg;
      rl = (sg.buf ? ss.idx1 : ss.idx0) + sg.start + begin;
      yl = (sg.buf ? ss.y1 : ss.y0) + sg.start + begin;
      fo = ss.fo; hb = ss.hb; slotCells = ss.cells;
    } else {
      sg = a.dseg[a.slot_pos[slot]];
      rl = (sg.buf ? a.idx1 : a.idx0) + sg.start + begin;
      yl = (sg.buf ? a.y1 : a.y0) + sg.start + begin;
      fo = a.slot_fo + (size_t)slot * a.Fsub;
      hb = a.hist + (size_t)a.slot_base[slot] * 4;
      slotCells = a.slot_cells[slot];
    }
    const int n = (int)min((uint32_t)a.CH, sg.count - begin);
    const int j = g * G + fl;
    const bool active = j < a.Fsub;
    const uint2 e = active ? fo[j] : make_uint2(0u, 0u);
    // ---- task scale: largest label magnitude of the task's rows (the high word is monotone in |y|) ----
    uint32_t ex = 0;
    if (!BIN) for (int i = tid; i < n; i += THREADS) ex = max(ex, (uint32_t)__double2hiint(yl[i]) & 0x7fffffffu);
    ex = __reduce_max_sync(FULL, ex);
    if (lane == 0) s_red[warp] = ex;
    bool direct = (int64_t)n * a.Fsub < (int64_t)slotCells;   // fewer updates than cells: skip the shared-memory round trip
    if (!direct) {
      uint4* h4 = reinterpret_cast<uint4*>(hist);
      for (int c = tid; c < (int)(histBytes >> 4); c += THREADS) h4[c] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < W; i++) ex = max(ex, s_red[i]);
    int ey = (int)(ex >> 20) - 1023;
    if (ex == 0u) ey = 0;                                             // all labels zero (e.g. a pure class-0 node)
    else if (ey > LIMB_EXP_MAX || ey < -LIMB_EXP_MAX) direct = true;   // non-finite labels, or y*y outside the normal fp64 range: plain fp64 path

    if (direct) {
      for (int base = warp * 32; base < n; base += W * 32) {
        const bool in = base + lane < n;
        const uint32_t ent = in ? rl[base + lane] : 0u;
        const double yv = in ? yl[base + lane] : 0.0;
        const double wy = (double)(ent >> 27) * yv;                    // MLlib: instanceWeight * label * label == (w*y)*y
        const double wyy = wy * yv;
#pragma unroll 4
        for (int u = 0; u < G; u++) {
          const int src = u * R + r;
          const uint32_t e2 = __shfl_sync(FULL, ent, src);
          const double wy2 = __shfl_sync(FULL, wy, src), wyy2 = __shfl_sync(FULL, wyy, src);
          if (active && (e2 >> 27)) {
            const uint32_t code = bins_rm[(size_t)(e2 & 0x7ffffffu) * ldr + e.x];
            double* c = hb + (size_t)(e.y + code) * 4;
            red_add_f64(c, (double)(e2 >> 27)); red_add_f64(c + 1, wy2); red_add_f64(c + 2, wyy2);
          }
        }
      }
      continue;   // the next iteration's __syncthreads (ticket) separates the tasks
    }

    const int Ea = ey + 1 + LIMB_WBITS, Eb = 2 * ey + 2 + LIMB_WBITS;
    LimbScale sc;
    sc.sa = pow2d((WIDE ? 39 : 59) - Ea);
    sc.sb = pow2d((WIDE ? 40 : 60) - Eb);
    const int nblk = (n + BR - 1) / BR;
    uint32_t ent = 0u; double yv = 0.0;
    auto load = [&](int blk) {
      const int i = blk * BR + lane;
      const bool in = blk < nblk && lane < BR && i < n;
      ent = in ? rl[i] : 0u;
      yv = in ? yl[i] : 0.0;
    };
    uint32_t* hcell = hist + fl;
    // packed values of the block held in (ent, yv) -> value slot s; returns the block's slow rows as a lane mask
    auto put_vals = [&](int s) -> uint32_t {
      unsigned char* vals = valsBase + s * valBytes;
      const uint32_t w = ent >> 27;
      bool slow = false;
      if (PS) {
        if (lane < BR) {
          uint32_t L[NP];
          limb_values<NP>(w, yv, sc, L, slow);
#pragma unroll
          for (int q = 0; q < H; q++) {
            uint32_t rec[4] = {0u, 0u, 0u, 0u};
            if constexpr (SQ) {
              if (q == 0) { rec[0] = L[2] | (L[0] << 20); rec[1] = L[4]; rec[2] = L[6]; rec[3] = L[8]; }
              else { rec[0] = L[1]; rec[1] = L[3]; rec[2] = L[5]; rec[3] = L[7]; }
            } else {
#pragma unroll
              for (int k = 0; k < KW; k++) if (q + k * H < NP) rec[k] = L[q + k * H];
            }
            unsigned char* d = vals + lane * VB + q * RB;
            if (RB == 16) *reinterpret_cast<uint4*>(d) = make_uint4(rec[0], rec[1], rec[2], rec[3]);
            else if (RB == 8) *reinterpret_cast<uint2*>(d) = make_uint2(rec[0], rec[1]);
            else *reinterpret_cast<uint32_t*>(d) = rec[0];
          }
        }
      } else if (lane < BR) {
        if (BIN) {
          *reinterpret_cast<uint2*>(vals + lane * VB) = make_uint2(w, yv != 0.0 ? w : 0u);
        } else if (WIDE) {
          uint4 p0; uint2 p1;
          limb_pack_wide(w, yv, sc, p0, p1, slow);
          *reinterpret_cast<uint4*>(vals + lane * VB) = p0;
          *reinterpret_cast<uint2*>(vals + lane * VB + 16) = p1;
        } else {
          *reinterpret_cast<uint4*>(vals + lane * VB) = limb_pack_narrow(w, yv, sc, slow);
        }
      }
      return __ballot_sync(FULL, slow && w != 0u);
    };
    // the adds of one staged block: lane = (row slot r, feature fl); tile / vals point at the block's code rows and packed values;
    // slowMask = rows of the block (first row: chunk row blk * BR) that are not on the grid
    auto consume = [&](const unsigned char* tile, const unsigned char* vals, uint32_t slowMask, int blk) {
      const unsigned char* tl = tile + r * ldr + e.x;
      const unsigned char* vv = vals + r * VB;
      if (PS) {
        if (active) {
          // one row per instruction; lane group r adds the planes r, r + H, ... of its feature's cell
          const unsigned char* tp = tile + e.x;
          const unsigned char* vp = vals + r * RB;
          uint32_t* cb = hcell + r * G;
          uint32_t code = tp[0];
          uint4 rc = make_uint4(0u, 0u, 0u, 0u);
          auto ldrec = [&](const unsigned char* a) {
            if (RB == 16) rc = *reinterpret_cast<const uint4*>(a);
            else if (RB == 8) { const uint2 t2 = *reinterpret_cast<const uint2*>(a); rc.x = t2.x; rc.y = t2.y; }
            else rc.x = *reinterpret_cast<const uint32_t*>(a);
          };
          ldrec(vp);
#pragma unroll
          for (int u = 0; u < BR; u++) {
            const uint32_t cc = code; const uint4 cur = rc;
            if (u + 1 < BR) { code = tp[(u + 1) * ldr]; ldrec(vp + (u + 1) * VB); }
            uint32_t* c = cb + cc * (NP * G);
            if (SQ) {
              const uint32_t v0 = r ? cur.x : cur.x >> 20, v1 = r ? cur.y : cur.x & LIMB_MASK, v2 = r ? cur.z : cur.y, v3 = r ? cur.w : cur.z;
              atomicAdd(c, v0); atomicAdd(c + 2 * G, v1); atomicAdd(c + 4 * G, v2); atomicAdd(c + 6 * G, v3);
              if (r == 0) atomicAdd(c + 8 * G, cur.w);
            } else {
              const uint32_t v[4] = {cur.x, cur.y, cur.z, cur.w};
#pragma unroll
              for (int k = 0; k < KW; k++) if (r + k * H < NP) atomicAdd(c + k * H * G, v[k]);
            }
          }
        }
      } else if (BIN) {
        if (active) {
          uint32_t code = tl[0];
          uint2 q = *reinterpret_cast<const uint2*>(vv);
#pragma unroll
          for (int u = 0; u < NU; u++) {
            const uint32_t cc = code; const uint2 p = q;
            if (u + 1 < NU) { code = tl[(u + 1) * R * ldr]; q = *reinterpret_cast<const uint2*>(vv + (u + 1) * R * VB); }
            uint32_t* c = hcell + cc * (NP * G);
            atomicAdd(c, p.x); atomicAdd(c + G, p.y);      // rows past the end of the chunk add zeros
          }
        }
      } else if (active) {
        // one row ahead: the next row's code and packed limbs are loaded before this row's adds are issued (the compiler keeps
        // shared loads behind earlier shared atomics, so the overlap has to be written out)
        uint32_t code = tl[0];
        uint4 q0 = *reinterpret_cast<const uint4*>(vv);
        uint2 q1 = WIDE ? *reinterpret_cast<const uint2*>(vv + 16) : make_uint2(0u, 0u);
#pragma unroll
        for (int u = 0; u < NU; u++) {
          const uint32_t cc = code; const uint4 p0 = q0; const uint2 p1 = q1;
          if (u + 1 < NU) {
            code = tl[(u + 1) * R * ldr];
            q0 = *reinterpret_cast<const uint4*>(vv + (u + 1) * R * VB);
            if (WIDE) q1 = *reinterpret_cast<const uint2*>(vv + (u + 1) * R * VB + 16);
          }
          uint32_t* c = hcell + cc * (NP * G);
          const unsigned long long x0 = ((unsigned long long)p0.y << 32) | p0.x, x1 = ((unsigned long long)p0.w << 32) | p0.z;
          if (WIDE) {
            const unsigned long long x2 = ((unsigned long long)p1.y << 32) | p1.x;
            const uint32_t hiBl = (uint32_t)(x2 >> 40) | ((p0.y & 0xff00u) << 16);         // low 32 bits of hiB: bits 24..31 come from x0 bits 40..47
            atomicAdd(c, p0.y >> 24);
            atomicAdd(c + G, p0.x & LIMB_MASK); atomicAdd(c + 2 * G, (uint32_t)(x0 >> 20) & LIMB_MASK);
            atomicAdd(c + 3 * G, p0.z & LIMB_MASK); atomicAdd(c + 4 * G, (uint32_t)((long long)x1 >> 20));
            atomicAdd(c + 5 * G, p1.x & LIMB_MASK); atomicAdd(c + 6 * G, (uint32_t)(x2 >> 20) & LIMB_MASK);
            atomicAdd(c + 7 * G, hiBl & LIMB_MASK); atomicAdd(c + 8 * G, (hiBl >> 20) | (((p0.y >> 16) & 0xffu) << 12));
          } else {
            atomicAdd(c, p0.y >> 27);
            atomicAdd(c + G, p0.x & LIMB_MASK); atomicAdd(c + 2 * G, (uint32_t)(x0 >> 20) & LIMB_MASK);
            atomicAdd(c + 3 * G, (uint32_t)((long long)(x0 << 5) >> 45));
            atomicAdd(c + 4 * G, p0.z & LIMB_MASK); atomicAdd(c + 5 * G, (uint32_t)(x1 >> 20) & LIMB_MASK);
            atomicAdd(c + 6 * G, (uint32_t)(x1 >> 40));
          }
        }
      }
      // slow pass (rare): rows whose labels are tiny next to the task's largest go to the global histogram as MLlib's own fp64 terms
      while (slowMask) {
        const int k = __ffs(slowMask) - 1;
        slowMask &= slowMask - 1u;
        const uint32_t e2 = rl[blk * BR + k];
        const double ys = yl[blk * BR + k];
        if (active && r == 0) {
          const uint32_t w = e2 >> 27;
          const double wy = (double)w * ys;
          double* gc = hb + (size_t)(e.y + tile[k * ldr + e.x]) * 4;
          red_add_f64(gc, (double)w); red_add_f64(gc + 1, wy); red_add_f64(gc + 2, wy * ys);
        }
      }
    };
    int b = warp;
    if (TMA) {
      // tile slot s <- the block held in (ent, yv): one bulk async copy per row, completion on the slot's mbarrier
      uint32_t slowM[2] = {0u, 0u};
      auto stage_tma = [&](int s, int blk) {
        slowM[s] = put_vals(s);
        unsigned char* tile = ring + s * tileBytes;
        const int nvalid = min(BR, n - blk * BR);
        if (lane == 0) mbar_expect_tx(&bars[s], (uint32_t)(nvalid * ldr));
        __syncwarp();
        if (lane < nvalid) bulk_g2s(tile + lane * ldr, bins_rm + (size_t)(ent & 0x7ffffffu) * ldr, (uint32_t)ldr, &bars[s]);
      };
      if (b < nblk) { load(b); stage_tma(0, b); load(b + W); }
      for (int k = 0; b < nblk; k++, b += W) {
        const int s = k & 1;
        if (b + W < nblk) { stage_tma(s ^ 1, b + W); load(b + 2 * W); }
        while (!mbar_try(&bars[s], (phase >> s) & 1u)) { }
        phase ^= 1u << s;
        __syncwarp();
        consume(ring + s * tileBytes, valsBase + s * valBytes, slowM[s], b);
        __syncwarp();
      }
    } else {
      // register staging: the 16-byte pieces of the NEXT block's code rows are in flight (LDG.128) while the warp adds the current one
      constexpr int CPR = (LDR > 0 ? LDR : 16) / 16;               // pieces per row
      uint4 creg[NPC];
      auto fetch = [&](int blk) {                                  // (ent, yv) hold block blk
        const int nvalid = min(BR, n - blk * BR);
#pragma unroll
        for (int q = 0; q < NPC; q++) {
          const int c = lane + 32 * q, row = c / CPR, seg = c - row * CPR;
          const uint32_t e2 = __shfl_sync(FULL, ent, row);
          creg[q] = row < nvalid ? ldg128_sparse(bins_rm + (size_t)(e2 & 0x7ffffffu) * ldr + seg * 16) : make_uint4(0u, 0u, 0u, 0u);
        }
      };
      uint32_t slowM[2] = {0u, 0u};
      if (b < nblk) { load(b); fetch(b); slowM[0] = put_vals(0); load(b + W); }
      for (int k = 0; b < nblk; k++, b += W) {
        const int s = k & 1;
#pragma unroll
        for (int q = 0; q < NPC; q++) reinterpret_cast<uint4*>(ring)[lane + 32 * q] = creg[q];
        __syncwarp();
        if (b + W < nblk) { fetch(b + W); slowM[s ^ 1] = put_vals(s ^ 1); load(b + 2 * W); }
        consume(ring, valsBase + s * valBytes, slowM[s], b);
        __syncwarp();
      }
    }
    __syncthreads();
    // ---- flush: exact limb sums -> fp64 (one rounding), added to the global histogram ----
    const double sA = BIN ? 1.0 : pow2d(Ea - (WIDE ? 79 : 59)), sB = BIN ? 1.0 : pow2d(Eb - (WIDE ? 80 : 60));
    for (int c = tid; c < padB * G; c += THREADS) {
      const int bin = c / G, f2 = c & (G - 1);
      const uint32_t* cell = hist + (size_t)bin * (NP * G) + f2;
      const uint32_t cnt = cell[0];
      if (!cnt) continue;                                                // untouched cell (also every cell of an idle lane)
      double A, B;
      if (BIN) {
        A = B = (double)cell[G];
      } else if (WIDE) {
        A = ((double)(int32_t)cell[4 * G] * 0x1p60 + (double)cell[3 * G] * 0x1p40) + ((double)cell[2 * G] * 0x1p20 + (double)cell[G]);
        B = ((double)cell[8 * G] * 0x1p60 + (double)cell[7 * G] * 0x1p40) + ((double)cell[6 * G] * 0x1p20 + (double)cell[5 * G]);
      } else {
        A = ((double)(int32_t)cell[3 * G] * 0x1p40 + (double)cell[2 * G] * 0x1p20) + (double)cell[G];
        B = ((double)cell[6 * G] * 0x1p40 + (double)cell[5 * G] * 0x1p20) + (double)cell[4 * G];
      }
      double* gp = hb + ((size_t)fo[g * G + f2].y + (size_t)bin) * 4;
      red_add_f64(gp, (double)cnt); red_add_f64(gp + 1, A * sA); red_add_f64(gp + 2, B * sB);
    }
    // the ticket __syncthreads at the top of the next iteration orders this flush before the next task's zeroing
  }
}

static std::atomic<long long> g_variant_count[B2RF_K4_VARIANTS];
void count_k4_variant(int v) { if (v >= 0 && v < B2RF_K4_VARIANTS) g_variant_count[v].fetch_add(1, std::memory_order_relaxed); }
long long k4_variant_count(int v) { return (v >= 0 && v < B2RF_K4_VARIANTS) ? g_variant_count[v].load(std::memory_order_relaxed) : 0; }

// Shared memory of one CTA: the limb histogram, the warps' staging rings (code-row tile slots + two packed-value slots) and, for the
// TMA form, two mbarriers per warp.
static size_t limb_smem(int padB, int G, int ldr, int BR, int threads, int NP, bool tma) {
  size_t vb = NP == 9 ? 32 : (NP == 2 ? 8 : 16);
  if (G < 32) {   // plane split: H records per row (see the kernel's VB)
    const int H = 32 / G, KW = (NP == 9 && H == 2) ? 4 : (NP + H - 1) / H, RB = KW <= 1 ? 4 : (KW == 2 ? 8 : 16);
    vb = (size_t)H * RB;
  }
  const size_t warp = (size_t)(tma ? 2 : 1) * BR * ldr + 2 * (size_t)BR * vb;
  return (size_t)padB * NP * G * 4 + (size_t)(threads / 32) * warp + (size_t)(threads / 32) * 16;
}

// Picks (G, block rows, threads, staging) for a launch, or returns false when the limb kernel does not apply (the caller then runs
// the fp64 compare-and-swap kernel of k4_hist.cu): uint8 codes, labels carried in the row lists, 16-byte aligned code rows, and a
// task histogram that leaves room for the staging ring.
bool limb_plan(const BucketLaunch& b, int maxBinsAll, int form, bool wantTma, LimbPlan* p) {
  if (b.codeBytes != 1 || ((b.dseg == nullptr || b.y0 == nullptr || b.y1 == nullptr) && b.slots == nullptr)) return false;
  if ((b.ldr & 15) != 0 || (reinterpret_cast<uintptr_t>(b.bins_rm) & 15) != 0 || b.ldr > 1024) return false;
  const size_t cap2 = (size_t)113 * 1024, cap1 = (size_t)226 * 1024;   // two CTAs per SM / one
  const int NP = form == 2 ? 2 : (form == 1 ? 7 : 9);   // form: 0 wide, 1 narrow (integer labels), 2 binary (labels 0 / 1)
  const bool tma = wantTma || !(b.ldr == 64 || b.ldr == 128);          // register staging is compiled for the 64- and 128-byte rows
  const int ldr = (int)b.ldr;
  int G = 32;
  while (G > 8 && G / 2 >= b.Fsub) G /= 2;
  for (; G >= 8; G /= 2) {
    // two CTAs per SM: 12 warps each when the ring fits (register staging; 384 threads x 2 fill the register file), else 8
    if (!tma && limb_smem(maxBinsAll, G, ldr, 32, 384, NP, false) <= cap2) { *p = {G, 32, 384, maxBinsAll, NP, 0}; return true; }
    if (!tma && limb_smem(maxBinsAll, G, ldr, 16, 384, NP, false) <= cap2) { *p = {G, 16, 384, maxBinsAll, NP, 0}; return true; }
    if (limb_smem(maxBinsAll, G, ldr, 32, 256, NP, tma) <= cap2) { *p = {G, 32, 256, maxBinsAll, NP, tma}; return true; }
    if (limb_smem(maxBinsAll, G, ldr, 16, 256, NP, tma) <= cap2) { *p = {G, 16, 256, maxBinsAll, NP, tma}; return true; }
    // one CTA per SM (large histograms): 24 warps when the rings fit beside the planes (register staging, 16-row blocks), else 16
    if (!tma && limb_smem(maxBinsAll, G, ldr, 16, 768, NP, false) <= cap1) { *p = {G, 16, 768, maxBinsAll, NP, 0}; return true; }
    if (limb_smem(maxBinsAll, G, ldr, 32, 512, NP, tma) <= cap1) { *p = {G, 32, 512, maxBinsAll, NP, tma}; return true; }
    if (limb_smem(maxBinsAll, G, ldr, 16, 512, NP, tma) <= cap1) { *p = {G, 16, 512, maxBinsAll, NP, tma}; return true; }
  }
  return false;
}

template <int G, int LDR, int BR, int THREADS, int NP, bool TMA>
static cudaError_t launch_limb_inst(const BucketLaunch& b, uint32_t* ticket, int grid, size_t smem, cudaStream_t s) {
  cudaError_t e = cudaFuncSetAttribute(limb_hist_kernel<G, LDR, BR, THREADS, NP, TMA>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  limb_hist_kernel<G, LDR, BR, THREADS, NP, TMA><<<grid, THREADS, smem, s>>>(b, ticket);
  return cudaGetLastError();
}

template <int G, int NP, int BR>
static cudaError_t launch_limb_g(const BucketLaunch& b, const LimbPlan& p, uint32_t* ticket, int grid, size_t smem, cudaStream_t s) {
  if (!p.tma) {   // register staging: 64- or 128-byte rows
    if constexpr (BR == 16) {
      if (p.threads == 768) {
        if (b.ldr == 64) return launch_limb_inst<G, 64, 16, 768, NP, false>(b, ticket, grid, smem, s);
        return launch_limb_inst<G, 128, 16, 768, NP, false>(b, ticket, grid, smem, s);
      }
    }
    if (b.ldr == 64) {
      if (p.threads == 384) return launch_limb_inst<G, 64, BR, 384, NP, false>(b, ticket, grid, smem, s);
      if (p.threads == 256) return launch_limb_inst<G, 64, BR, 256, NP, false>(b, ticket, grid, smem, s);
      return launch_limb_inst<G, 64, BR, 512, NP, false>(b, ticket, grid, smem, s);
    }
    if (p.threads == 384) return launch_limb_inst<G, 128, BR, 384, NP, false>(b, ticket, grid, smem, s);
    if (p.threads == 256) return launch_limb_inst<G, 128, BR, 256, NP, false>(b, ticket, grid, smem, s);
    return launch_limb_inst<G, 128, BR, 512, NP, false>(b, ticket, grid, smem, s);
  }
  // TMA staging: any row width; the two named widths get compile-time offsets in the add loop
  if (p.threads == 256) {
    if (b.ldr == 64) return launch_limb_inst<G, 64, BR, 256, NP, true>(b, ticket, grid, smem, s);
    return launch_limb_inst<G, 0, BR, 256, NP, true>(b, ticket, grid, smem, s);
  }
  if (b.ldr == 128) return launch_limb_inst<G, 128, BR, 512, NP, true>(b, ticket, grid, smem, s);
  return launch_limb_inst<G, 0, BR, 512, NP, true>(b, ticket, grid, smem, s);
}

template <int G, int NP>
static cudaError_t launch_limb_br(const BucketLaunch& b, const LimbPlan& p, uint32_t* ticket, int grid, size_t smem, cudaStream_t s) {
  if (p.BR == 32) return launch_limb_g<G, NP, 32>(b, p, ticket, grid, smem, s);
  return launch_limb_g<G, NP, 16>(b, p, ticket, grid, smem, s);
}

// Task table (seg_tasks, which also zeroes the ticket behind the prefix array) + the limb kernel.
cudaError_t launch_limb_hist(const BucketLaunch& b0, const LimbPlan& p, cudaStream_t s) {
  BucketLaunch b = b0;
  b.padB = p.padB; b.G = p.G;
  cudaError_t e = b.slots ? launch_seg_tasks_slots(b.slots, b.nSlots, b.CH, const_cast<uint32_t*>(b.chunk_prefix), s)
                          : launch_seg_tasks(b.dseg, b.slot_pos, b.nSlots, b.CH, const_cast<uint32_t*>(b.chunk_prefix), s, /*ticket=*/true, b.rows_acc);
  if (e != cudaSuccess) return e;
  uint32_t* ticket = const_cast<uint32_t*>(b.chunk_prefix) + b.nSlots + 1;
  const size_t smem = limb_smem(p.padB, p.G, (int)b.ldr, p.BR, p.threads, p.NP, p.tma != 0);
  // the two-plane form is bound by its gathers (two adds per update): a third CTA per SM when the shared memory allows (its launch
  // bounds keep it at 56 registers)
  const int perSM = p.threads > 384 ? 1 : (p.NP == 2 && (smem + 1024) * 3 <= (size_t)227 * 1024 ? 3 : 2);
  const int grid = device_sm_count() * perSM;
  count_k4_variant(B2RF_K4_LIMB);
  if (p.NP == 2) {
    if (p.G == 32) return launch_limb_br<32, 2>(b, p, ticket, grid, smem, s);
    if (p.G == 16) return launch_limb_br<16, 2>(b, p, ticket, grid, smem, s);
    return launch_limb_br<8, 2>(b, p, ticket, grid, smem, s);
  }
  if (p.NP == 7) {
    if (p.G == 32) return launch_limb_br<32, 7>(b, p, ticket, grid, smem, s);
    if (p.G == 16) return launch_limb_br<16, 7>(b, p, ticket, grid, smem, s);
    return launch_limb_br<8, 7>(b, p, ticket, grid, smem, s);
  }
  if (p.G == 32) return launch_limb_br<32, 9>(b, p, ticket, grid, smem, s);
  if (p.G == 16) return launch_limb_br<16, 9>(b, p, ticket, grid, smem, s);
  return launch_limb_br<8, 9>(b, p, ticket, grid, smem, s);
}

// flag[0]: 1 when every label is an integer of magnitude <= 2^24 (then w*y and w*y*y are integers below 2^53 and the narrow windows
// hold every row exactly), else 0.  flag[1]: 1 when every label is 0 or 1 (binary classification: the two-plane form).  Both must be
// 1 on entry.
__global__ void labels_integer_kernel(const double* y, int64_t n, int32_t* flag) {
  bool ok = true, ok01 = true;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double v = y[i];
    ok = ok && (fabs(v) <= 16777216.0) && (v == rint(v));
    ok01 = ok01 && (v == 0.0 || v == 1.0);
  }
  if (!__all_sync(0xffffffffu, ok) && (threadIdx.x & 31) == 0) flag[0] = 0;
  if (!__all_sync(0xffffffffu, ok01) && (threadIdx.x & 31) == 0) flag[1] = 0;
}
cudaError_t launch_labels_integer(const double* y, int64_t n, int32_t* d_flag, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  int64_t bx = (n + 255) / 256;
  if (bx > 2048) bx = 2048;
  labels_integer_kernel<<<(unsigned)bx, 256, 0, s>>>(y, n, d_flag);
  return cudaGetLastError();
}

}  // namespace b2rf

extern "C" {

/* How many times each K4 implementation has been launched by this process (test / evidence hook):
 * 0 = limb kernel (k4_limb.cu), 1 = fp64 kernel for <= 32 features (bucket_hist32_kernel), 2 = general fp64 kernel,
 * 3 = multiclass class-count launches, 4 = column-scan kernels (hist_kernel<0|1|2>), 5 = multiclass launches in the staged form. */
int b2rf_k4_launch_counts(int64_t* h_counts, int32_t n) {
  if (!h_counts || n <= 0) return B2RF_ERR_ARG;
  for (int i = 0; i < n; i++) h_counts[i] = (int64_t)b2rf::k4_variant_count(i);
  return B2RF_OK;
}

}  // extern "C"
