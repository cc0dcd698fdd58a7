// The two 27-cell neighbour passes of the SPH step: density + pressure, and forces.
//
// Reference: libclsph/kernels/sphb.cl:3-35 (density_pressure -> forces.cl:14-42 ->
// smoothing.cl:1-4) and sphb.cl:37-126 (forces, with smoothing.cl:6-34).  Both visit, for
// every particle, the 27 cells of side 2h around its own cell, although the smoothing
// kernels vanish beyond h: ~1000 candidate pairs per particle for ~20-40 true neighbours.
// A pair with r >= h contributes exactly zero to every sum, so the kernels here (a) search a
// grid of side h instead and (b) skip pairs with r >= h before doing any work on them; results
// are unchanged up to summation order.
//
// Particles are sorted by the row-major index of their fine cell — side h in y and z, h / kXSub
// along x (common.cuh) — so the x-adjacent cells of each of the 9 (dy, dz) rows are ONE contiguous
// run [ftable[row + first], ftable[row + last + 1]), and that run is trimmed to the chord of the
// support sphere at the row (row_reach): ~70 candidates per particle at kXSub = 4.
//
// BOTH passes give one thread a PAIR of consecutive particles (2t, 2t + 1).  Two particles that
// are neighbours in the sorted array sit in the same row of cells a few sub-cells apart, so their
// candidate runs and their neighbour sets largely coincide:
//   * the density pass walks the UNION of the two particles' runs once — every 16-byte candidate
//     load serves two distance tests, done together in packed fp32x2 arithmetic — and records
//     ONE list per pair: the candidates that are inside h of either particle;
//   * the forces pass evaluates every listed neighbour against both particles at once, again in
//     packed fp32x2 arithmetic, with (h - r) and (h^2 - r^2) clamped at zero so that a neighbour
//     of only one of the two contributes exactly nothing to the other.
// A pair that straddles two rows (or sits far apart in one) is "split": the thread walks each
// particle's own rows in turn (row ids 0-8, then 9-17), and the list says where the second part
// begins.
//
// This translation unit is compiled with FMA contraction ON and uses rsqrt for 1/r; it is
// checked against the CPU oracle to a stated tolerance (tests/test_gpu_parity.py), unlike
// kernels_exact.cu which is bit-exact.
//
// Data the two passes exchange (both are bound by L1 data-pipe wavefronts — one per 128-byte
// line a warp-wide load touches — so both read 32-byte records with ONE 256-bit load):
//   pairs[t]  {x0, x1, y0, y1 | z0, z1, -, -} of particles 2t, 2t + 1, written by the reorder
//             pass: the density pass's candidates, two per load, coordinates already paired
//             for fp32x2 arithmetic;
//   pv[i]     {x, y, z, p/rho^2 | vx, vy, vz, m/rho}, written by the density pass: everything
//             the forces pass needs of a neighbour in one gather.  The reference recomputes
//             p_j/rho_j^2 and m/rho_j for every PAIR (sphb.cl:80-106); they depend only on j.
#include <cstdlib>

#include "common.cuh"

namespace sphb {

namespace {

#include "neighbors_common.cuh"
#include "neighbors_v1.cuh"

// ---- the pair list --------------------------------------------------------------------------
// One pool of 32-bit words per engine, addressed by pair index t = i / 2 with np = ceil(n / 2):
//   meta[t]                  bits 0-7 entries of the first part, 8-15 entries in all, bit 31 =
//                            "list unusable" (it overflowed): the pair is queued instead
//   qcount[0], queue[]       the pairs whose list is unusable, in no particular order: a follow-up
//                            kernel (k_forces_overflow) lets their particles search on their own,
//                            so that no warp of the forces pass mixes the two kinds of work
//   start[r * np + t]        first particle of the run walked for row id r (0-8; 9-17 for the
//                            second part of a split pair — written only then)
//   codes[(k / 2) * np + t]  entries k and k + 1 as two 16-bit codes: row id << 11 | offset in
//                            that row's run; kNullCode pads every part to a multiple of 4
// Everything is laid out pair-minor, so the warp's stores and the forces pass's loads coalesce.
constexpr int kUB = 64;                              // threads per CTA (128 particles)
constexpr uint32_t kRowShift = 11;
constexpr uint32_t kOffMask = (1u << kRowShift) - 1u;
constexpr uint32_t kNullCode = kOffMask;             // row 0, offset 2047: never recorded
constexpr uint32_t kMaxRun = kOffMask;               // a longer run cannot be coded: unusable
constexpr uint32_t kSpanTogether = 2u * kXSub;       // sub-cells two particles may be apart (2h)
constexpr uint32_t kUDump = 8;                       // rows past the capacity that absorb hits
constexpr uint32_t kUStride = 2u * kUB;              // bytes between two rows of the hit columns
constexpr uint32_t kMetaUnusable = 0x80000000u;
constexpr uint32_t kStartRows = 18;

struct PairLists {
  uint32_t* meta;
  uint32_t* qcount;  // 4 words (16-byte aligned rest)
  uint32_t* queue;
  uint32_t* start;
  uint32_t* codes;
  uint32_t np;
};
__host__ __device__ inline PairLists pair_lists(uint32_t* pool, uint32_t n) {
  PairLists L;
  L.np = (n + 1u) >> 1;
  L.meta = pool;
  L.qcount = pool + L.np;
  L.queue = L.qcount + 4;
  L.start = L.queue + L.np;
  L.codes = L.start + (size_t)kStartRows * L.np;
  return L;
}

__device__ __forceinline__ float2 bcast2(float v) { return make_float2(v, v); }

// 32 bytes with one load (LDG.E.256): a, b the first float4, c, d the second
struct Rec8 {
  float2 a, b, c, d;
};
__device__ __forceinline__ Rec8 ldg256(const float4* p) {
  Rec8 r;
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.a.x), "=f"(r.a.y), "=f"(r.b.x), "=f"(r.b.y), "=f"(r.c.x), "=f"(r.c.y), "=f"(r.d.x),
                 "=f"(r.d.y)
               : "l"(p));
  return r;
}

// TWO candidates (one record of `pairs`) against the thread's TWO particles: per particle
// u = r^2 - h^2 for both candidates in packed fp32x2 arithmetic (FADD2 / FFMA2 / FMUL2, the
// particle's coordinates as broadcast operands); where u < 0 the particle's sum takes u^3 (the
// caller negates the total), and where either particle's is, the candidate's code goes to the
// pair's list: a predicated store at a cursor that advances only on hits.
template <uint32_t STRIDE>  // bytes between two entries of a hit column
__device__ __forceinline__ void density_test_record(const float2& qx, const float2& qy, const float2& qz,
                                                    float neg_h2, const float2& X, const float2& Y,
                                                    const float2& Z, float& sum_a, float& sum_b,
                                                    uint32_t& cur, uint32_t code) {
  float2 d = __fadd2_rn(X, bcast2(-qx.x));
  float2 ua = __ffma2_rn(d, d, bcast2(neg_h2));
  d = __fadd2_rn(Y, bcast2(-qy.x));
  ua = __ffma2_rn(d, d, ua);
  d = __fadd2_rn(Z, bcast2(-qz.x));
  ua = __ffma2_rn(d, d, ua);
  d = __fadd2_rn(X, bcast2(-qx.y));
  float2 ub = __ffma2_rn(d, d, bcast2(neg_h2));
  d = __fadd2_rn(Y, bcast2(-qy.y));
  ub = __ffma2_rn(d, d, ub);
  d = __fadd2_rn(Z, bcast2(-qz.y));
  ub = __ffma2_rn(d, d, ub);
  const float2 ua2 = __fmul2_rn(ua, ua), ub2 = __fmul2_rn(ub, ub);
  asm volatile(
      "{ .reg .pred pa, pb, pc;\n"
      "  .reg .b16 c1;\n"
      "  setp.lt.f32 pa, %3, 0f00000000;\n"
      "  setp.lt.f32 pb, %5, 0f00000000;\n"
      "  @pa fma.rn.f32 %0, %7, %3, %0;\n"
      "  @pb fma.rn.f32 %1, %9, %5, %1;\n"
      "  or.pred pc, pa, pb;\n"
      "  @pc st.shared.u16 [%2], %11;\n"
      "  @pc add.u32 %2, %2, %12;\n"
      "  setp.lt.f32 pa, %4, 0f00000000;\n"
      "  setp.lt.f32 pb, %6, 0f00000000;\n"
      "  @pa fma.rn.f32 %0, %8, %4, %0;\n"
      "  @pb fma.rn.f32 %1, %10, %6, %1;\n"
      "  or.pred pc, pa, pb;\n"
      "  add.u16 c1, %11, 1;\n"
      "  @pc st.shared.u16 [%2], c1;\n"
      "  @pc add.u32 %2, %2, %12; }"
      : "+f"(sum_a), "+f"(sum_b), "+r"(cur)
      : "f"(ua.x), "f"(ua.y), "f"(ub.x), "f"(ub.y), "f"(ua2.x), "f"(ua2.y), "f"(ub2.x), "f"(ub2.y),
        "h"((uint16_t)code), "n"(STRIDE)
      : "memory");
}

// The density pass of ONE pair of particles (2t, 2t + 1): walks the union of their runs, sums both
// densities, records the pair's list (staged in the shared-memory column `col`, entries STRIDE
// bytes apart) and writes rho, p and the pv records.
template <uint32_t STRIDE>
__device__ __forceinline__ void density_pair(uint32_t t, uint16_t* col, const float4* __restrict__ pairs,
                                             const float4* __restrict__ vel, const uint32_t* __restrict__ keys,
                                             const uint32_t* __restrict__ ftable, float* __restrict__ rho,
                                             float* __restrict__ prs, float4* __restrict__ pv, const PairLists& L,
                                             uint32_t cap, const StepConsts& c) {
  const uint32_t ia = 2u * t;
  if (ia >= c.n) return;
  const bool has_b = ia + 1 < c.n;
  const uint32_t ib = has_b ? ia + 1 : ia;
  const Rec8 own = ldg256(pairs + ia);  // {xa, xb, ya, yb | za, zb, -, -}
  const float4 pa = make_float4(own.a.x, own.b.x, own.c.x, 0.f);
  const float4 pb = has_b ? make_float4(own.a.y, own.b.y, own.c.y, 0.f) : pa;
  uint32_t fxa, fya, fza, fxb, fyb, fzb;
  fine_coords_of(__ldg(keys + ia), c, fxa, fya, fza);
  fine_coords_of(__ldg(keys + ib), c, fxb, fyb, fzb);
  const CellFrac fa = cell_frac(pa, fxa, fya, fza, c), fb = cell_frac(pb, fxb, fyb, fzb, c);
  // sorted by row-major cell: inside a row, b's cell is never left of a's
  const bool split = has_b && !(fya == fyb && fza == fzb && fxb - fxa <= kSpanTogether);
  const int npass = split ? 2 : 1;
  const float nan = __int_as_float(0x7fc00000);
  const float4 nowhere = make_float4(nan, nan, nan, 0.f);  // fails every distance test

  // (shared-window byte addresses, so that a store is one predicated STS)
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(col);
  const uint32_t end = base + cap * STRIDE;
  uint32_t cur = base, cnt_a = 0;
  float sum_a = 0.f, sum_b = 0.f;
  bool unusable = false;
#pragma unroll 1
  for (int pass = 0; pass < npass; ++pass) {
    const bool la = pass == 0, lb = has_b && (split ? pass == 1 : true);
    const float4 qa = la ? pa : nowhere, qb = lb ? pb : nowhere;
    const float2 qx = make_float2(qa.x, qb.x), qy = make_float2(qa.y, qb.y), qz = make_float2(qa.z, qb.z);
    const uint32_t ry = la ? fya : fyb, rz = la ? fza : fzb;
    uint32_t rowid = 9u * pass;
#pragma unroll 1
    for (int dz = -1; dz <= 1; ++dz, rowid += 3u) {
      const uint32_t z = rz + dz;  // wraps to a huge value below 0
      if (z >= c.fnz) continue;
      // the three rows of this plane are planned together, so that their six table loads
      // are in flight at once (one load latency per plane, not one per row)
      uint32_t rs[3], re[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int dy = k - 1;
        const uint32_t y = ry + dy;
        const Reach A = row_reach(c, fxa, fa, dy, dz), B = row_reach(c, fxb, fb, dy, dz);
        const bool ona = la && A.on, onb = lb && B.on;
        const uint32_t x_lo = ona ? (onb ? min(A.lo, B.lo) : A.lo) : B.lo;
        const uint32_t x_hi = (ona ? (onb ? max(A.hi, B.hi) : A.hi) : B.hi) + 1u;
        rs[k] = re[k] = 0;
        if (y < c.fny && (ona || onb)) {
          const uint32_t rowbase = (z * c.fny + y) * c.fnx;
          rs[k] = __ldg(ftable + rowbase + x_lo);
          re[k] = __ldg(ftable + rowbase + x_hi);
        }
      }
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const uint32_t s = rs[k], e = re[k];
        L.start[(size_t)(rowid + k) * L.np + t] = s;
        if (e - s > kMaxRun) unusable = true;
        // The run [s, e) is walked as whole records — candidates 2r, 2r + 1 — from the record
        // that holds s to the one that holds e - 1; a candidate outside the run (only the first
        // record's first, when s is odd, and the last record's second, when e is odd) is moved
        // out of reach.  code = row id << 11 | (candidate - s).  The body is branch-free; the
        // cursor is pulled back into the dump rows once per record.
        const uint32_t j_end = s < e ? e : s & ~1u;  // (an empty run walks nothing)
        uint32_t code = ((rowid + k) << kRowShift) - (s & 1u);
#pragma unroll 2
        for (uint32_t j = s & ~1u; j < j_end; j += 2u, code += 2u) {
          Rec8 r = ldg256(pairs + j);
          if (j < s) r.a.x = nan;
          if (j + 1u >= e) r.a.y = nan;
          density_test_record<STRIDE>(qx, qy, qz, -c.h2, r.a, r.b, r.c, sum_a, sum_b, cur, code);
          cur = min(cur, end);
        }
      }
    }
    // every part ends on a multiple of four entries (the forces pass takes four per trip)
#pragma unroll
    for (int u = 0; u < 3; ++u)
      if (((cur - base) / STRIDE) & 3u) {
        asm volatile("st.shared.u16 [%0], %1;" ::"r"(cur), "h"((uint16_t)kNullCode) : "memory");
        cur += STRIDE;
      }
    cur = min(cur, end);
    if (pass == 0) cnt_a = (cur - base) / STRIDE;
  }
  const uint32_t cnt = (cur - base) / STRIDE;
  unusable = unusable || cnt >= cap;  // the dump rows were reached
#pragma unroll
  for (int w = 0; w < 2; ++w) {
    if (w == 1 && !has_b) break;
    const uint32_t i = w ? ib : ia;
    const float density = (c.mass * c.c_poly6) * -(w ? sum_b : sum_a);  // the sums hold -(h^2 - r^2)^3
    const float pressure = tait_pressure(density, c);
    rho[i] = density;
    prs[i] = pressure;
    // the forces pass's record of this particle, with the two per-particle factors of its pair terms
    const float4 p = w ? pb : pa, v = __ldg(vel + i);
    pv[2u * i] = make_float4(p.x, p.y, p.z, pressure / (density * density));
    pv[2u * i + 1u] = make_float4(v.x, v.y, v.z, c.mass / density);
  }
  L.meta[t] = unusable ? kMetaUnusable : (cnt_a | (cnt << 8));
  if (unusable) L.queue[atomicAdd(L.qcount, 1u)] = t;
  if (!unusable) {
    uint32_t* out = L.codes + t;
    for (uint32_t k = 0; k < cnt; k += 2u)
      out[(size_t)(k >> 1) * L.np] = (uint32_t)col[k * (STRIDE / 2u)] | ((uint32_t)col[(k + 1u) * (STRIDE / 2u)] << 16);
  }
}

// Residency of the density pass, measured at 16.7 M (tools/ab_builds.sh): 16 CTAs of 64 threads per
// SM = 64 registers: 1.20 ms; 12 CTAs (80 registers) 1.23; 20 CTAs (48 registers, spills) 1.53; left
// to the compiler with a minimum of 1 (94 registers) 1.33.
#ifndef SPHB_DENSITY_MINB
#define SPHB_DENSITY_MINB 16
#endif
template <bool Q>
__global__ void __launch_bounds__(kUB, SPHB_DENSITY_MINB)
    k_density_union(const float4* __restrict__ pairs, const float4* __restrict__ vel,
                    const uint32_t* __restrict__ keys, const uint32_t* __restrict__ ftable,
                    float* __restrict__ rho, float* __restrict__ prs, float4* __restrict__ pv,
                    uint32_t* __restrict__ pool, uint32_t cap, StepConsts c_arg) {
  if (Q && !g_hot.active) return;
  const StepConsts& c = Q ? g_hot.c : c_arg;
  extern __shared__ uint16_t hit[];  // (cap + kUDump) rows of kUB columns
  const PairLists L = pair_lists(pool, c.n);
  density_pair<kUStride>(blockIdx.x * kUB + threadIdx.x, hit + threadIdx.x, pairs, vel, keys, ftable, rho, prs, pv, L,
                         cap, c);
}

// ---- experiment (sphb_set_option("neighbor_kernel", 7)): four particles per thread -----------------
// The passes are bound by bytes delivered from L1 to registers; a 32-byte record serves 4 distance
// tests in k_density_union.  Here a thread owns FOUR consecutive particles (two pairs): when all
// four sit in one row of cells within 2 h of each other it walks the union of their runs ONCE and
// every record serves 8 tests; it still records one list per PAIR (two shared-memory columns), so
// the forces pass is unchanged.  Any other quad (row ends, n not a multiple of 4) is done as two
// pair walks.
constexpr int kQB = 64;                          // threads per CTA (256 particles)
constexpr uint32_t kQStride = 2u * 2u * kQB;     // bytes between two entries of a column (2 columns per thread)

template <uint32_t STRIDE>
__device__ __forceinline__ void density_test_record_quad(const float2 (&qx)[2], const float2 (&qy)[2],
                                                         const float2 (&qz)[2], float neg_h2, const float2& X,
                                                         const float2& Y, const float2& Z, float (&sum)[4],
                                                         uint32_t& cur0, uint32_t& cur1, uint32_t code) {
  float2 u[4], u2[4];  // per particle: (candidate 0, candidate 1)
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    const float px = (w & 1) ? qx[w >> 1].y : qx[w >> 1].x, py = (w & 1) ? qy[w >> 1].y : qy[w >> 1].x,
                pz = (w & 1) ? qz[w >> 1].y : qz[w >> 1].x;
    float2 d = __fadd2_rn(X, bcast2(-px));
    u[w] = __ffma2_rn(d, d, bcast2(neg_h2));
    d = __fadd2_rn(Y, bcast2(-py));
    u[w] = __ffma2_rn(d, d, u[w]);
    d = __fadd2_rn(Z, bcast2(-pz));
    u[w] = __ffma2_rn(d, d, u[w]);
    u2[w] = __fmul2_rn(u[w], u[w]);
  }
  asm volatile(
      "{ .reg .pred pa, pb, pc, pd, p0, p1;\n"
      "  .reg .b16 c1;\n"
      "  setp.lt.f32 pa, %6, 0f00000000;\n"
      "  setp.lt.f32 pb, %8, 0f00000000;\n"
      "  setp.lt.f32 pc, %10, 0f00000000;\n"
      "  setp.lt.f32 pd, %12, 0f00000000;\n"
      "  @pa fma.rn.f32 %0, %14, %6, %0;\n"
      "  @pb fma.rn.f32 %1, %16, %8, %1;\n"
      "  @pc fma.rn.f32 %2, %18, %10, %2;\n"
      "  @pd fma.rn.f32 %3, %20, %12, %3;\n"
      "  or.pred p0, pa, pb;\n"
      "  or.pred p1, pc, pd;\n"
      "  @p0 st.shared.u16 [%4], %22;\n"
      "  @p0 add.u32 %4, %4, %23;\n"
      "  @p1 st.shared.u16 [%5], %22;\n"
      "  @p1 add.u32 %5, %5, %23;\n"
      "  setp.lt.f32 pa, %7, 0f00000000;\n"
      "  setp.lt.f32 pb, %9, 0f00000000;\n"
      "  setp.lt.f32 pc, %11, 0f00000000;\n"
      "  setp.lt.f32 pd, %13, 0f00000000;\n"
      "  @pa fma.rn.f32 %0, %15, %7, %0;\n"
      "  @pb fma.rn.f32 %1, %17, %9, %1;\n"
      "  @pc fma.rn.f32 %2, %19, %11, %2;\n"
      "  @pd fma.rn.f32 %3, %21, %13, %3;\n"
      "  or.pred p0, pa, pb;\n"
      "  or.pred p1, pc, pd;\n"
      "  add.u16 c1, %22, 1;\n"
      "  @p0 st.shared.u16 [%4], c1;\n"
      "  @p0 add.u32 %4, %4, %23;\n"
      "  @p1 st.shared.u16 [%5], c1;\n"
      "  @p1 add.u32 %5, %5, %23; }"
      : "+f"(sum[0]), "+f"(sum[1]), "+f"(sum[2]), "+f"(sum[3]), "+r"(cur0), "+r"(cur1)
      : "f"(u[0].x), "f"(u[0].y), "f"(u[1].x), "f"(u[1].y), "f"(u[2].x), "f"(u[2].y), "f"(u[3].x), "f"(u[3].y),
        "f"(u2[0].x), "f"(u2[0].y), "f"(u2[1].x), "f"(u2[1].y), "f"(u2[2].x), "f"(u2[2].y), "f"(u2[3].x),
        "f"(u2[3].y), "h"((uint16_t)code), "n"(STRIDE)
      : "memory");
}

__device__ __noinline__ void density_pair_outlined(uint32_t t, uint16_t* col, const float4* __restrict__ pairs,
                                                  const float4* __restrict__ vel, const uint32_t* __restrict__ keys,
                                                  const uint32_t* __restrict__ ftable, float* __restrict__ rho,
                                                  float* __restrict__ prs, float4* __restrict__ pv, const PairLists& L,
                                                  uint32_t cap, const StepConsts& c) {
  density_pair<kQStride>(t, col, pairs, vel, keys, ftable, rho, prs, pv, L, cap, c);
}

#ifndef SPHB_QUAD_MINB
#define SPHB_QUAD_MINB 10
#endif
__global__ void __launch_bounds__(kQB, SPHB_QUAD_MINB)
    k_density_quad(const float4* __restrict__ pairs, const float4* __restrict__ vel,
                   const uint32_t* __restrict__ keys, const uint32_t* __restrict__ ftable,
                   float* __restrict__ rho, float* __restrict__ prs, float4* __restrict__ pv,
                   uint32_t* __restrict__ pool, uint32_t cap, StepConsts c) {
  extern __shared__ uint16_t hit[];  // (cap + kUDump) rows of 2 * kQB columns
  const uint32_t q = blockIdx.x * kQB + threadIdx.x;
  const uint32_t i0 = 4u * q;
  if (i0 >= c.n) return;
  const PairLists L = pair_lists(pool, c.n);
  uint16_t* const col0 = hit + 2u * threadIdx.x;
  uint32_t fx[4], fy[4], fz[4];
  bool quad = i0 + 3u < c.n;
  if (quad) {
#pragma unroll
    for (int w = 0; w < 4; ++w) fine_coords_of(__ldg(keys + i0 + w), c, fx[w], fy[w], fz[w]);
    quad = fy[0] == fy[3] && fz[0] == fz[3] && fy[1] == fy[0] && fy[2] == fy[0] && fz[1] == fz[0] && fz[2] == fz[0] &&
           fx[3] - fx[0] <= kSpanTogether;
  }
  if (!quad) {  // two pair walks (rare: out of line, one copy)
#pragma unroll 1
    for (uint32_t pr = 0; pr < 2u; ++pr)
      density_pair_outlined(2u * q + pr, col0 + pr, pairs, vel, keys, ftable, rho, prs, pv, L, cap, c);
    return;
  }
  const Rec8 o0 = ldg256(pairs + i0), o1 = ldg256(pairs + i0 + 2u);
  const float2 qx[2] = {o0.a, o1.a}, qy[2] = {o0.b, o1.b}, qz[2] = {o0.c, o1.c};
  CellFrac f[4];
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    const float4 p = make_float4((w & 1) ? qx[w >> 1].y : qx[w >> 1].x, (w & 1) ? qy[w >> 1].y : qy[w >> 1].x,
                                 (w & 1) ? qz[w >> 1].y : qz[w >> 1].x, 0.f);
    f[w] = cell_frac(p, fx[w], fy[w], fz[w], c);
  }
  const float nan = __int_as_float(0x7fc00000);
  const uint32_t base0 = (uint32_t)__cvta_generic_to_shared(col0), base1 = base0 + 2u;
  const uint32_t end0 = base0 + cap * kQStride, end1 = base1 + cap * kQStride;
  uint32_t cur0 = base0, cur1 = base1;
  float sum[4] = {0.f, 0.f, 0.f, 0.f};
  bool unusable = false;
  const uint32_t t0 = 2u * q, t1 = t0 + 1u;
  uint32_t rowid = 0;
#pragma unroll 1
  for (int dz = -1; dz <= 1; ++dz, rowid += 3u) {
    const uint32_t z = fz[0] + dz;
    if (z >= c.fnz) {
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        L.start[(size_t)(rowid + k) * L.np + t0] = 0;
        L.start[(size_t)(rowid + k) * L.np + t1] = 0;
      }
      continue;
    }
    uint32_t rs[3], re[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int dy = k - 1;
      const uint32_t y = fy[0] + dy;
      uint32_t x_lo = 0xFFFFFFFFu, x_hi = 0;
      bool any = false;
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const Reach R = row_reach(c, fx[w], f[w], dy, dz);
        if (R.on) { any = true; x_lo = min(x_lo, R.lo); x_hi = max(x_hi, R.hi); }
      }
      rs[k] = re[k] = 0;
      if (y < c.fny && any) {
        const uint32_t rowbase = (z * c.fny + y) * c.fnx;
        rs[k] = __ldg(ftable + rowbase + x_lo);
        re[k] = __ldg(ftable + rowbase + x_hi + 1u);
      }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const uint32_t s = rs[k], e = re[k];
      L.start[(size_t)(rowid + k) * L.np + t0] = s;
      L.start[(size_t)(rowid + k) * L.np + t1] = s;
      if (e - s > kMaxRun) unusable = true;
      const uint32_t j_end = s < e ? e : s & ~1u;
      uint32_t code = ((rowid + k) << kRowShift) - (s & 1u);
#pragma unroll 1
      for (uint32_t j = s & ~1u; j < j_end; j += 2u, code += 2u) {
        Rec8 r = ldg256(pairs + j);
        if (j < s) r.a.x = nan;
        if (j + 1u >= e) r.a.y = nan;
        density_test_record_quad<kQStride>(qx, qy, qz, -c.h2, r.a, r.b, r.c, sum, cur0, cur1, code);
        cur0 = min(cur0, end0);
        cur1 = min(cur1, end1);
      }
    }
  }
#pragma unroll
  for (int u = 0; u < 3; ++u) {  // both lists end on a multiple of four entries
    if (((cur0 - base0) / kQStride) & 3u) {
      asm volatile("st.shared.u16 [%0], %1;" ::"r"(cur0), "h"((uint16_t)kNullCode) : "memory");
      cur0 += kQStride;
    }
    if (((cur1 - base1) / kQStride) & 3u) {
      asm volatile("st.shared.u16 [%0], %1;" ::"r"(cur1), "h"((uint16_t)kNullCode) : "memory");
      cur1 += kQStride;
    }
  }
  cur0 = min(cur0, end0);
  cur1 = min(cur1, end1);
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    const uint32_t i = i0 + w;
    const float density = (c.mass * c.c_poly6) * -sum[w];
    const float pressure = tait_pressure(density, c);
    rho[i] = density;
    prs[i] = pressure;
    const float4 v = __ldg(vel + i);
    pv[2u * i] = make_float4((w & 1) ? qx[w >> 1].y : qx[w >> 1].x, (w & 1) ? qy[w >> 1].y : qy[w >> 1].x,
                             (w & 1) ? qz[w >> 1].y : qz[w >> 1].x, pressure / (density * density));
    pv[2u * i + 1u] = make_float4(v.x, v.y, v.z, c.mass / density);
  }
#pragma unroll
  for (int pr = 0; pr < 2; ++pr) {
    const uint32_t t = pr ? t1 : t0;
    const uint32_t cnt = ((pr ? cur1 : cur0) - (pr ? base1 : base0)) / kQStride;
    const bool bad = unusable || cnt >= cap;
    L.meta[t] = bad ? kMetaUnusable : (cnt | (cnt << 8));
    if (bad) {
      L.queue[atomicAdd(L.qcount, 1u)] = t;
    } else {
      const uint16_t* col = col0 + pr;
      uint32_t* out = L.codes + t;
      for (uint32_t k = 0; k < cnt; k += 2u)
        out[(size_t)(k >> 1) * L.np] = (uint32_t)col[k * (kQStride / 2u)] | ((uint32_t)col[(k + 1u) * (kQStride / 2u)] << 16);
    }
  }
}

// ---- forces from the pair list ---------------------------------------------------------------
struct ForceAcc2 {
  float2 px, py, pz;  // lane x: particle a, lane y: particle b (sums as in ForceAcc)
  float2 vx, vy, vz;
  float2 nx, ny, nz;
  float2 lap;
  float quirk_a, quirk_b;
};

__device__ __forceinline__ float rsqrt_fast(float x) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

struct PairQuery {
  float2 x, y, z, w;     // positions, p/rho^2
  float2 vx, vy, vz;     // velocities
};

// One listed neighbour j against both particles of the thread.  Nothing is skipped: a particle
// with r >= h (and a masked one, parked at 1e18) gets (h - r) = (h^2 - r^2) = 0 and so exactly
// zero from every term; the particle itself (r = 0) gets zero from the pressure, viscosity and
// normal terms (x_i - x_j = v_j - v_i = 0) and its own term of the colour-field laplacian, which
// the reference also sums (sphb.cl:95-106).
__device__ __forceinline__ void force_pair2(ForceAcc2& A, const PairQuery& q, const float4& pj,
                                            const float4& vj, uint32_t j, uint32_t ia, uint32_t ib,
                                            const StepConsts& c) {
  const float2 dx = __fadd2_rn(q.x, bcast2(-pj.x));
  const float2 dy = __fadd2_rn(q.y, bcast2(-pj.y));
  const float2 dz = __fadd2_rn(q.z, bcast2(-pj.z));
  float2 r2 = __fmul2_rn(dx, dx);
  r2 = __ffma2_rn(dy, dy, r2);
  r2 = __ffma2_rn(dz, dz, r2);
  const float2 r2c = make_float2(fmaxf(r2.x, 1.0e-30f), fmaxf(r2.y, 1.0e-30f));
  float2 inv_r = make_float2(rsqrt_fast(r2c.x), rsqrt_fast(r2c.y));
  const float2 pp = __fadd2_rn(q.w, bcast2(pj.w));
  if (fminf(r2.x, r2.y) < 2.0e-14f) {
    // smoothing.cl:21-25: for |x_i - x_j| < 1e-7 (and j != i) spiky_gradient returns the SCALAR
    // -45/(pi h^6) broadcast to xyz, ungated.  Decided with the reference's own test; reached
    // for the two self pairs of the thread and for coincident particles only.
    if (r2.x < 2.0e-14f && j != ia) {
      const float r = sqrtf(r2.x);
      if ((r - 1.0e-7f < 0.f) && (0.f < r + 1.0e-7f)) { A.quirk_a += pp.x; inv_r.x = 0.f; }
    }
    if (r2.y < 2.0e-14f && j != ib) {
      const float r = sqrtf(r2.y);
      if ((r - 1.0e-7f < 0.f) && (0.f < r + 1.0e-7f)) { A.quirk_b += pp.y; inv_r.y = 0.f; }
    }
  }
  const float2 r = __fmul2_rn(r2c, inv_r);
  float2 hr = __fadd2_rn(bcast2(c.h), make_float2(-r.x, -r.y));
  hr = make_float2(fmaxf(hr.x, 0.f), fmaxf(hr.y, 0.f));
  // pressure: (p_j/rho_j^2 + p_i/rho_i^2) (h - r)^2 (x_i - x_j) / r
  const float2 sp = __fmul2_rn(__fmul2_rn(pp, hr), __fmul2_rn(hr, inv_r));
  A.px = __ffma2_rn(sp, dx, A.px);
  A.py = __ffma2_rn(sp, dy, A.py);
  A.pz = __ffma2_rn(sp, dz, A.pz);
  // viscosity: (v_j - v_i) (m/rho_j) (h - r)
  const float2 vw = __fmul2_rn(bcast2(vj.w), hr);
  A.vx = __ffma2_rn(__fadd2_rn(bcast2(vj.x), make_float2(-q.vx.x, -q.vx.y)), vw, A.vx);
  A.vy = __ffma2_rn(__fadd2_rn(bcast2(vj.y), make_float2(-q.vy.x, -q.vy.y)), vw, A.vy);
  A.vz = __ffma2_rn(__fadd2_rn(bcast2(vj.z), make_float2(-q.vz.x, -q.vz.y)), vw, A.vz);
  // colour field: normal (m/rho_j)(h^2 - r^2)^2 (x_i - x_j), laplacian (m/rho_j)(h^2 - r^2)(3h^2 - 7r^2)
  float2 tt = __fadd2_rn(bcast2(c.h2), make_float2(-r2.x, -r2.y));
  tt = make_float2(fmaxf(tt.x, 0.f), fmaxf(tt.y, 0.f));
  const float2 mt = __fmul2_rn(bcast2(vj.w), tt);
  const float2 nw = __fmul2_rn(mt, tt);
  A.nx = __ffma2_rn(nw, dx, A.nx);
  A.ny = __ffma2_rn(nw, dy, A.ny);
  A.nz = __ffma2_rn(nw, dz, A.nz);
  A.lap = __ffma2_rn(mt, __ffma2_rn(r2, bcast2(-7.f), bcast2(3.f * c.h2)), A.lap);
}

// Tuning of the forces pass, measured on the B200 at 16.7 M particles:
// two neighbours per trip and 80 registers (12 CTAs of 64 threads per SM) 1.18 ms; four per trip
// at 123 registers 1.23 ms, at 96 registers (spills) 1.44 ms; a double-buffered software pipeline
// of the gathers 1.31 ms at best (the register copies, or the second register set, cost more than
// the latency they hide); CTAs of 128 threads 1.43-1.53 ms.
#ifndef SPHB_FORCES_TRIP
#define SPHB_FORCES_TRIP 2
#endif
#ifndef SPHB_FORCES_BLOCK
#define SPHB_FORCES_BLOCK 64
#endif
#ifndef SPHB_FORCES_MINB
#define SPHB_FORCES_MINB 12
#endif
constexpr int kTrip = SPHB_FORCES_TRIP;  // neighbours per trip (2 or 4): their gathers are issued together
static_assert(kTrip == 2 || kTrip == 4, "a trip is one or two words of codes");
constexpr int kFB = SPHB_FORCES_BLOCK;   // threads per CTA
constexpr float kFar = 1.0e18f;

template <bool Q>
__global__ void __launch_bounds__(kFB, SPHB_FORCES_MINB)
    k_forces_union(const float4* __restrict__ pv, const uint32_t* __restrict__ keys,
                   const uint32_t* __restrict__ ftable, const float* __restrict__ rho,
                   const uint32_t* __restrict__ pool, float4* __restrict__ acc, StepConsts c_arg,
                   const uint32_t* __restrict__ ghost_ids) {
  if (Q && !g_hot.active) return;
  const StepConsts& c = Q ? g_hot.c : c_arg;
  __shared__ uint32_t start[kStartRows][kFB];
  const uint32_t t = blockIdx.x * kFB + threadIdx.x;
  const uint32_t ia = 2u * t;
  if (ia >= c.n) return;
  const PairLists L = pair_lists(const_cast<uint32_t*>(pool), c.n);
  const bool has_b = ia + 1 < c.n;
  const uint32_t ib = has_b ? ia + 1 : ia;
  // slab mode: another rank's particles get no forces
  const bool ghost_a = ghost_ids && (__ldg(ghost_ids + ia) & 0x80000000u);
  const bool ghost_b = ghost_ids && (__ldg(ghost_ids + ib) & 0x80000000u);
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  const uint32_t meta = __ldg(L.meta + t);
  if (meta & kMetaUnusable) return;  // no list: the pair is in the queue of k_forces_overflow
  const uint32_t cnt_a = meta & 255u, cnt = (meta >> 8) & 255u;
  const bool split = cnt > cnt_a;
#pragma unroll
  for (int r = 0; r < 9; ++r) start[r][threadIdx.x] = __ldg(L.start + (size_t)r * L.np + t);
  if (split) {
#pragma unroll
    for (int r = 9; r < 18; ++r) start[r][threadIdx.x] = __ldg(L.start + (size_t)r * L.np + t);
  }
  const float4 pa = __ldg(pv + 2u * ia), va = __ldg(pv + 2u * ia + 1u);
  const float4 pb = __ldg(pv + 2u * ib), vb = __ldg(pv + 2u * ib + 1u);
  PairQuery q;
  q.w = make_float2(pa.w, pb.w);
  q.vx = make_float2(va.x, vb.x); q.vy = make_float2(va.y, vb.y); q.vz = make_float2(va.z, vb.z);
  ForceAcc2 A = {};
  const uint32_t* mine = L.codes + t;
#pragma unroll 1
  for (int part = 0; part < 2; ++part) {
    const uint32_t k0 = part ? cnt_a : 0u, k1 = part ? cnt : cnt_a;
    if (k0 >= k1) continue;
    // a particle that is not part of this walk is parked far away: it collects exact zeros
    const bool on_a = part == 0, on_b = has_b && (split ? part == 1 : true);
    q.x = make_float2(on_a ? pa.x : kFar, on_b ? pb.x : kFar);
    q.y = make_float2(on_a ? pa.y : kFar, on_b ? pb.y : kFar);
    q.z = make_float2(on_a ? pa.z : kFar, on_b ? pb.z : kFar);
    constexpr int W = kTrip / 2;  // words of codes per trip
    const uint32_t null_word = kNullCode | (kNullCode << 16);
    uint32_t next[W];
#pragma unroll
    for (int u = 0; u < W; ++u) next[u] = __ldg(mine + (size_t)((k0 >> 1) + u) * L.np);
    for (uint32_t k = k0; k < k1; k += kTrip) {  // (every part holds a multiple of 4 entries)
      uint32_t word[W];
#pragma unroll
      for (int u = 0; u < W; ++u) {
        // the next trip's codes are requested before this trip's gathers are
        word[u] = next[u];
        const uint32_t kn = k + kTrip + 2u * u;
        next[u] = kn < k1 ? __ldg(mine + (size_t)(kn >> 1) * L.np) : null_word;
      }
      uint32_t j[kTrip];
      float4 pj[kTrip], vj[kTrip];
#pragma unroll
      for (int u = 0; u < kTrip; ++u) {
        const uint32_t code = (word[u >> 1] >> (16 * (u & 1))) & 0xffffu;
        // a null entry (padding) reads the thread's own particle and is moved out of reach below
        j[u] = code == kNullCode ? 0xffffffffu : start[code >> kRowShift][threadIdx.x] + (code & kOffMask);
        const uint32_t jj = j[u] == 0xffffffffu ? ia : j[u];
        const Rec8 r = ldg256(pv + 2u * (size_t)jj);  // one gather per neighbour
        pj[u] = make_float4(r.a.x, r.a.y, r.b.x, r.b.y);
        vj[u] = make_float4(r.c.x, r.c.y, r.d.x, r.d.y);
      }
#pragma unroll
      for (int u = 0; u < kTrip; ++u) {
        if (j[u] == 0xffffffffu) pj[u].x = -kFar;  // out of reach of both, parked or not
        force_pair2(A, q, pj[u], vj[u], j[u], ia, has_b ? ib : 0xfffffffeu, c);
      }
    }
  }
  ForceAcc Aa, Ab;
  Aa.px = A.px.x; Aa.py = A.py.x; Aa.pz = A.pz.x; Aa.vx = A.vx.x; Aa.vy = A.vy.x; Aa.vz = A.vz.x;
  Aa.nx = A.nx.x; Aa.ny = A.ny.x; Aa.nz = A.nz.x; Aa.lap = A.lap.x; Aa.quirk = A.quirk_a;
  Ab.px = A.px.y; Ab.py = A.py.y; Ab.pz = A.pz.y; Ab.vx = A.vx.y; Ab.vy = A.vy.y; Ab.vz = A.vz.y;
  Ab.nx = A.nx.y; Ab.ny = A.ny.y; Ab.nz = A.nz.y; Ab.lap = A.lap.y; Ab.quirk = A.quirk_b;
  acc[ia] = ghost_a ? zero4 : force_finish(Aa, __ldg(rho + ia), va.w, c, true);
  if (has_b) acc[ib] = ghost_b ? zero4 : force_finish(Ab, __ldg(rho + ib), vb.w, c, true);
}

// ---- experiment (sphb_set_option("neighbor_kernel", 2 + log2 G): 4, 5, 6 for G = 4, 8, 16) ----------
// Sub-warp-cooperative density: G lanes share ONE particle (a warp = 32 / G consecutive particles).
// Lane g plans rows g, g + G, ...; the G lanes stride each row's run with consecutive 16-byte
// loads (8 lanes = one 128-byte line), the row's hits come out of __ballot_sync as a bit mask per
// (particle, row) — stored as the neighbour list, 9 words per particle — and the partial sums are
// reduced with __shfl_xor_sync.  Built to be MEASURED against the pair kernels above (DESIGN.md
// section 3): every candidate load serves ONE distance test instead of four, nothing is packed, and
// a run of ~8 candidates fills a group's lanes only partly.  The forces pass that goes with it
// searches on its own (k_forces_list).
template <int G>
__global__ void __launch_bounds__(128)
    k_density_coop(float4* pos, float4* vel, const uint32_t* __restrict__ keys,
                   const uint32_t* __restrict__ ftable, float* __restrict__ rho, float* __restrict__ prs,
                   uint32_t* __restrict__ masks, StepConsts c) {
  constexpr uint32_t kPer = 32 / G;  // particles per warp
  constexpr int kOwn = (9 + G - 1) / G;
  const uint32_t lane = threadIdx.x & 31u, g = lane % G, grp = lane / G;
  const uint32_t warp = (blockIdx.x * 128u + threadIdx.x) >> 5;
  const uint32_t i_raw = warp * kPer + grp;
  const bool live = i_raw < c.n;
  const uint32_t i = live ? i_raw : c.n - 1u;  // (idle groups shadow the last particle; nothing is written)
  const float4 pi = pos[i];
  uint32_t fx, fy, fz;
  fine_coords_of(__ldg(keys + i), c, fx, fy, fz);
  const CellFrac f = cell_frac(pi, fx, fy, fz, c);
  uint32_t my_s[kOwn], my_e[kOwn];
#pragma unroll
  for (int q = 0; q < kOwn; ++q) {
    const int row = q * G + (int)g;
    my_s[q] = my_e[q] = 0;
    if (row < 9) row_range(ftable, c, fx, fy, fz, f, row % 3 - 1, row / 3 - 1, my_s[q], my_e[q]);
  }
  float sum = 0.f;
  const uint32_t gshift = grp * G;
#pragma unroll
  for (int row = 0; row < 9; ++row) {
    const int owner = (int)(grp * G) + row % G;
    const uint32_t s = __shfl_sync(0xffffffffu, my_s[row / G], owner),
                   e = __shfl_sync(0xffffffffu, my_e[row / G], owner);
    uint32_t rowmask = 0, shift = 0;
    for (uint32_t j0 = s; __any_sync(0xffffffffu, j0 < e); j0 += G, shift += G) {
      const uint32_t j = j0 + g;
      bool hit = false;
      if (j < e) {
        const float4 pj = __ldg(reinterpret_cast<const float4*>(pos) + j);
        const float dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
        const float u = c.h2 - (dx * dx + dy * dy + dz * dz);
        hit = u > 0.f;
        if (hit) sum += u * u * u;
      }
      const uint32_t b = __ballot_sync(0xffffffffu, hit);
      if (shift < 32u) rowmask |= ((b >> gshift) & ((1u << G) - 1u)) << shift;
    }
    if (live && g == 0) masks[(size_t)row * c.n + i] = rowmask;
  }
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (live && g == 0) {
    const float density = (c.mass * c.c_poly6) * sum;
    const float pressure = tait_pressure(density, c);
    rho[i] = density;
    prs[i] = pressure;
    reinterpret_cast<float*>(pos + i)[3] = pressure / (density * density);
    reinterpret_cast<float*>(vel + i)[3] = c.mass / density;
  }
}

// The particles of the queued pairs (their lists overflowed: compressed fluid), each by its own
// search of the 9 row runs; a fixed, small grid strides over the queue, whose length only the
// device knows.
constexpr int kOverflowBlocks = 592, kOverflowThreads = 64;
template <bool Q>
__global__ void __launch_bounds__(kOverflowThreads)
    k_forces_overflow(const float4* __restrict__ pv, const uint32_t* __restrict__ keys,
                      const uint32_t* __restrict__ ftable, const float* __restrict__ rho,
                      const uint32_t* __restrict__ pool, float4* __restrict__ acc, StepConsts c_arg,
                      const uint32_t* __restrict__ ghost_ids) {
  if (Q && !g_hot.active) return;
  const StepConsts& c = Q ? g_hot.c : c_arg;
  const PairLists L = pair_lists(const_cast<uint32_t*>(pool), c.n);
  const uint32_t total = 2u * min(__ldg(L.qcount), L.np);
  for (uint32_t k = blockIdx.x * kOverflowThreads + threadIdx.x; k < total; k += kOverflowBlocks * kOverflowThreads) {
    const uint32_t i = 2u * __ldg(L.queue + (k >> 1)) + (k & 1u);
    if (i >= c.n) continue;
    const bool ghost = ghost_ids && (__ldg(ghost_ids + i) & 0x80000000u);
    acc[i] = ghost ? make_float4(0.f, 0.f, 0.f, 0.f) : forces_by_search<2>(i, pv, pv + 1, keys, ftable, rho, c);
  }
}

inline unsigned blocks_for(uint32_t n) { return (n + kBlock - 1) / kBlock; }
inline unsigned pair_blocks(uint32_t n, int block = kUB) { return ((n + 1u) / 2u + block - 1) / block; }
inline size_t union_smem(uint32_t cap) { return (size_t)(cap + kUDump) * kUB * sizeof(uint16_t); }

// round-1 layout inside the same pool: 64 + 2 u16 per particle, then one count byte per particle
inline uint16_t* v1_list(uint32_t* pool) { return reinterpret_cast<uint16_t*>(pool); }
inline uint8_t* v1_count(uint32_t* pool, uint32_t n) {
  return reinterpret_cast<uint8_t*>(reinterpret_cast<uint16_t*>(pool) + (size_t)(kListCap + 2) * n);
}

}  // namespace

// 32-bit words of the list pool for n particles with room for `cap` entries per pair
// (also enough for the round-1 kernels' per-particle lists)
size_t neighbor_pool_words(uint32_t n, uint32_t cap) {
  const size_t np = ((size_t)n + 1) / 2;
  const size_t pairs = np * (2 + kStartRows + cap / 2) + 4;
  const size_t v1 = ((size_t)(kListCap + 2) * n * 2 + n + 3) / 4;
  return (pairs > v1 ? pairs : v1) + 4;
}

// Entries a pair's list should have room for, from the particles expected inside one support
// sphere at rest density (4/3 pi h^3 rho0 / m): a pair's union is ~1.5 spheres; the margin absorbs
// compression.  64 keeps the hit columns at 9 KB per CTA; denser fluids (mucus) take 128.
uint32_t neighbor_list_cap(float h, float rho0, float mass) {
  if (const char* env = getenv("SPHB_LIST_CAP")) {  // experiments (tools/ab_neighbors.py): 8 ... 248, a multiple of 8
    const int v = atoi(env);
    if (v >= 8 && v <= 248) return (uint32_t)(v & ~7);
  }
  const float in_sphere = 4.18879f * h * h * h * rho0 / mass;
  return 1.5f * in_sphere * 1.6f <= 64.f ? 64u : 128u;
}

int launch_density(float4* pos, float4* vel, const float4* pairs, float4* pv, const uint32_t* keys,
                   const uint32_t* ftable, float* rho, float* prs, uint32_t* pool, uint32_t cap,
                   const StepConsts& c, int variant, cudaStream_t s, bool queued) {
  if (variant >= 4 && variant <= 6 && !queued) {  // experiment: G = 4 / 8 / 16 lanes per particle
    const int G = 1 << (variant - 2);
    const unsigned grid = (unsigned)(((size_t)c.n * G + 127) / 128);
    if (G == 4) k_density_coop<4><<<grid, 128, 0, s>>>(pos, vel, keys, ftable, rho, prs, pool, c);
    else if (G == 8) k_density_coop<8><<<grid, 128, 0, s>>>(pos, vel, keys, ftable, rho, prs, pool, c);
    else k_density_coop<16><<<grid, 128, 0, s>>>(pos, vel, keys, ftable, rho, prs, pool, c);
    return 1;
  }
  if (variant == 7 && !queued) {  // experiment: four particles per thread, the same pair lists
    const unsigned quads = (c.n + 3u) / 4u;
    cudaMemsetAsync(pair_lists(pool, c.n).qcount, 0, 4 * sizeof(uint32_t), s);
    k_density_quad<<<(quads + kQB - 1) / kQB, kQB, (size_t)(cap + kUDump) * 2 * kQB * sizeof(uint16_t), s>>>(
        pairs, vel, keys, ftable, rho, prs, pv, pool, cap, c);
    return 1;
  }
  if (variant != 0) {  // round-1 kernel (A/B; it leaves p/rho^2 and m/rho in the w lanes of pos / vel)
    const unsigned pairs_n = (c.n + 1u) / 2u, grid = (pairs_n + kPairBlock - 1) / kPairBlock;
    uint16_t* nlist = v1_list(pool);
    uint8_t* ncount = v1_count(pool, c.n);
    if (queued) k_density_pair<true><<<grid, kPairBlock, 0, s>>>(pos, vel, keys, ftable, rho, prs, nlist, ncount, c);
    else k_density_pair<false><<<grid, kPairBlock, 0, s>>>(pos, vel, keys, ftable, rho, prs, nlist, ncount, c);
    return 1;
  }
  cudaMemsetAsync(pair_lists(pool, c.n).qcount, 0, 4 * sizeof(uint32_t), s);  // (c.n: the launch-time count, also when queued)
  if (queued)
    k_density_union<true><<<pair_blocks(c.n), kUB, union_smem(cap), s>>>(pairs, vel, keys, ftable, rho, prs, pv, pool, cap, c);
  else
    k_density_union<false><<<pair_blocks(c.n), kUB, union_smem(cap), s>>>(pairs, vel, keys, ftable, rho, prs, pv, pool, cap, c);
  return 1;
}

// What the density pass recorded, for measurement (sphb_neighbor_list_stats): out[0] = pairs whose
// list is unusable (overflow: their particles search on their own in the forces pass), out[1] =
// split pairs, out[2] = list entries over all pairs (padding included), out[3] = longest list.
__global__ void __launch_bounds__(256) k_list_stats(const uint32_t* __restrict__ meta, uint32_t np,
                                                    unsigned long long* __restrict__ out) {
  const uint32_t t = blockIdx.x * 256u + threadIdx.x;
  uint32_t bad = 0, split = 0, entries = 0;
  if (t < np) {
    const uint32_t m = __ldg(meta + t);
    bad = (m & kMetaUnusable) ? 1u : 0u;
    if (!bad) {
      entries = (m >> 8) & 255u;
      split = entries > (m & 255u) ? 1u : 0u;
    }
  }
  const uint32_t nb = __reduce_add_sync(0xffffffffu, bad), ns = __reduce_add_sync(0xffffffffu, split),
                 ne = __reduce_add_sync(0xffffffffu, entries), mx = __reduce_max_sync(0xffffffffu, entries);
  if ((threadIdx.x & 31u) == 0) {
    if (nb) atomicAdd(out + 0, (unsigned long long)nb);
    if (ns) atomicAdd(out + 1, (unsigned long long)ns);
    if (ne) atomicAdd(out + 2, (unsigned long long)ne);
    atomicMax(out + 3, (unsigned long long)mx);
  }
}

int launch_list_stats(const uint32_t* pool, uint32_t n, unsigned long long* out_dev, cudaStream_t s) {
  const uint32_t np = (n + 1u) / 2u;
  cudaMemsetAsync(out_dev, 0, 4 * sizeof(unsigned long long), s);
  k_list_stats<<<(np + 255u) / 256u, 256, 0, s>>>(pool, np, out_dev);
  return 1;
}

int queue_refresh_neighbors(const QueueCtl* ctl, cudaStream_t s) {
  // the symbol's address is looked up once per device: a plain device-to-device copy after that
  static thread_local void* addr[64] = {};
  int dev = 0;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return (int)err;
  void* local = nullptr;
  void*& slot = (dev >= 0 && dev < 64) ? addr[dev] : local;
  if (!slot) {
    err = cudaGetSymbolAddress(&slot, g_hot);
    if (err != cudaSuccess) return (int)err;
  }
  return (int)cudaMemcpyAsync(slot, &ctl->hot, sizeof(QueueHot), cudaMemcpyDeviceToDevice, s);
}

int launch_forces(const float4* pos, const float4* vel, const float4* pv, const uint32_t* keys,
                  const uint32_t* ftable, const float* rho, const uint32_t* pool, float4* acc,
                  const StepConsts& c, int variant, cudaStream_t s, const uint32_t* ghost_ids,
                  bool queued) {
  if ((variant == 1 || (variant >= 4 && variant <= 6)) && !ghost_ids && !queued) {  // forces by their own search
    k_forces_list<<<blocks_for(c.n), kBlock, 0, s>>>(pos, vel, keys, ftable, rho, acc, c);
    return 1;
  }
  if (variant != 0 && variant != 7) {  // round-1 kernel, from the round-1 density pass's per-particle lists
    const uint16_t* nlist = v1_list(const_cast<uint32_t*>(pool));
    const uint8_t* ncount = v1_count(const_cast<uint32_t*>(pool), c.n);
    if (queued)
      k_forces_nlist<true><<<blocks_for(c.n), kBlock, 0, s>>>(pos, vel, keys, ftable, rho, nlist, ncount, acc, c, ghost_ids);
    else
      k_forces_nlist<false><<<blocks_for(c.n), kBlock, 0, s>>>(pos, vel, keys, ftable, rho, nlist, ncount, acc, c, ghost_ids);
    return 1;
  }
  if (queued) {
    k_forces_union<true><<<pair_blocks(c.n, kFB), kFB, 0, s>>>(pv, keys, ftable, rho, pool, acc, c, ghost_ids);
    k_forces_overflow<true><<<kOverflowBlocks, kOverflowThreads, 0, s>>>(pv, keys, ftable, rho, pool, acc, c, ghost_ids);
  } else {
    k_forces_union<false><<<pair_blocks(c.n, kFB), kFB, 0, s>>>(pv, keys, ftable, rho, pool, acc, c, ghost_ids);
    k_forces_overflow<false><<<kOverflowBlocks, kOverflowThreads, 0, s>>>(pv, keys, ftable, rho, pool, acc, c, ghost_ids);
  }
  return 2;
}

}  // namespace sphb
