// The round-1 neighbour passes, kept selectable (sphb_set_option("neighbor_kernel", 2)) so that
// the current kernels can be measured against them on one box: one thread per pair of particles
// walking the runs with a 16-bit code list PER PARTICLE (k_density_pair), and one thread per
// particle evaluating its own list (k_forces_nlist).  k_forces_list (independent search) is
// also the fallback of the current forces pass.  Included inside kernels_neighbors.cu.
#pragma once

// ---- density + pressure: one thread per PAIR of consecutive particles, walking the 9 rows ----
// Two particles that are neighbours in the sorted array sit in the same or in x-adjacent cells
// four times out of five, so their candidate runs coincide: one thread tests every candidate
// against both, which halves the 16-byte candidate loads (the pass is bound by L1 wavefronts
// and issue slots, not by HBM).  A pair that is NOT in the same row / adjacent cells is walked
// in two passes, one particle at a time (the idle one is given a NaN position, which fails
// every distance test).
//
// Besides the density sum, the pass RECORDS which candidates were inside h, so that the forces
// pass does not have to search again: a neighbour is stored as a 16-bit code
// (row 0..8) << 12 | (offset of the neighbour inside the run that was walked for that row),
// staged per particle in a shared-memory column and flushed k-major (nlist[k * n + i]) so that
// the stores — and the forces pass's loads — are coalesced across the warp.  Where the walked
// run began is recorded as kBackBits bits per row (nback: own sub-cell x - first sub-cell x).  count 255 =
// "list unusable" (more than kListCap neighbours, or a run longer than 4096): the forces pass
// then searches itself.
// Where a walked run began is recorded per row as (own sub-cell - first sub-cell) in kBackBits
// bits: at most kTogether (the partner's lead) + kXSub + 1 (the chord, rounded out).
constexpr uint32_t kTogether = kXSub >= 2 ? 2u : 1u;  // sub-cells two particles of a thread may be apart
constexpr uint32_t kBackBits = 3;
static_assert(kTogether + kXSub + 1u < (1u << kBackBits), "run starts do not fit their bit field");
constexpr int kListCap = 64;
constexpr uint32_t kListOverflow = 255;
constexpr int kPairBlock = 64;                  // threads per CTA; 2 particles each
constexpr int kPairCols = 2 * kPairBlock;       // shared-memory columns (one per particle)
constexpr int kDumpRows = 4;                    // rows past the capacity that absorb a group's hits
constexpr uint32_t kColStride = 2u * kPairCols; // bytes between two entries of a column

__device__ __forceinline__ uint32_t* nback_of(const uint16_t* nlist, uint32_t n) {
  return reinterpret_cast<uint32_t*>(const_cast<uint16_t*>(nlist) + (size_t)kListCap * n);
}

// One candidate against the thread's TWO particles at once, in packed fp32x2 arithmetic (FADD2 /
// FFMA2 / FMUL2: Blackwell's fp32 pipe retires two lanes per instruction only in this form; the
// candidate's coordinates enter as broadcast scalar operands).  u = r^2 - h^2 per particle, then,
// where u < 0: sum += u^3 (the caller negates the total), a predicated store of the code and the
// cursor's advance — one predicate per particle drives all three (NaN compares false).
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void density_test2(uint64_t qx, uint64_t qy, uint64_t qz, uint64_t neg_h2,
                                              const float4& pj, float& sum_a, float& sum_b,
                                              uint32_t& cur_a, uint32_t& cur_b, uint32_t code) {
  asm volatile(
      "{ .reg .b64 bx, by, bz, dx, dy, dz, u, u2;\n"
      "  .reg .f32 ua, ub, va, vb;\n"
      "  .reg .pred pa, pb;\n"
      "  mov.b64 bx, {%9, %9};\n"
      "  mov.b64 by, {%10, %10};\n"
      "  mov.b64 bz, {%11, %11};\n"
      "  sub.rn.f32x2 dx, %5, bx;\n"
      "  sub.rn.f32x2 dy, %6, by;\n"
      "  sub.rn.f32x2 dz, %7, bz;\n"
      "  fma.rn.f32x2 u, dx, dx, %8;\n"
      "  fma.rn.f32x2 u, dy, dy, u;\n"
      "  fma.rn.f32x2 u, dz, dz, u;\n"
      "  mul.rn.f32x2 u2, u, u;\n"
      "  mov.b64 {ua, ub}, u;\n"
      "  mov.b64 {va, vb}, u2;\n"
      "  setp.lt.f32 pa, ua, 0f00000000;\n"
      "  setp.lt.f32 pb, ub, 0f00000000;\n"
      "  @pa fma.rn.f32 %0, va, ua, %0;\n"
      "  @pb fma.rn.f32 %1, vb, ub, %1;\n"
      "  @pa st.shared.u16 [%2], %4;\n"
      "  @pb st.shared.u16 [%3], %4;\n"
      "  @pa add.u32 %2, %2, %12;\n"
      "  @pb add.u32 %3, %3, %12; }"
      : "+f"(sum_a), "+f"(sum_b), "+r"(cur_a), "+r"(cur_b)
      : "h"((uint16_t)code), "l"(qx), "l"(qy), "l"(qz), "l"(neg_h2), "f"(pj.x), "f"(pj.y), "f"(pj.z),
        "n"(kColStride)
      : "memory");
}

template <bool Q>
__global__ void __launch_bounds__(kPairBlock)
    k_density_pair(float4* pos, float4* vel, const uint32_t* __restrict__ keys,
                   const uint32_t* __restrict__ ftable, float* __restrict__ rho,
                   float* __restrict__ prs, uint16_t* __restrict__ nlist,
                   uint8_t* __restrict__ ncount, StepConsts c_arg) {
  if (Q && !g_hot.active) return;
  const StepConsts& c = Q ? g_hot.c : c_arg;
  __shared__ uint16_t hit[(kListCap + kDumpRows) * kPairCols];
  const uint32_t ia = 2u * (blockIdx.x * kPairBlock + threadIdx.x);
  if (ia >= c.n) return;
  const bool has_b = ia + 1 < c.n;
  const uint32_t ib = has_b ? ia + 1 : ia;
  const float4 pa = pos[ia], pb = pos[ib];
  uint32_t fxa, fya, fza, fxb, fyb, fzb;
  fine_coords_of(__ldg(keys + ia), c, fxa, fya, fza);
  fine_coords_of(__ldg(keys + ib), c, fxb, fyb, fzb);
  const CellFrac fa = cell_frac(pa, fxa, fya, fza, c), fb = cell_frac(pb, fxb, fyb, fzb, c);
  // sorted by row-major cell: inside a row, b's cell is never left of a's
  const bool together = has_b && fya == fyb && fza == fzb && fxb - fxa <= kTogether;
  const int npass = (has_b && !together) ? 2 : 1;
  const float nan = __int_as_float(0x7fc00000);
  const float4 nowhere = make_float4(nan, nan, nan, 0.f);

  // (shared-window byte addresses, so that a store is one predicated STS)
  const uint32_t base_a = (uint32_t)__cvta_generic_to_shared(hit) + 4u * threadIdx.x, base_b = base_a + 2u;
  const uint32_t end_a = base_a + kListCap * kColStride, end_b = base_b + kListCap * kColStride;
  uint32_t cur_a = base_a, cur_b = base_b, back_a = 0, back_b = 0;
  float sum_a = 0.f, sum_b = 0.f;
  bool overflow = false;
  const float4* cand = reinterpret_cast<const float4*>(pos);
#pragma unroll 1
  for (int pass = 0; pass < npass; ++pass) {
    const bool la = pass == 0, lb = together || pass == 1;
    const float4 qa = la ? pa : nowhere, qb = lb ? pb : nowhere;
    const uint64_t qx = pack2(qa.x, qb.x), qy = pack2(qa.y, qb.y), qz = pack2(qa.z, qb.z);
    const uint64_t neg_h2 = pack2(-c.h2, -c.h2);
    const uint32_t ry = la ? fya : fyb, rz = la ? fza : fzb;
    uint32_t tag = 0, shift = 0;
#pragma unroll 1
    for (int dz = -1; dz <= 1; ++dz, tag += 3u * 4096u, shift += 3u * kBackBits) {
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
          // (a particle that cannot reach the row records nothing there: the union's start may
          // lie beyond its own sub-cell)
          if (ona) back_a |= (fxa - x_lo) << (shift + kBackBits * k);
          if (onb) back_b |= (fxb - x_lo) << (shift + kBackBits * k);
        }
      }
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const uint32_t e = re[k];
        uint32_t j = rs[k], code = tag + 4096u * k;
        if (e - j > 4096u) overflow = true;
        // The loop body is branch-free: the cubic term is clamped instead of skipped, and a hit
        // (the particle itself included — the forces pass skips it) is recorded with a
        // predicated store at a cursor that advances only on hits; the cursors are pulled
        // back into the dump rows once per group of four candidates.
        for (; j + 4u <= e; j += 4u, code += 4u) {
          const float4 p0 = __ldg(cand + j), p1 = __ldg(cand + j + 1), p2 = __ldg(cand + j + 2),
                       p3 = __ldg(cand + j + 3);
          density_test2(qx, qy, qz, neg_h2, p0, sum_a, sum_b, cur_a, cur_b, code);
          density_test2(qx, qy, qz, neg_h2, p1, sum_a, sum_b, cur_a, cur_b, code + 1u);
          density_test2(qx, qy, qz, neg_h2, p2, sum_a, sum_b, cur_a, cur_b, code + 2u);
          density_test2(qx, qy, qz, neg_h2, p3, sum_a, sum_b, cur_a, cur_b, code + 3u);
          cur_a = min(cur_a, end_a);
          cur_b = min(cur_b, end_b);
        }
        for (; j < e; ++j, ++code) {
          const float4 p0 = __ldg(cand + j);
          density_test2(qx, qy, qz, neg_h2, p0, sum_a, sum_b, cur_a, cur_b, code);
          cur_a = min(cur_a, end_a);
          cur_b = min(cur_b, end_b);
        }
      }
    }
  }
  uint32_t* const nback = nback_of(nlist, c.n);
  const uint16_t* col = hit + 2 * threadIdx.x;
  uint32_t cnt2[2];
  bool over2[2];
#pragma unroll
  for (int w = 0; w < 2; ++w) {
    if (w == 1 && !has_b) break;
    const uint32_t i = w ? ib : ia;
    const float density = (c.mass * c.c_poly6) * -(w ? sum_b : sum_a);  // the sums hold -(h^2 - r^2)^3
    const float pressure = tait_pressure(density, c);
    rho[i] = density;
    prs[i] = pressure;
    // derived per-particle factors for the forces pass (only the w lanes are written)
    reinterpret_cast<float*>(pos + i)[3] = pressure / (density * density);
    reinterpret_cast<float*>(vel + i)[3] = c.mass / density;
    cnt2[w] = ((w ? cur_b : cur_a) - (w ? base_b : base_a)) / kColStride;
    over2[w] = overflow || cnt2[w] >= (uint32_t)kListCap;  // the dump rows were reached
    ncount[i] = over2[w] ? (uint8_t)kListOverflow : (uint8_t)cnt2[w];
    nback[i] = w ? back_b : back_a;
  }
  if (has_b && !over2[0] && !over2[1] && (c.n & 1u) == 0u) {
    // both columns of the thread leave as one aligned 32-bit store per entry (entries past the
    // shorter list's count are never read)
    const uint32_t m = max(cnt2[0], cnt2[1]);
    const uint32_t* col32 = reinterpret_cast<const uint32_t*>(col);
    uint32_t* out = reinterpret_cast<uint32_t*>(nlist + ia);
    for (uint32_t k = 0; k < m; ++k) out[(size_t)k * (c.n / 2u)] = col32[k * (kPairCols / 2)];
  } else {
#pragma unroll
    for (int w = 0; w < 2; ++w) {
      if ((w == 1 && !has_b) || over2[w]) continue;
      const uint32_t i = w ? ib : ia;
      for (uint32_t k = 0; k < cnt2[w]; ++k) nlist[(size_t)k * c.n + i] = col[k * kPairCols + w];
    }
  }
}

// Forces in two loops per particle.  Loop A tests the ~100 candidates and only RECORDS the
// ~20-40 with r < h in a per-thread column of shared memory; loop B then evaluates the pair
// terms for recorded neighbours only.  In a single loop nearly every iteration has at least
// one of the 32 lanes inside h, so the whole warp would pay for the ~45-instruction pair
// evaluation on every candidate.
constexpr int kHitCap = 40;
__global__ void __launch_bounds__(kBlock)
    k_forces_list(const float4* __restrict__ pos, const float4* __restrict__ vel,
                  const uint32_t* __restrict__ keys, const uint32_t* __restrict__ ftable,
                  const float* __restrict__ rho, float4* __restrict__ acc, StepConsts c) {
  __shared__ uint32_t hit[kHitCap][kBlock];
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= c.n) return;
  const float4 pi = __ldg(pos + i), vi = __ldg(vel + i);
  uint32_t fx, fy, fz;
  fine_coords_of(__ldg(keys + i), c, fx, fy, fz);
  const CellFrac f = cell_frac(pi, fx, fy, fz, c);
  ForceAcc A = {};
  int cnt = 0;
  auto drain = [&]() {
    for (int k = 0; k < cnt; ++k) {
      const uint32_t j = hit[k][threadIdx.x];
      const float4 pj = __ldg(pos + j);
      const float4 vj = __ldg(vel + j);
      const float dx = pi.x - pj.x, dy_ = pi.y - pj.y, dz_ = pi.z - pj.z;
      const float r2 = dx * dx + dy_ * dy_ + dz_ * dz_;
      force_pair(A, pi, vi, pj, vj, dx, dy_, dz_, r2, c);
    }
    cnt = 0;
  };
#pragma unroll 1
  for (int dz = -1; dz <= 1; ++dz)
#pragma unroll 1
    for (int dy = -1; dy <= 1; ++dy) {
      uint32_t s, e;
      row_range(ftable, c, fx, fy, fz, f, dy, dz, s, e);
      for (uint32_t j = s; j < e; ++j) {
        const float4 pj = __ldg(pos + j);
        const float dx = pi.x - pj.x, dy_ = pi.y - pj.y, dz_ = pi.z - pj.z;
        const float r2 = dx * dx + dy_ * dy_ + dz_ * dz_;
        if (r2 < c.h2 && j != i) {
          hit[cnt][threadIdx.x] = j;
          if (++cnt == kHitCap) drain();
        }
      }
    }
  drain();
  acc[i] = force_finish(A, __ldg(rho + i), vi.w, c);
}

// Forces from the neighbour list the density pass recorded.
// (experiments: -DSPHB_FORCES_MINBLOCKS=<n> asks for n resident CTAs per SM)
#ifdef SPHB_FORCES_MINBLOCKS
#define SPHB_FORCES_BOUNDS __launch_bounds__(kBlock, SPHB_FORCES_MINBLOCKS)
#else
#define SPHB_FORCES_BOUNDS __launch_bounds__(kBlock)
#endif
template <bool Q>
__global__ void SPHB_FORCES_BOUNDS
    k_forces_nlist(const float4* __restrict__ pos, const float4* __restrict__ vel,
                   const uint32_t* __restrict__ keys, const uint32_t* __restrict__ ftable,
                   const float* __restrict__ rho, const uint16_t* __restrict__ nlist,
                   const uint8_t* __restrict__ ncount, float4* __restrict__ acc, StepConsts c_arg,
                   const uint32_t* __restrict__ ghost_ids) {
  if (Q && !g_hot.active) return;
  const StepConsts& c = Q ? g_hot.c : c_arg;
  __shared__ uint32_t row_start[9][kBlock];
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= c.n) return;
  if (ghost_ids && (__ldg(ghost_ids + i) & 0x80000000u)) {  // slab mode: another rank's particle
    acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    return;
  }
  const float4 pi = __ldg(pos + i), vi = __ldg(vel + i);
  uint32_t fx, fy, fz;
  fine_coords_of(__ldg(keys + i), c, fx, fy, fz);
  const uint32_t cnt = ncount[i];
  ForceAcc A = {};
  if (cnt != kListOverflow) {
    // where the run the density pass walked for each row began: kBackBits bits per row
    uint32_t back = __ldg(nback_of(nlist, c.n) + i);
#pragma unroll
    for (int dz = -1; dz <= 1; ++dz)
#pragma unroll
      for (int dy = -1; dy <= 1; ++dy, back >>= kBackBits) {
        const uint32_t y = fy + dy, z = fz + dz;
        uint32_t s = 0;
        if (y < c.fny && z < c.fnz)
          s = __ldg(ftable + ((z * c.fny + y) * c.fnx + fx - (back & ((1u << kBackBits) - 1u))));
        row_start[(dz + 1) * 3 + (dy + 1)][threadIdx.x] = s;
      }
    // four neighbours per trip: their codes, then their eight float4 loads, are issued together
    // (each is a dependent chain code -> row start -> position / velocity -> ~45 flops)
    const uint16_t* mine = nlist + i;
    uint32_t code[4], next[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) next[u] = (uint32_t)u < cnt ? mine[(size_t)u * c.n] : 0u;
    for (uint32_t k = 0; k < cnt; k += 4) {
      uint32_t j[4];
      // the next trip's codes are requested before this trip's gathers are
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        code[u] = next[u];
        next[u] = k + 4 + u < cnt ? mine[(size_t)(k + 4 + u) * c.n] : 0u;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        // a dead slot points at the particle itself, which is skipped like its own list entry
        j[u] = k + u < cnt ? row_start[code[u] >> 12][threadIdx.x] + (code[u] & 4095u) : i;
      }
      float4 pj[4], vj[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        pj[u] = __ldg(pos + j[u]);
        vj[u] = __ldg(vel + j[u]);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (j[u] == i) continue;  // the list holds the particle itself
        const float dx = pi.x - pj[u].x, dy_ = pi.y - pj[u].y, dz_ = pi.z - pj[u].z;
        const float r2 = dx * dx + dy_ * dy_ + dz_ * dz_;
        force_pair(A, pi, vi, pj[u], vj[u], dx, dy_, dz_, r2, c);
      }
    }
  } else {  // the list overflowed: search again
    const CellFrac f = cell_frac(pi, fx, fy, fz, c);
#pragma unroll 1
    for (int dz = -1; dz <= 1; ++dz)
#pragma unroll 1
      for (int dy = -1; dy <= 1; ++dy) {
        uint32_t s, e;
        row_range(ftable, c, fx, fy, fz, f, dy, dz, s, e);
        for (uint32_t j = s; j < e; ++j) {
          const float4 pj = __ldg(pos + j);
          const float dx = pi.x - pj.x, dy_ = pi.y - pj.y, dz_ = pi.z - pj.z;
          const float r2 = dx * dx + dy_ * dy_ + dz_ * dz_;
          if (r2 < c.h2 && j != i) force_pair(A, pi, vi, pj, __ldg(vel + j), dx, dy_, dz_, r2, c);
        }
      }
  }
  acc[i] = force_finish(A, __ldg(rho + i), vi.w, c);
}
