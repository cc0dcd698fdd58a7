// The two 27-cell neighbour passes of the SPH step: density + pressure, and forces.
//
// Reference: libclsph/kernels/sphb.cl:3-35 (density_pressure -> forces.cl:14-42 ->
// smoothing.cl:1-4) and sphb.cl:37-126 (forces, with smoothing.cl:6-34).  Both visit, for
// every particle, the 27 cells of side 2h around its own cell, although the smoothing
// kernels vanish beyond h: ~1000 candidate pairs per particle for ~20-40 true neighbours.
// A pair with r >= h contributes exactly zero to every sum, so the kernels here (a) search a
// grid of side h instead — 27 cells of side h hold ~130 candidates — and (b) skip pairs with
// r >= h before doing any work on them; results are unchanged up to summation order.
//
// Particles are sorted by the row-major index of their fine cell, so the three x-adjacent
// cells of each of the 9 (dy, dz) rows are ONE contiguous run [ftable[row], ftable[row + 3]).
//
// This translation unit is compiled with FMA contraction ON and uses rsqrt for 1/r; it is
// checked against the CPU oracle to a stated tolerance (tests/test_gpu_parity.py), unlike
// kernels_exact.cu which is bit-exact.
//
// Data layout (sorted SoA): pos = (x, y, z, p/rho^2), vel = (vx, vy, vz, m/rho).  The two
// w lanes are written by the density pass for the forces pass: the reference recomputes
// p_j/rho_j^2 and m/rho_j for every PAIR (sphb.cl:80-106); they depend only on j.
#include "common.cuh"

namespace sphb {

namespace {

constexpr int kBlock = 128;

// Tait equation, sphb.cl:32-34: K * ((rho/rho0)^7 - 1), pown as repeated multiplication
__device__ __forceinline__ float tait_pressure(float rho, const StepConsts& c) {
  const float q = rho / c.rho0;
  float q7 = q;
#pragma unroll
  for (int k = 1; k < 7; ++k) q7 = q7 * q;
  return c.K * (q7 - 1.f);
}

__device__ __forceinline__ void fine_coords_of(uint32_t key, const StepConsts& c, uint32_t& fx,
                                               uint32_t& fy, uint32_t& fz) {
  fx = key % c.fnx;
  const uint32_t t = key / c.fnx;
  fy = t % c.fny;
  fz = t / c.fny;
}

// Where a particle sits inside its fine cell, in units of h: lo = distance to the cell's lower
// face, hi = distance to its upper face (both clamped to [0, 1]).
struct CellFrac {
  float lo[3], hi[3];
};
__device__ __forceinline__ CellFrac cell_frac(const float4& p, uint32_t fx, uint32_t fy, uint32_t fz,
                                              const StepConsts& c) {
  CellFrac f;
  const float inv_h = 1.f / c.h;
  const float u[3] = {(p.x - c.minx) * inv_h - (float)fx, (p.y - c.miny) * inv_h - (float)fy,
                      (p.z - c.minz) * inv_h - (float)fz};
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    f.lo[a] = fminf(fmaxf(u[a], 0.f), 1.f);
    f.hi[a] = 1.f - f.lo[a];
  }
  return f;
}

// First / one-past-last particle of the x-run of row (fy + dy, fz + dz) that can hold a
// neighbour of a particle at `f` inside cell (fx, fy, fz): a neighbouring cell is visited only
// if its nearest point is closer than h.  An empty range when nothing in the row can be.
constexpr float kReach2 = 1.0f + 1.0e-3f;  // (h * (1 + slack))^2 in units of h^2
__device__ __forceinline__ void row_range(const uint32_t* __restrict__ ftable, const StepConsts& c,
                                          uint32_t fx, uint32_t fy, uint32_t fz, const CellFrac& f,
                                          int dy, int dz, uint32_t& s, uint32_t& e) {
  s = e = 0;
  const uint32_t y = fy + dy, z = fz + dz;  // wraps to a huge value below 0
  if (y >= c.fny || z >= c.fnz) return;
  const float gy = dy < 0 ? f.lo[1] : (dy > 0 ? f.hi[1] : 0.f);
  const float gz = dz < 0 ? f.lo[2] : (dz > 0 ? f.hi[2] : 0.f);
  const float q = gy * gy + gz * gz;
  if (q >= kReach2) return;
  const bool left = fx > 0 && q + f.lo[0] * f.lo[0] < kReach2;
  const bool right = fx + 1 < c.fnx && q + f.hi[0] * f.hi[0] < kReach2;
  const uint32_t row = (z * c.fny + y) * c.fnx + fx;
  s = __ldg(ftable + row - (left ? 1u : 0u));
  e = __ldg(ftable + row + (right ? 2u : 1u));
}

// ---- density + pressure: one thread per particle walking the 9 rows -----------------------
// Besides the density sum, the pass RECORDS which candidates were inside h, so that the forces
// pass does not have to search again: a neighbour is stored as a 16-bit code
// (row 0..8) << 12 | (offset of the neighbour inside that row's run), staged per thread in
// shared memory and flushed k-major (nlist[k * n + i]) so that the stores — and the forces
// pass's loads — are coalesced across the warp.  count 255 = "list unusable" (more than
// kListCap neighbours, or a row run longer than 4096): the forces pass then searches itself.
constexpr int kListCap = 64;
constexpr uint32_t kListOverflow = 255;

__global__ void __launch_bounds__(kBlock)
    k_density_list(float4* pos, float4* vel, const uint32_t* __restrict__ keys,
                   const uint32_t* __restrict__ ftable, float* __restrict__ rho,
                   float* __restrict__ prs, uint16_t* __restrict__ nlist,
                   uint8_t* __restrict__ ncount, StepConsts c) {
  // one column per thread; row kListCap is a dump row that absorbs writes past the capacity
  __shared__ uint16_t hit[(kListCap + 1) * kBlock];
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= c.n) return;
  const float4 pi = pos[i];
  uint32_t fx, fy, fz;
  fine_coords_of(__ldg(keys + i), c, fx, fy, fz);
  const CellFrac f = cell_frac(pi, fx, fy, fz, c);
  float sum = 0.f;
  // The inner loop is branch-free: the cubic term is clamped instead of skipped, and a hit
  // (the particle itself included — the forces pass skips it) is recorded with a predicated
  // store at a cursor that advances only on hits.
  // (shared-window byte addresses, so that the store is one predicated STS)
  const uint32_t hit_base = (uint32_t)__cvta_generic_to_shared(hit) + 2u * threadIdx.x;
  uint32_t cursor = hit_base;
  const uint32_t cursor_end = hit_base + 2u * kListCap * kBlock;
  bool overflow = false;
  uint32_t tag = 0;
#pragma unroll 1
  for (int dz = -1; dz <= 1; ++dz)
#pragma unroll 1
    for (int dy = -1; dy <= 1; ++dy, tag += 4096u) {
      uint32_t s, e;
      row_range(ftable, c, fx, fy, fz, f, dy, dz, s, e);
      if (e - s > 4096u) overflow = true;
      uint32_t code = tag;
#pragma unroll 4
      for (uint32_t j = s; j < e; ++j, ++code) {
        const float4 pj = __ldg(reinterpret_cast<const float4*>(pos) + j);
        const float dx = pi.x - pj.x, dy_ = pi.y - pj.y, dz_ = pi.z - pj.z;
        const float t = c.h2 - (dx * dx + dy_ * dy_ + dz_ * dz_);
        const bool inside = t > 0.f;
        const float tt = fmaxf(t, 0.f);
        sum = fmaf(tt * tt, tt, sum);
        asm volatile("{ .reg .pred p; setp.ne.u32 p, %2, 0; @p st.shared.u16 [%0], %1; }"
                     :: "r"(cursor), "h"((uint16_t)code), "r"((uint32_t)inside) : "memory");
        cursor = min(cursor + (inside ? 2u * kBlock : 0u), cursor_end);
      }
    }
  const float density = (c.mass * c.c_poly6) * sum;
  const float pressure = tait_pressure(density, c);
  rho[i] = density;
  prs[i] = pressure;
  // derived per-particle factors for the forces pass (only the w lanes are written)
  reinterpret_cast<float*>(pos + i)[3] = pressure / (density * density);
  reinterpret_cast<float*>(vel + i)[3] = c.mass / density;
  const uint32_t cnt = (cursor - hit_base) / (2u * kBlock);
  if (cnt >= (uint32_t)kListCap) overflow = true;  // the dump row was reached
  ncount[i] = overflow ? (uint8_t)kListOverflow : (uint8_t)cnt;
  if (!overflow)
    for (uint32_t k = 0; k < cnt; ++k) nlist[(size_t)k * c.n + i] = hit[k * kBlock + threadIdx.x];
}

struct ForceAcc {
  float px, py, pz;  // sum (p_j/rho_j^2 + p_i/rho_i^2) (h-r)^2 (x_i-x_j)/r
  float vx, vy, vz;  // sum (v_j - v_i) (m/rho_j) (h-r)
  float nx, ny, nz;  // sum (m/rho_j) (h^2-r^2)^2 (x_i-x_j)
  float lap;         // sum (m/rho_j) (h^2-r^2)(3h^2-7r^2)
  float quirk;       // sum (p_j/rho_j^2 + p_i/rho_i^2) over pairs with r < 1e-7
};

// One (i, j) pair with j != i and r^2 < h^2.
__device__ __forceinline__ void force_pair(ForceAcc& A, const float4& pi, const float4& vi,
                                           const float4& pj, const float4& vj, float dx, float dy,
                                           float dz, float r2, const StepConsts& c) {
  float r, inv_r;
  bool quirk = false;
  if (r2 < 2.0e-14f) {
    // smoothing.cl:21-25: for |x_i - x_j| < 1e-7 spiky_gradient returns the SCALAR
    // -45/(pi h^6) broadcast to xyz, ungated.  Decide with the reference's own test.
    r = sqrtf(r2);
    quirk = (r - 1.0e-7f < 0.f) && (0.f < r + 1.0e-7f);
    inv_r = quirk ? 0.f : 1.f / r;
  } else {
    inv_r = rsqrtf(r2);
    r = r2 * inv_r;
  }
  const float hr = c.h - r;
  const float pp = pj.w + pi.w;
  if (quirk) {
    A.quirk += pp;
  } else {
    const float sp = pp * hr * hr * inv_r;
    A.px += sp * dx; A.py += sp * dy; A.pz += sp * dz;
  }
  const float vw = vj.w * hr;
  A.vx += (vj.x - vi.x) * vw; A.vy += (vj.y - vi.y) * vw; A.vz += (vj.z - vi.z) * vw;
  const float t = c.h2 - r2;
  const float nw = vj.w * t * t;
  A.nx += nw * dx; A.ny += nw * dy; A.nz += nw * dz;
  A.lap += vj.w * t * (3.f * c.h2 - 7.f * r2);
}

// sphb.cl:108-122: combine the four sums into the acceleration.
__device__ __forceinline__ float4 force_finish(const ForceAcc& A, float rho_i, float mrho_i,
                                               const StepConsts& c) {
  // the self pair contributes only to the colour-field laplacian (r = 0)
  const float lap = c.c_poly6_lap * (A.lap + mrho_i * c.h2 * (3.f * c.h2));
  const float ms = c.mass * c.c_spiky, mq = c.mass * c.spiky_eps * A.quirk;
  const float prx = ms * A.px + mq, pry = ms * A.py + mq, prz = ms * A.pz + mq;
  const float nx = c.c_poly6_grad * A.nx, ny = c.c_poly6_grad * A.ny, nz = c.c_poly6_grad * A.nz;
  const float mv = c.mu * c.c_visc;
  float sx = -rho_i * prx + mv * A.vx, sy = -rho_i * pry + mv * A.vy, sz = -rho_i * prz + mv * A.vz;
  const float nl = sqrtf(nx * nx + ny * ny + nz * nz);
  if (nl > c.st_threshold) {
    const float k = -c.st * lap / nl;
    sx += k * nx; sy += k * ny; sz += k * nz;
  }
  return make_float4(sx / rho_i + c.gx, sy / rho_i + c.gy, sz / rho_i + c.gz, 0.f);
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
__global__ void __launch_bounds__(kBlock)
    k_forces_nlist(const float4* __restrict__ pos, const float4* __restrict__ vel,
                   const uint32_t* __restrict__ keys, const uint32_t* __restrict__ ftable,
                   const float* __restrict__ rho, const uint16_t* __restrict__ nlist,
                   const uint8_t* __restrict__ ncount, float4* __restrict__ acc, StepConsts c,
                   const uint32_t* __restrict__ ghost_ids) {
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
  const CellFrac f = cell_frac(pi, fx, fy, fz, c);
  const uint32_t cnt = ncount[i];
  ForceAcc A = {};
  if (cnt != kListOverflow) {
    int row = 0;
#pragma unroll
    for (int dz = -1; dz <= 1; ++dz)
#pragma unroll
      for (int dy = -1; dy <= 1; ++dy, ++row) {
        uint32_t s, e;
        row_range(ftable, c, fx, fy, fz, f, dy, dz, s, e);
        row_start[row][threadIdx.x] = s;
      }
    // four neighbours per trip: their codes, then their eight float4 loads, are issued together
    // (each is a dependent chain code -> row start -> position / velocity -> ~45 flops)
    const uint16_t* mine = nlist + i;
    for (uint32_t k = 0; k < cnt; k += 4) {
      uint32_t j[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const bool live = k + u < cnt;
        const uint32_t code = live ? mine[(size_t)(k + u) * c.n] : 0u;
        // a dead slot points at the particle itself, which is skipped like its own list entry
        j[u] = live ? row_start[code >> 12][threadIdx.x] + (code & 4095u) : i;
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

inline unsigned blocks_for(uint32_t n) { return (n + kBlock - 1) / kBlock; }

}  // namespace

size_t neighbor_list_entries(uint32_t n) { return (size_t)kListCap * n; }

int launch_density(float4* pos, float4* vel, const uint32_t* keys, const uint32_t* ftable,
                   float* rho, float* prs, uint16_t* nlist, uint8_t* ncount, const StepConsts& c,
                   int variant, cudaStream_t s) {
  (void)variant;  // the list is recorded either way; variant 1 just does not use it
  k_density_list<<<blocks_for(c.n), kBlock, 0, s>>>(pos, vel, keys, ftable, rho, prs, nlist, ncount, c);
  return 1;
}

int launch_forces(const float4* pos, const float4* vel, const uint32_t* keys,
                  const uint32_t* ftable, const float* rho, const uint16_t* nlist,
                  const uint8_t* ncount, float4* acc, const StepConsts& c, int variant,
                  cudaStream_t s, const uint32_t* ghost_ids) {
  if (variant == 1 && !ghost_ids)
    k_forces_list<<<blocks_for(c.n), kBlock, 0, s>>>(pos, vel, keys, ftable, rho, acc, c);
  else
    k_forces_nlist<<<blocks_for(c.n), kBlock, 0, s>>>(pos, vel, keys, ftable, rho, nlist, ncount, acc, c, ghost_ids);
  return 1;
}

}  // namespace sphb
