// Helpers shared by the neighbour passes (kernels_neighbors.cu and the round-1 kernels kept for
// A/B in neighbors_v1.cuh): the search-grid geometry of a particle and the pair terms of the
// forces pass.  Included inside namespace sphb { namespace { ... } } of kernels_neighbors.cu.
#pragma once

constexpr int kBlock = 128;

// device-driven stepping (common.cuh: QueueCtl): see kernels_exact.cu
__constant__ QueueHot g_hot;

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

// Where a particle sits inside its fine cell in y and z, in units of h: lo = distance to the
// cell's lower face, hi = distance to its upper face (both clamped to [0, 1]); ux = its x
// coordinate in units of the x sub-cell (h / kXSub), from the grid origin.
struct CellFrac {
  float lo[3], hi[3];
  float ux;
};
__device__ __forceinline__ CellFrac cell_frac(const float4& p, uint32_t fx, uint32_t fy, uint32_t fz,
                                              const StepConsts& c) {
  CellFrac f;
  const float inv_h = 1.f / c.h;
  (void)fx;
  const float u[3] = {0.f, (p.y - c.miny) * inv_h - (float)fy, (p.z - c.minz) * inv_h - (float)fz};
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    f.lo[a] = fminf(fmaxf(u[a], 0.f), 1.f);
    f.hi[a] = 1.f - f.lo[a];
  }
  f.ux = (p.x - c.minx) * (inv_h * (float)kXSub);
  return f;
}

// Which x sub-cells of row (fy + dy, fz + dz) can hold a neighbour of a particle at `f` inside
// cell (fx, fy, fz): the row is visited only if its nearest line is closer than h, and then only
// over the chord of the support sphere at that distance, [x - w, x + w] with
// w = sqrt(h^2 - gy^2 - gz^2), rounded out to whole sub-cells.  The radius carries a slack, so
// that rounding in this arithmetic (or in the sub-cell a candidate was sorted into) can only add
// candidates, never lose a pair with r < h.
constexpr float kReach2 = 1.0f + 1.0e-3f;  // (h * (1 + slack))^2 in units of h^2
struct Reach {
  bool on;
  uint32_t lo, hi;  // first / last sub-cell (inclusive); always lo <= fx <= hi
};
__device__ __forceinline__ Reach row_reach(const StepConsts& c, uint32_t fx, const CellFrac& f, int dy,
                                           int dz) {
  const float gy = dy < 0 ? f.lo[1] : (dy > 0 ? f.hi[1] : 0.f);
  const float gz = dz < 0 ? f.lo[2] : (dz > 0 ? f.hi[2] : 0.f);
  const float q = gy * gy + gz * gz;
  Reach r;
  r.on = q < kReach2;
  const float w = sqrtf(fmaxf(kReach2 - q, 0.f)) * (float)kXSub;
  // float -> unsigned conversion saturates at 0 for negatives and NaN
  r.lo = min((uint32_t)fmaxf(f.ux - w, 0.f), fx);
  r.hi = min(max((uint32_t)fmaxf(f.ux + w, 0.f), fx), c.fnx - 1u);
  return r;
}

// First / one-past-last particle of that run; an empty range when nothing in the row can be.
__device__ __forceinline__ void row_range(const uint32_t* __restrict__ ftable, const StepConsts& c,
                                          uint32_t fx, uint32_t fy, uint32_t fz, const CellFrac& f,
                                          int dy, int dz, uint32_t& s, uint32_t& e) {
  s = e = 0;
  const uint32_t y = fy + dy, z = fz + dz;  // wraps to a huge value below 0
  if (y >= c.fny || z >= c.fnz) return;
  const Reach r = row_reach(c, fx, f, dy, dz);
  if (!r.on) return;
  const uint32_t row = (z * c.fny + y) * c.fnx;
  s = __ldg(ftable + row + r.lo);
  e = __ldg(ftable + row + r.hi + 1u);
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
// `self_summed`: the sums already hold the j = i term (the pair kernels walk the list, which
// holds the particle itself); otherwise it is added here.
__device__ __forceinline__ float4 force_finish(const ForceAcc& A, float rho_i, float mrho_i,
                                               const StepConsts& c, bool self_summed = false) {
  // the self pair contributes only to the colour-field laplacian (r = 0)
  const float lap = c.c_poly6_lap * (A.lap + (self_summed ? 0.f : mrho_i * c.h2 * (3.f * c.h2)));
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

// Forces of one particle by its own search of the 9 row runs (no recorded list): what a thread
// falls back to when the density pass could not record its neighbours.
// (S = 1: separate position / velocity arrays; S = 2: the interleaved records pv, pos = pv, vel = pv + 1)
template <int S>
__device__ __noinline__ float4 forces_by_search(uint32_t i, const float4* __restrict__ pos,
                                                const float4* __restrict__ vel,
                                                const uint32_t* __restrict__ keys,
                                                const uint32_t* __restrict__ ftable,
                                                const float* __restrict__ rho, const StepConsts& c) {
  const float4 pi = __ldg(pos + (size_t)S * i), vi = __ldg(vel + (size_t)S * i);
  uint32_t fx, fy, fz;
  fine_coords_of(__ldg(keys + i), c, fx, fy, fz);
  const CellFrac f = cell_frac(pi, fx, fy, fz, c);
  ForceAcc A = {};
#pragma unroll 1
  for (int dz = -1; dz <= 1; ++dz)
#pragma unroll 1
    for (int dy = -1; dy <= 1; ++dy) {
      uint32_t s, e;
      row_range(ftable, c, fx, fy, fz, f, dy, dz, s, e);
      for (uint32_t j = s; j < e; ++j) {
        const float4 pj = __ldg(pos + (size_t)S * j);
        const float dx = pi.x - pj.x, dy_ = pi.y - pj.y, dz_ = pi.z - pj.z;
        const float r2 = dx * dx + dy_ * dy_ + dz_ * dz_;
        if (r2 < c.h2 && j != i) force_pair(A, pi, vi, pj, __ldg(vel + (size_t)S * j), dx, dy_, dz_, r2, c);
      }
    }
  return force_finish(A, __ldg(rho + i), vi.w, c);
}
