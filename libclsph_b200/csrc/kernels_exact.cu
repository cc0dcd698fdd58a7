// Per-particle streaming kernels of the SPH step: bbox reduce, Morton locate, SoA reorder,
// cell table, advection + distance-field collision, distance-field build, AoS<->SoA.
//
// This translation unit is compiled with -fmad=false (and CUDA's default IEEE '/' and
// sqrt): every expression below evaluates exactly as the reference's expression tree does
// without FMA contraction, so these passes are BIT-IDENTICAL to the CPU checker
// (oracle/sph_oracle.c), not merely within tolerance.  They are all HBM-bound streaming
// passes; the arithmetic is free.
#include "common.cuh"

namespace sphb {

namespace {

constexpr int kBlock = 256;

// Device-driven stepping (common.cuh: QueueCtl): a queued step's kernels (template argument
// Q = true) take their constants from this copy, refreshed per step by queue_refresh_exact, and
// return at once when the step is not to run.  Q = false is the host-driven path: constants by
// value, no checks.
__constant__ QueueHot g_hot;
#define SPHB_Q_PROLOGUE(c)               \
  if (Q && !g_hot.active) return;        \
  const StepConsts& cc = Q ? g_hot.c : (c)

__device__ __forceinline__ float warp_min(float v) {
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-level min/max of up to 8 values per thread, then one atomic per value per block.
// vals[0..2] are minimised into red->min_enc, vals[3..5] maximised into red->max_enc,
// vals[6], vals[7] maximised into max_v2 / max_a2.  `mask` selects which are live.
__device__ __forceinline__ void block_reduce_to(Reduce* red, float (&vals)[8], unsigned mask) {
  __shared__ float sh[8][kBlock / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    if (!(mask & (1u << k))) continue;
    float v = (k < 3) ? warp_min(vals[k]) : warp_max(vals[k]);
    if (lane == 0) sh[k][warp] = v;
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (!(mask & (1u << k))) continue;
      float v = (lane < kBlock / 32) ? sh[k][lane] : (k < 3 ? INFINITY : -INFINITY);
      v = (k < 3) ? warp_min(v) : warp_max(v);
      if (lane == 0) {
        if (k < 3) atomicMin(&red->min_enc[k], enc_float(v));
        else if (k < 6) atomicMax(&red->max_enc[k - 3], enc_float(v));
        else if (k == 6) atomicMax(&red->max_v2, enc_float(v));
        else atomicMax(&red->max_a2, enc_float(v));
      }
    }
  }
}

template <bool Q>
__global__ void k_reduce_init(Reduce* red, const QueueCtl* q) {
  if (Q && !q->attempt) return;
  if (threadIdx.x == 0) {
    for (int a = 0; a < 3; ++a) {
      red->min_enc[a] = enc_float(INFINITY);
      red->max_enc[a] = enc_float(-INFINITY);
    }
    red->max_v2 = enc_float(-INFINITY);
    red->max_a2 = enc_float(-INFINITY);
  }
}

// libclsph/kernels/reduction.cl:2-64 (minimum_pos + maximum_pos in one pass).
__global__ void __launch_bounds__(kBlock) k_bbox(const float4* __restrict__ pos, uint32_t n,
                                                 Reduce* red) {
  float v[8] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY, 0.f, 0.f};
  for (uint32_t i = blockIdx.x * kBlock + threadIdx.x; i < n; i += gridDim.x * kBlock) {
    const float4 p = __ldg(pos + i);
    v[0] = fminf(v[0], p.x); v[1] = fminf(v[1], p.y); v[2] = fminf(v[2], p.z);
    v[3] = fmaxf(v[3], p.x); v[4] = fmaxf(v[4], p.y); v[5] = fmaxf(v[5], p.z);
  }
  block_reduce_to(red, v, 0x3f);
}

// libclsph/kernels/grid.cl:40-64 computes cell = (uint)((p - min) / (h*2)) per axis — that
// expression is evaluated here exactly (IEEE subtract and divide, truncation), because the
// reference's sorted order and cell table are functions of it.  The engine's own search grid
// halves the cell: fine = 2*cell + (the particle lies in the upper half of the cell), and the
// sort key is the ROW-MAJOR index of the fine cell, so that three x-adjacent cells are one
// contiguous run of particles.  (The upper-half test may be off by an ulp at the mid-plane;
// pairs that could be missed because of it have |r| within an ulp of h, where every
// smoothing kernel is zero.)
__device__ __forceinline__ uint32_t fine_coord(float p, float mn, float cell, float h) {
  const float rel = p - mn;
  const uint32_t c = (uint32_t)(rel / cell);
  const float within = rel - (float)c * cell;
  return 2u * c + (within >= h ? 1u : 0u);
}

// Along x the cell of side 2h is cut into kXPerRef sub-cells (common.cuh: kXSub); the reference
// cell is still exactly (fine x) / kXPerRef.
__device__ __forceinline__ uint32_t fine_coord_x(float p, float mn, float cell, float h) {
  const float rel = p - mn;
  const uint32_t c = (uint32_t)(rel / cell);
  const float within = rel - (float)c * cell;
  const uint32_t sub = (uint32_t)(within / (h / (float)kXSub));  // negative / NaN convert to 0
  return kXPerRef * c + min(sub, kXPerRef - 1u);
}

__device__ __forceinline__ uint32_t fine_key_of(const float4& p, const StepConsts& c) {
  uint32_t fx = fine_coord_x(p.x, c.minx, c.cell, c.h);
  uint32_t fy = fine_coord(p.y, c.miny, c.cell, c.h);
  uint32_t fz = fine_coord(p.z, c.minz, c.cell, c.h);
  // NaN / runaway positions: park them in the last cell instead of indexing out of bounds
  fx = min(fx, c.fnx - 1); fy = min(fy, c.fny - 1); fz = min(fz, c.fnz - 1);
  return (fz * c.fny + fy) * c.fnx + fx;
}

__global__ void __launch_bounds__(kBlock) k_locate(const float4* __restrict__ pos,
                                                   uint32_t* __restrict__ keys, StepConsts c) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= c.n) return;
  keys[i] = fine_key_of(__ldg(pos + i), c);
}

// ---- counting sort by search cell ------------------------------------------------------------
// The search grid has about as many cells as there are particles / 4, so ordering particles by
// cell is a COUNTING sort: (1) key + a slot inside the cell from one atomic per (warp, cell),
// (2) exclusive scan of the counts — which IS the cell table —, (3) placement, (4) a
// canonicalisation that orders each cell's members by their previous index, so that the
// result is exactly the stable sort a radix sort would give (atomics decide the slots, so
// without it the order inside a cell would vary from run to run).
constexpr int kLocPer = 4;  // particles per thread of k_locate_count: four atomic round trips in flight
template <bool Q>
__global__ void __launch_bounds__(kBlock)
    k_locate_count(const float4* __restrict__ pos, uint32_t* __restrict__ keys, uint32_t* __restrict__ slot,
                   uint32_t* __restrict__ count, StepConsts c) {
  SPHB_Q_PROLOGUE(c);
  // the kernel is bound by the latency of the atomic's round trip to L2 (ncu: long-scoreboard 16
  // cycles per issue at 85 % occupancy), so every thread carries kLocPer independent particles,
  // kBlock apart (coalesced), through it at once
  const uint32_t i0 = blockIdx.x * (kBlock * kLocPer) + threadIdx.x;
  const int lane = threadIdx.x & 31;
  uint32_t key[kLocPer], base[kLocPer];
  unsigned peers[kLocPer];
#pragma unroll
  for (int u = 0; u < kLocPer; ++u) {
    const uint32_t i = i0 + u * kBlock;
    key[u] = i < cc.n ? fine_key_of(__ldg(pos + i), cc) : 0xFFFFFFFFu;
  }
#pragma unroll
  for (int u = 0; u < kLocPer; ++u) {
    peers[u] = __match_any_sync(0xffffffffu, key[u]);
    base[u] = 0;
    if (key[u] != 0xFFFFFFFFu && lane == __ffs(peers[u]) - 1) base[u] = atomicAdd(count + key[u], __popc(peers[u]));
  }
#pragma unroll
  for (int u = 0; u < kLocPer; ++u) {
    const uint32_t i = i0 + u * kBlock;
    const uint32_t b = __shfl_sync(0xffffffffu, base[u], __ffs(peers[u]) - 1);
    if (i < cc.n) {
      keys[i] = key[u];
      // bits 24-31: how many particles of the cell this warp instruction ranked together (their slots
      // ascend with their index); k_place compares it with the cell's final count.  0 = unknown.
      const uint32_t together = cc.n <= (1u << 24) ? min((uint32_t)__popc(peers[u]), 255u) : 0u;
      slot[i] = (b + __popc(peers[u] & ((1u << lane) - 1u))) | (together << 24);
    }
  }
}

// Placement: sorted position = first position of the particle's cell + its slot inside the cell.
// The slots came from atomics, so the order INSIDE a cell is arbitrary here; k_order_cells fixes it.
template <bool Q>
__global__ void __launch_bounds__(kBlock)
    k_place(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ slot,
            const uint32_t* __restrict__ start, uint32_t n, uint32_t* __restrict__ keys_sorted,
            uint32_t* __restrict__ perm, uint8_t* __restrict__ cell_flags) {
  if (Q && !g_hot.active) return;
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= n) return;
  const uint32_t key = __ldg(keys + i);
  const uint32_t w = __ldg(slot + i);
  const bool packed = n <= (1u << 24);
  const uint32_t s = __ldg(start + key), d = s + (packed ? (w & 0xFFFFFFu) : w);
  keys_sorted[d] = key;
  perm[d] = i;
  if (cell_flags) {
    // A cell whose members were ALL ranked by one warp instruction is already in ascending order
    // of their index; any other cell with two or more members is flagged for k_order_cells.
    const uint32_t cnt = __ldg(start + key + 1u) - s, together = packed ? (w >> 24) : 0u;
    if (cnt > 1u && !(cnt < 255u && together == cnt)) cell_flags[key] = 1;
  }
}

// One thread per CELL: the members of a cell that holds two or more particles are put in
// ascending order of their previous index, which makes the result exactly the stable sort (and
// run-to-run deterministic).  A search cell holds ~1.2 particles, so most threads read two table
// entries and leave; small cells are insertion-sorted, a pile-up (hundreds of particles pressed
// into one cell) is heap-sorted: O(m log m) whatever the occupancy.
__device__ void order_members(uint32_t* __restrict__ a, uint32_t m) {
  if (m <= 4u) {  // the common cases: all loads issued together, a 5-comparator network, stores only if needed
    uint32_t v0 = a[0], v1 = a[1], v2 = m > 2u ? a[2] : 0xFFFFFFFFu, v3 = m > 3u ? a[3] : 0xFFFFFFFFu;
    const uint32_t o0 = v0, o1 = v1, o2 = v2, o3 = v3;
#define SPHB_CSWAP(x, y) { const uint32_t lo = min(x, y), hi = max(x, y); x = lo; y = hi; }
    SPHB_CSWAP(v0, v1) SPHB_CSWAP(v2, v3) SPHB_CSWAP(v0, v2) SPHB_CSWAP(v1, v3) SPHB_CSWAP(v1, v2)
#undef SPHB_CSWAP
    if (v0 != o0) a[0] = v0;
    if (v1 != o1) a[1] = v1;
    if (m > 2u && v2 != o2) a[2] = v2;
    if (m > 3u && v3 != o3) a[3] = v3;
    return;
  }
  if (m <= 16u) {
    for (uint32_t k = 1; k < m; ++k) {
      const uint32_t v = a[k];
      uint32_t j = k;
      for (; j > 0 && a[j - 1] > v; --j) a[j] = a[j - 1];
      a[j] = v;
    }
    return;
  }
  auto sift = [&](uint32_t root, uint32_t end) {
    for (;;) {
      uint32_t child = 2u * root + 1u;
      if (child >= end) return;
      if (child + 1u < end && a[child] < a[child + 1u]) ++child;
      if (a[root] >= a[child]) return;
      const uint32_t t = a[root]; a[root] = a[child]; a[child] = t;
      root = child;
    }
  };
  for (uint32_t k = m / 2u; k-- > 0;) sift(k, m);
  for (uint32_t end = m; end-- > 1;) {
    const uint32_t t = a[0]; a[0] = a[end]; a[end] = t;
    sift(0, end);
  }
}
constexpr uint32_t kCellsPerThread = 16;  // with flags: one 16-byte load decides for 16 cells
template <bool Q>
__global__ void __launch_bounds__(kBlock)
    k_order_cells(const uint32_t* __restrict__ start, uint32_t cells_arg, uint32_t* __restrict__ perm,
                  const uint8_t* __restrict__ cell_flags) {
  if (Q && !g_hot.active) return;
  const uint32_t cells = Q ? g_hot.c.nfine : cells_arg;
  if (cell_flags) {
    // (one thread per cell would be bound by launching 17 M threads that read one byte each)
    const uint32_t first = (blockIdx.x * kBlock + threadIdx.x) * kCellsPerThread;
    if (first >= cells) return;
    const uint4 f = __ldg(reinterpret_cast<const uint4*>(cell_flags + first));  // (the flags are padded to 16)
    if (!(f.x | f.y | f.z | f.w)) return;  // nothing to order: single members, or ranked in order already
    const uint32_t w[4] = {f.x, f.y, f.z, f.w};
#pragma unroll
    for (uint32_t k = 0; k < kCellsPerThread; ++k) {
      const uint32_t cell = first + k;
      if (((w[k >> 2] >> (8u * (k & 3u))) & 0xffu) && cell < cells) {
        const uint32_t s = __ldg(start + cell), e = __ldg(start + cell + 1u);
        if (e - s >= 2u) order_members(perm + s, e - s);
      }
    }
    return;
  }
  const uint32_t cell = blockIdx.x * kBlock + threadIdx.x;
  if (cell >= cells) return;
  const uint32_t s = __ldg(start + cell), e = __ldg(start + cell + 1u);
  if (e - s >= 2u) order_members(perm + s, e - s);
}

// The reference moves whole 80-byte structs in each of its 4 radix passes
// (kernels/sort.cl:41-59); here only (key, index) pairs are sorted and the carried state
// (position, velocity, intermediate_velocity: 36 of the 80 bytes, + the 4-byte reference
// rank) is gathered ONCE.
template <bool Q>
__global__ void __launch_bounds__(kBlock)
    k_reorder(const uint32_t* __restrict__ perm, uint32_t n, const float4* __restrict__ pos_in,
              const float4* __restrict__ vel_in, const float4* __restrict__ vhf_in,
              const uint32_t* __restrict__ rank_in, float4* __restrict__ pos_out,
              float4* __restrict__ vel_out, float4* __restrict__ vhf_out,
              uint32_t* __restrict__ rank_out, float4* __restrict__ pairs_out) {
  if (Q && !g_hot.active) return;
  const uint32_t k = blockIdx.x * kBlock + threadIdx.x;
  const bool live = k < n;
  float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
  if (live) {
    const uint32_t src = __ldg(perm + k);
    p = __ldg(pos_in + src);
    float4 v = __ldg(vel_in + src), w = __ldg(vhf_in + src);
    p.w = 0.f; v.w = 0.f; w.w = 0.f;
    pos_out[k] = p;
    vel_out[k] = v;
    vhf_out[k] = w;
    rank_out[k] = __ldg(rank_in + src);
  }
  // the positions once more as one 32-byte record per PAIR of consecutive particles,
  // {x0, x1, y0, y1 | z0, z1, 0, 0}: what the density pass walks (kernels_neighbors.cu) — one
  // 256-bit load brings two candidates with their coordinates already paired for fp32x2
  if (pairs_out) {  // (kernel argument: uniform; every lane of the warp reaches the shuffles)
    const float x1 = __shfl_down_sync(0xffffffffu, p.x, 1), y1 = __shfl_down_sync(0xffffffffu, p.y, 1),
                z1 = __shfl_down_sync(0xffffffffu, p.z, 1);
    if (live && !(k & 1u)) {
      pairs_out[k] = make_float4(p.x, x1, p.y, y1);   // record k / 2 = float4s k and k + 1
      pairs_out[k + 1] = make_float4(p.z, z1, 0.f, 0.f);
    }
  }
}

// Fine cell table, step 1: ends[f] = (index of the last particle of fine cell f) + 1 (else 0).
// An exclusive running maximum then turns `ends` into ftable[f] = first particle with key >= f.
__global__ void __launch_bounds__(kBlock)
    k_mark_cell_ends(const uint32_t* __restrict__ keys, uint32_t n, uint32_t* __restrict__ ends) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= n) return;
  const uint32_t k = __ldg(keys + i);
  if (i + 1 == n || __ldg(keys + i + 1) != k) ends[k] = i + 1;
}

// ---- the reference's order ---------------------------------------------------------------
// The reference keeps its particle array stably sorted by the Morton key of the 2h cell
// (sph_simulation.cpp:110-198): position of a particle = (number of particles in cells with a
// smaller key) + (number of particles of its own cell that preceded it in the previous array).
// The engine computes exactly that number for every particle — its "reference rank" — without
// moving any data: downloads scatter through it, and the parity hooks report it.
__device__ __forceinline__ void fine_coords_of(uint32_t key, const StepConsts& c, uint32_t& fx,
                                               uint32_t& fy, uint32_t& fz) {
  fx = key % c.fnx;
  const uint32_t t = key / c.fnx;
  fy = t % c.fny;
  fz = t / c.fny;
}

// histogram over Morton cells; lanes of a warp that share a cell add once
template <bool Q>
__global__ void __launch_bounds__(kBlock)
    k_morton_hist(const uint32_t* __restrict__ fine_keys, uint32_t* __restrict__ hist,
                  uint32_t* __restrict__ ckey, StepConsts c) {
  SPHB_Q_PROLOGUE(c);
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  const bool valid = i < cc.n;
  uint32_t m = 0xFFFFFFFFu;
  if (valid) {
    uint32_t fx, fy, fz;
    fine_coords_of(__ldg(fine_keys + i), cc, fx, fy, fz);
    m = morton3(fx / kXPerRef, fy >> 1, fz >> 1);
    ckey[i] = m;
  }
  const unsigned peers = __match_any_sync(0xffffffffu, m);
  const int lane = threadIdx.x & 31;
  if (valid && lane == __ffs(peers) - 1 && m < cc.grid_cell_count) atomicAdd(hist + m, __popc(peers));
}

// One thread per PAIR of consecutive particles: they share their Morton cell nine times out of
// ten, and then every previous rank that is loaded serves two comparisons.  A pair that straddles
// two cells is walked twice, one particle at a time.
template <bool Q>
__global__ void __launch_bounds__(kBlock)
    k_reference_rank(const uint32_t* __restrict__ fine_keys, const uint32_t* __restrict__ ftable,
                     const uint32_t* __restrict__ rank_prev, const uint32_t* __restrict__ ref_table,
                     uint32_t* __restrict__ rank_new, StepConsts c) {
  SPHB_Q_PROLOGUE(c);
  const uint32_t ia = 2u * (blockIdx.x * kBlock + threadIdx.x);
  if (ia >= cc.n) return;
  const bool has_b = ia + 1 < cc.n;
  const uint32_t ib = has_b ? ia + 1 : ia;
  uint32_t fx[2], fy[2], fz[2], m[2];
  fine_coords_of(__ldg(fine_keys + ia), cc, fx[0], fy[0], fz[0]);
  fine_coords_of(__ldg(fine_keys + ib), cc, fx[1], fy[1], fz[1]);
#pragma unroll
  for (int w = 0; w < 2; ++w) m[w] = morton3(fx[w] / kXPerRef, fy[w] >> 1, fz[w] >> 1);
  const uint32_t mine_a = __ldg(rank_prev + ia), mine_b = __ldg(rank_prev + ib);
  const bool together = has_b && m[0] == m[1];
  const int npass = (has_b && !together) ? 2 : 1;
  uint32_t before_a = 0, before_b = 0;
#pragma unroll 1
  for (int pass = 0; pass < npass; ++pass) {
    // a particle that sits out a pass compares against 0: nothing is smaller
    const uint32_t qa = pass == 0 ? mine_a : 0u, qb = (together || pass == 1) ? mine_b : 0u;
    const uint32_t x0 = fx[pass] / kXPerRef * kXPerRef, y0 = fy[pass] & ~1u, z0 = fz[pass] & ~1u;
    uint32_t rs[4], re[4];
#pragma unroll
    for (uint32_t k = 0; k < 4; ++k) {
      const uint32_t row = ((z0 + (k >> 1)) * cc.fny + (y0 + (k & 1u))) * cc.fnx + x0;
      rs[k] = __ldg(ftable + row);
      re[k] = __ldg(ftable + row + kXPerRef);
    }
#pragma unroll
    for (uint32_t k = 0; k < 4; ++k) {
      const uint32_t e = re[k];
      // four independent loads in flight; a slot past the run is not loaded and counts as "huge"
      for (uint32_t j = rs[k]; j < e; j += 4) {
        uint32_t r[4];
#pragma unroll
        for (uint32_t u = 0; u < 4; ++u) {
          r[u] = 0xFFFFFFFFu;
          if (j + u < e) r[u] = __ldg(rank_prev + j + u);  // (one predicated load)
        }
#pragma unroll
        for (uint32_t u = 0; u < 4; ++u) {
          before_a += r[u] < qa ? 1u : 0u;
          before_b += r[u] < qb ? 1u : 0u;
        }
      }
    }
  }
  rank_new[ia] = (m[0] < cc.grid_cell_count ? __ldg(ref_table + m[0]) : 0u) + before_a;
  if (has_b) rank_new[ib] = (m[1] < cc.grid_cell_count ? __ldg(ref_table + m[1]) : 0u) + before_b;
}

__global__ void __launch_bounds__(kBlock)
    k_reference_debug(const uint32_t* __restrict__ rank, const uint32_t* __restrict__ rank_prev,
                      const uint32_t* __restrict__ ckey, uint32_t n, uint32_t* __restrict__ keys_ref,
                      uint32_t* __restrict__ perm_ref) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = __ldg(rank + i);
  if (r < n) {
    keys_ref[r] = __ldg(ckey + i);
    perm_ref[r] = __ldg(rank_prev + i);
  }
}

// libclsph/kernels/collisionsv2.cl:140-156
__device__ __forceinline__ float weighted_average(float x, float x1, float x2, float d1, float d2) {
  return ((x2 - x) / (x2 - x1)) * d1 + ((x - x1) / (x2 - x1)) * d2;
}
__device__ __forceinline__ float bilinear(float x, float y, float xmin, float ymin, float xmax,
                                          float ymax, float d00, float d01, float d10, float d11) {
  const float R1 = weighted_average(x, xmin, xmax, d00, d10);
  const float R2 = weighted_average(x, xmin, xmax, d01, d11);
  return weighted_average(y, ymin, ymax, R1, R2);
}
__device__ __forceinline__ float len3(float x, float y, float z) {
  return sqrtf((x * x + y * y) + z * z);
}

// libclsph/kernels/sphb.cl:177-223 (advection_collision) = advection.cl:6-18 (leapfrog) +
// collisionsv2.cl:249-336 (handle_collisions) + :9-22 (respond), with the reductions of
// computeTimeStep (reduction.cl:66-124) and of the NEXT step's findMinMaxPosition
// (reduction.cl:2-64) fused in, so that a step ends with one 32-byte read-back.
#ifndef SPHB_ADVECT_MINB
#define SPHB_ADVECT_MINB 4  // 64 registers: 4 CTAs of 256 threads per SM (72 registers uncapped: 0.38 ms; capped: 0.328; no prefetch: 0.342)
#endif
template <bool Q>
__global__ void __launch_bounds__(kBlock, SPHB_ADVECT_MINB)
    k_advect(const float4* __restrict__ pos, const float4* __restrict__ vhf,
             const float4* __restrict__ acc, float4* __restrict__ pos_out,
             float4* __restrict__ vel_out, float4* __restrict__ vhf_out, uint32_t n, float dt,
             float restitution, const Face* __restrict__ faces, uint32_t face_count,
             const float* __restrict__ df, Reduce* red, SlabAdvect slab, FaceBin bin,
             const QueueCtl* __restrict__ q) {
  if (Q) {  // device-driven: this attempt runs only if the last one did not settle dt
    if (!q->attempt) return;
    dt = q->dt;
  }
  // grid-stride over a grid sized to the machine: one set of 8 atomics per CTA, not per 256
  // particles (65 536 CTAs hammering the same 8 words serialise in L2)
  float rv[8] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY, -INFINITY, -INFINITY};
  // (the pass is bound by load latency, not by issue slots: the NEXT particle's three loads are
  // requested before this one is worked on — six 16-byte loads in flight per thread instead of three)
  const uint32_t stride = gridDim.x * kBlock;
  uint32_t i_next = blockIdx.x * kBlock + threadIdx.x;
  float4 x0_n = make_float4(0.f, 0.f, 0.f, 0.f), vh_n = x0_n, a_n = x0_n;
  if (i_next < n) { x0_n = __ldg(pos + i_next); vh_n = __ldg(vhf + i_next); a_n = __ldg(acc + i_next); }
  for (uint32_t i = i_next; i < n; i = i_next) {
    const float4 x0 = x0_n, vh = vh_n, a = a_n;
    i_next = i + stride;
    if (i_next < n) { x0_n = __ldg(pos + i_next); vh_n = __ldg(vhf + i_next); a_n = __ldg(acc + i_next); }
    uint32_t o = i;  // where the result goes
    if (slab.ids) {
      // slab mode: a ghost belongs to another rank, which advects it; owned particles are
      // written densely (dst = exclusive scan of the owned flags)
      const uint32_t id = __ldg(slab.ids + i);
      if (id & 0x80000000u) continue;
      o = __ldg(slab.dst + i);
      if (slab.write_aux) {
        slab.ids_out[o] = id;
        slab.src_of[o] = i;
      }
    }
    // leapfrog
    const float nx = vh.x + a.x * dt, ny = vh.y + a.y * dt, nz = vh.z + a.z * dt;
    const float px = x0.x + nx * dt, py = x0.y + ny * dt, pz = x0.z + nz * dt;

    // LAST face whose inflated AABB contains the advected position
    int indice = -1;
    const float4* f4 = reinterpret_cast<const float4*>(faces);
    if (bin.start) {
      // candidates of the point's bin cell, highest face index first
      const float bx_ = (px - bin.ox) * bin.inv, by_ = (py - bin.oy) * bin.inv, bz_ = (pz - bin.oz) * bin.inv;
      if (bx_ >= 0.f && by_ >= 0.f && bz_ >= 0.f) {
        const uint32_t cx = (uint32_t)bx_, cy = (uint32_t)by_, cz = (uint32_t)bz_;
        if (cx < bin.nx && cy < bin.ny && cz < bin.nz) {
          const uint32_t cell = (cz * bin.ny + cy) * bin.nx + cx;
          const uint32_t k1 = __ldg(bin.start + cell + 1);
          for (uint32_t k = __ldg(bin.start + cell); k < k1; ++k) {
            const int f = (int)__ldg(bin.faces + k);
            const float4 b0 = __ldg(f4 + 3 * f);
            const float4 b1 = __ldg(f4 + 3 * f + 1);
            if (px <= b0.x && px >= b0.w && py <= b0.y && py >= b1.x && pz <= b0.z && pz >= b1.y) {
              indice = f;
              break;
            }
          }
        }
      }
    } else {
      for (int f = 0; f < (int)face_count; ++f) {
        const float4 b0 = __ldg(f4 + 3 * f);
        const float4 b1 = __ldg(f4 + 3 * f + 1);
        if (px <= b0.x && px >= b0.w && py <= b0.y && py >= b1.x && pz <= b0.z && pz >= b1.y)
          indice = f;
      }
    }
    float rx = px, ry = py, rz = pz, wx = nx, wy = ny, wz = nz;
    if (indice > -1) {
      const Face F = faces[indice];
      const float sidex = (F.maxx - F.minx) / F.sxm1;
      const float sidey = (F.maxy - F.miny) / F.sym1;
      const float sidez = (F.maxz - F.minz) / F.szm1;
      const int x = (int)((px - F.minx) / sidex);
      const int y = (int)((py - F.miny) / sidey);
      const int z = (int)((pz - F.minz) / sidez);
      const float bx = x * sidex + F.minx;
      const float by = y * sidey + F.miny;
      const float bz = z * sidez + F.minz;
      const float xf = (float)x, yf = (float)y, zf = (float)z;
      // getDFindex: fp32 index arithmetic, left to right (collisionsv2.cl:152-156)
#define DFI(a_, b_, c_) \
  __ldg(df + (int)(F.off + (yf + (float)(b_)) * F.sx * F.sz + F.sx * (zf + (float)(c_)) + xf + (float)(a_)))
      const float d000 = DFI(0, 0, 0), d001 = DFI(0, 0, 1), d100 = DFI(1, 0, 0), d101 = DFI(1, 0, 1);
      const float d010 = DFI(0, 1, 0), d011 = DFI(0, 1, 1), d110 = DFI(1, 1, 0), d111 = DFI(1, 1, 1);
#undef DFI
      const float facedown = bilinear(px, pz, bx, bz, bx + sidex, bz + sidez, d000, d001, d100, d101);
      const float faceup = bilinear(px, pz, bx, bz, bx + sidex, bz + sidez, d010, d011, d110, d111);
      const float d = weighted_average(py, by, by + sidey, facedown, faceup);
      if ((double)d < 0.02) {  // compared in double, as the reference's literal makes it
        const float faceright = bilinear(py, pz, by, bz, by + sidey, bz + sidez, d100, d101, d110, d111);
        const float faceleft = bilinear(py, pz, by, bz, by + sidey, bz + sidez, d000, d001, d010, d011);
        const float faceback = bilinear(px, py, bx, by, bx + sidex, by + sidey, d000, d010, d100, d110);
        const float facefront = bilinear(px, py, bx, by, bx + sidex, by + sidey, d001, d011, d101, d111);
        float mx = faceright - faceleft, my = faceup - facedown, mz = facefront - faceback;
        const float lenn = len3(mx, my, mz);
        mx = mx / lenn; my = my / lenn; mz = mz / lenn;
        // respond()
        const float ad = fabsf(d);
        rx = px + ad * mx; ry = py + ad * my; rz = pz + ad * mz;
        const float dotvn = (nx * mx + ny * my) + nz * mz;
        const float f = (1.f + restitution * ad / (dt * len3(nx, ny, nz))) * dotvn;
        wx = nx - f * mx; wy = ny - f * my; wz = nz - f * mz;
      }
    }
    const float vx = (vh.x + wx) / 2.f, vy = (vh.y + wy) / 2.f, vz = (vh.z + wz) / 2.f;
    pos_out[o] = make_float4(rx, ry, rz, 0.f);
    vel_out[o] = make_float4(vx, vy, vz, 0.f);
    vhf_out[o] = make_float4(wx, wy, wz, 0.f);
    // fminf / fmaxf return the non-NaN operand: NaNs are skipped, as OpenCL's min/max do
    rv[0] = fminf(rv[0], rx); rv[1] = fminf(rv[1], ry); rv[2] = fminf(rv[2], rz);
    rv[3] = fmaxf(rv[3], rx); rv[4] = fmaxf(rv[4], ry); rv[5] = fmaxf(rv[5], rz);
    rv[6] = fmaxf(rv[6], vx * vx + vy * vy + vz * vz);
    rv[7] = fmaxf(rv[7], a.x * a.x + a.y * a.y + a.z * a.z);
  }
  block_reduce_to(red, rv, 0xff);
}

// libclsph/kernels/collisionsv2.cl:24-55
__device__ __forceinline__ float det2(float x1, float y1, float x2, float y2) {
  return x1 * y2 - y1 * x2;
}
__device__ float dist_point_segment(float x, float y, float z, float x1, float y1, float x2,
                                    float y2) {
  const float A = y - x1, B = z - y1, C = x2 - x1, D = y2 - y1;
  const float dot = A * C + B * D;
  const float len_sq = C * C + D * D;
  float param = -1;
  if (len_sq != 0) param = dot / len_sq;
  float xx, yy;
  if (param < 0) {
    xx = x1; yy = y1;
  } else if (param > 1) {
    xx = x2; yy = y2;
  } else {
    xx = x1 + param * C; yy = y1 + param * D;
  }
  const float dy = y - xx, dz = z - yy;
  return sqrtf(x * x + dz * dz + dy * dy);
}

// libclsph/kernels/collisionsv2.cl:57-138 — one thread per distance-field grid point.
// The reference finds the owning face by a linear walk down the offsets (:58-68); offsets
// are strictly increasing, so an upper_bound gives the same face.
// The reference then tests EVERY face's box against the point (:70-76); here only the faces
// whose box meets the owning face's box are walked (cand_start / cand_faces: ascending face
// index per owner, built by sphb_set_scene with a margin), with the same per-point test — the
// faces left out fail it for every point of the owner's grid, so the field is unchanged.
__global__ void __launch_bounds__(kBlock)
    k_distance_field(float* __restrict__ df, const Face* __restrict__ faces,
                     const uint32_t* __restrict__ offsets, const float* __restrict__ tr,
                     const float* __restrict__ rv, const uint32_t* __restrict__ cand_start,
                     const uint32_t* __restrict__ cand_faces, uint32_t face_count,
                     uint32_t gridcount) {
  const uint32_t g = blockIdx.x * kBlock + threadIdx.x;
  if (g >= gridcount) return;
  uint32_t lo = 0, hi = face_count;  // largest index with offsets[index] <= g (index 0 if none)
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (__ldg(offsets + mid) <= g) lo = mid; else hi = mid;
  }
  const Face F = faces[lo];
  const uint32_t rel = g - __ldg(offsets + lo);
  const uint32_t sx = (uint32_t)F.sx, sz = (uint32_t)F.sz;
  const int x = (int)((rel % (sx * sz)) % sx);
  const int z = (int)((rel % (sx * sz)) / sx);
  const int y = (int)(rel / (sx * sz));
  const float px = x * (F.maxx - F.minx) / F.sxm1 + F.minx;
  const float py = y * (F.maxy - F.miny) / F.sym1 + F.miny;
  const float pz = z * (F.maxz - F.minz) / F.szm1 + F.minz;

  float temd = 20;
  const float4* f4 = reinterpret_cast<const float4*>(faces);
  const uint32_t k0 = cand_start ? __ldg(cand_start + lo) : 0u;
  const uint32_t k1 = cand_start ? __ldg(cand_start + lo + 1) : face_count;
  for (uint32_t k = k0; k < k1; ++k) {
    const int i = cand_start ? (int)__ldg(cand_faces + k) : (int)k;
    const float4 b0 = __ldg(f4 + 3 * i);
    const float4 b1 = __ldg(f4 + 3 * i + 1);
    if (px <= b0.x && px >= b0.w && py <= b0.y && py >= b1.x && pz <= b0.z && pz >= b1.y) {
      const float4 t0 = __ldg(reinterpret_cast<const float4*>(tr) + 3 * i);
      const float4 t1 = __ldg(reinterpret_cast<const float4*>(tr) + 3 * i + 1);
      const float4 t2 = __ldg(reinterpret_cast<const float4*>(tr) + 3 * i + 2);
      const float4 R = __ldg(reinterpret_cast<const float4*>(rv) + i);
      const float tpx = px + t0.w, tpy = py + t1.w, tpz = pz + t2.w;
      const float rpx = t0.x * tpx + t0.y * tpy + t0.z * tpz;
      const float rpy = t1.x * tpx + t1.y * tpy + t1.z * tpz;
      const float rpz = t2.x * tpx + t2.y * tpy + t2.z * tpz;
      const float v1y = R.y, v2x = R.z, v2y = R.w;
      const float a = det2(rpy, rpz, 0, v1y) / det2(v2x, v2y, 0, v1y);
      const float b = -det2(rpy, rpz, v2x, v2y) / det2(v2x, v2y, 0, v1y);
      float d, td;
      if (a > 0 && b > 0 && a + b < 1) {
        d = fabsf(rpx);
      } else {
        d = dist_point_segment(rpx, rpy, rpz, 0, 0, R.x, R.y);
        td = dist_point_segment(rpx, rpy, rpz, R.x, R.y, R.z, R.w);
        if (td < d) d = td;
        td = dist_point_segment(rpx, rpy, rpz, 0, 0, R.z, R.w);
        if (td < d) d = td;
      }
      if (d < fabsf(temd)) temd = copysignf(d, rpx);
    }
  }
  df[g] = temd;
}

// 80-byte AoS `particle` (libclsph/common/structures.h:40-44) <-> the engine's SoA.  The AoS
// array is in the reference's order; the SoA is in the engine's compute order, and rank[i] is
// where particle i sits in the reference's order.
__global__ void __launch_bounds__(kBlock)
    k_aos_to_soa(const sphb_particle* __restrict__ aos, uint32_t n, float4* pos, float4* vel,
                 float4* vhf, float4* acc, float* rho, float* prs, uint32_t* ckey, uint32_t* rank) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= n) return;
  const float4* src = reinterpret_cast<const float4*>(aos + i);
  float4 p = src[0], v = src[1], w = src[2], a = src[3];
  const float4 tail = src[4];
  p.w = 0.f; v.w = 0.f; w.w = 0.f; a.w = 0.f;
  pos[i] = p; vel[i] = v; vhf[i] = w; acc[i] = a;
  rho[i] = tail.x;
  prs[i] = tail.y;
  ckey[i] = __float_as_uint(tail.z);
  rank[i] = i;
}

__global__ void __launch_bounds__(kBlock)
    k_soa_to_aos(sphb_particle* __restrict__ aos, uint32_t n, const float4* __restrict__ pos,
                 const float4* __restrict__ vel, const float4* __restrict__ vhf,
                 const float4* __restrict__ acc, const float* __restrict__ rho,
                 const float* __restrict__ prs, const uint32_t* __restrict__ ckey,
                 const uint32_t* __restrict__ rank) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = __ldg(rank + i);
  if (r >= n) return;
  float4 p = pos[i], v = vel[i], w = vhf[i], a = acc[i];
  p.w = 0.f; v.w = 0.f; w.w = 0.f; a.w = 0.f;
  float4* dst = reinterpret_cast<float4*>(aos + r);
  dst[0] = p; dst[1] = v; dst[2] = w; dst[3] = a;
  dst[4] = make_float4(rho[i], prs[i], __uint_as_float(ckey[i]), 0.f);
}

inline unsigned blocks_for(uint32_t n) { return (n + kBlock - 1) / kBlock; }

}  // namespace

int launch_reduce_init(Reduce* red, cudaStream_t s, const QueueCtl* q) {
  if (q) k_reduce_init<true><<<1, 32, 0, s>>>(red, q);
  else k_reduce_init<false><<<1, 32, 0, s>>>(red, nullptr);
  return 1;
}

// ---- device-driven stepping: the two one-thread kernels that stand in for the host ----------
__global__ void k_queue_plan(QueueCtl* q, const Reduce* __restrict__ red) {
  if (threadIdx.x) return;
  q->hot.active = 0;
  q->attempt = 0;
  if (q->status != SPHB_Q_RUNNING) return;
  if (!(q->timeleft > 0.0)) {  // while (timeleft > 0.0), sph_simulation.cpp:384
    q->status = SPHB_Q_DONE;
    return;
  }
  if (q->budget == 0) {
    q->status = SPHB_Q_BUDGET;
    return;
  }
  // findMinMaxPosition's host half (:668-727) on the bbox the last advection reduced
  float mn[3], mx[3];
  for (int a = 0; a < 3; ++a) {
    mn[a] = dec_float(red->min_enc[a]);
    mx[a] = dec_float(red->max_enc[a]);
  }
  sphb_params P = q->params;
  if (!grid_from_bbox_hd(P, mn, mx)) {
    q->params = P;
    q->status = SPHB_Q_GRID;
    return;
  }
  const StepConsts c = make_consts_hd(P, q->terms, q->spiky_eps, q->hot.c.n);
  const unsigned long long fine = (unsigned long long)c.fnx * c.fny * c.fnz + 1ull;
  if (fine > q->fcap || (unsigned long long)P.grid_cell_count + 1ull > q->rcap) {
    q->status = SPHB_Q_GROW;  // nothing of this step has been committed
    return;
  }
  q->params = P;
  q->hot.c = c;
  q->hot.active = 1;
  q->attempt = 1;
  q->attempts = 0;
  q->budget -= 1;
}

__global__ void k_queue_attempt(QueueCtl* q, const Reduce* __restrict__ red, float* __restrict__ log, int last) {
  if (threadIdx.x) return;
  if (!q->attempt) return;
  // computeTimeStep's host half (:598-602) and the retry test (:260)
  const float dt = q->dt;
  const float computed = dt_from_maxima(q->params, dec_float(red->max_v2), dec_float(red->max_a2));
  q->attempts += 1;
  if (dt - computed > 0.00001) {
    q->dt = computed;  // the next attempt's dt
    if (last) {
      q->status = SPHB_Q_ATTEMPTS;
      q->attempt = 0;
    }
    return;
  }
  // the step is done; the caller's loop (:390-394)
  q->attempt = 0;
  float timeleft = q->timeleft;
  float next = computed;
  timeleft -= next;
  if (timeleft < next) next = timeleft;
  q->timeleft = timeleft;
  q->dt = next;
  if (q->steps < q->log_cap) log[q->steps] = timeleft;
  q->steps += 1;
}

int launch_queue_plan(QueueCtl* ctl, const Reduce* red, cudaStream_t s) {
  k_queue_plan<<<1, 32, 0, s>>>(ctl, red);
  return 1;
}

int launch_queue_attempt(QueueCtl* ctl, const Reduce* red, float* timeleft_log, int last, cudaStream_t s) {
  k_queue_attempt<<<1, 32, 0, s>>>(ctl, red, timeleft_log, last);
  return 1;
}

int queue_refresh_exact(const QueueCtl* ctl, cudaStream_t s) {
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

int launch_bbox(const float4* pos, uint32_t n, Reduce* red, cudaStream_t s) {
  unsigned grid = blocks_for(n);
  if (grid > (unsigned)sm_count() * 8u) grid = (unsigned)sm_count() * 8u;  // grid-stride: 8 resident CTAs of 256 threads per SM
  k_bbox<<<grid, kBlock, 0, s>>>(pos, n, red);
  return 1;
}

int launch_locate(const float4* pos, uint32_t* keys, const StepConsts& c, cudaStream_t s) {
  k_locate<<<blocks_for(c.n), kBlock, 0, s>>>(pos, keys, c);
  return 1;
}

// `table_entries`: how much of ftable is cleared and scanned — nfine + 1 on the host-driven path;
// on the device-driven path (queued) the whole capacity, because the live length is only known
// on the device (entries past it scan to n and are never read).
int launch_cell_sort(const float4* pos, const StepConsts& c, uint32_t* keys_tmp, uint32_t* slot_tmp,
                     uint32_t* ftable, uint32_t* scan_scratch, uint32_t* keys_sorted, uint32_t* src_tmp,
                     uint32_t* perm, cudaStream_t s, bool queued, size_t table_entries, uint8_t* cell_flags) {
  if (!queued) table_entries = (size_t)c.nfine + 1;
  cudaMemsetAsync(ftable, 0, sizeof(uint32_t) * table_entries, s);
  if (cell_flags) cudaMemsetAsync(cell_flags, 0, (table_entries + 15) & ~(size_t)15, s);
  const unsigned loc_grid = (c.n + kBlock * kLocPer - 1) / (kBlock * kLocPer);
  // The slots live in src_tmp, NOT in slot_tmp: the callers hand the same buffer as slot_tmp and as
  // perm, and the placement writes perm while other threads still read their slots.
  (void)slot_tmp;
  uint32_t* const slots = src_tmp;
  if (queued) k_locate_count<true><<<loc_grid, kBlock, 0, s>>>(pos, keys_tmp, slots, ftable, c);
  else k_locate_count<false><<<loc_grid, kBlock, 0, s>>>(pos, keys_tmp, slots, ftable, c);
  int l = 1 + launch_exclusive_sum(ftable, (uint32_t)table_entries, scan_scratch, s);
  // (the placement writes the permutation directly; k_order_cells then orders each cell in place)
  const size_t cell_threads = cell_flags ? (table_entries - 1 + kCellsPerThread - 1) / kCellsPerThread : table_entries - 1;
  const unsigned cell_grid = (unsigned)((cell_threads + kBlock - 1) / kBlock);
  if (queued) {
    k_place<true><<<blocks_for(c.n), kBlock, 0, s>>>(keys_tmp, slots, ftable, c.n, keys_sorted, perm, cell_flags);
    k_order_cells<true><<<cell_grid, kBlock, 0, s>>>(ftable, 0u, perm, cell_flags);
  } else {
    k_place<false><<<blocks_for(c.n), kBlock, 0, s>>>(keys_tmp, slots, ftable, c.n, keys_sorted, perm, cell_flags);
    k_order_cells<false><<<cell_grid, kBlock, 0, s>>>(ftable, c.nfine, perm, cell_flags);
  }
  return l + 2;
}

int launch_reorder(const uint32_t* perm, uint32_t n, const float4* pos_in, const float4* vel_in,
                   const float4* vhf_in, const uint32_t* rank_in, float4* pos_out, float4* vel_out,
                   float4* vhf_out, uint32_t* rank_out, cudaStream_t s, bool queued, float4* pairs_out) {
  if (queued)
    k_reorder<true><<<blocks_for(n), kBlock, 0, s>>>(perm, n, pos_in, vel_in, vhf_in, rank_in, pos_out,
                                                     vel_out, vhf_out, rank_out, pairs_out);
  else
    k_reorder<false><<<blocks_for(n), kBlock, 0, s>>>(perm, n, pos_in, vel_in, vhf_in, rank_in, pos_out,
                                                      vel_out, vhf_out, rank_out, pairs_out);
  return 1;
}

int launch_fine_table(const uint32_t* sorted_keys, uint32_t n, uint32_t* ftable, uint32_t nfine,
                      uint32_t* scratch, cudaStream_t s) {
  cudaMemsetAsync(ftable, 0, sizeof(uint32_t) * ((size_t)nfine + 1), s);
  k_mark_cell_ends<<<blocks_for(n), kBlock, 0, s>>>(sorted_keys, n, ftable);
  return 1 + launch_exclusive_max(ftable, nfine + 1, scratch, s);
}

int launch_reference_rank(const uint32_t* fine_keys, const uint32_t* ftable, const uint32_t* rank_prev,
                          uint32_t* ref_table, uint32_t* ckey, uint32_t* rank_new, const StepConsts& c,
                          uint32_t* scratch, cudaStream_t s, bool queued, size_t table_entries) {
  if (!queued) table_entries = (size_t)c.grid_cell_count + 1;
  cudaMemsetAsync(ref_table, 0, sizeof(uint32_t) * table_entries, s);
  if (queued) k_morton_hist<true><<<blocks_for(c.n), kBlock, 0, s>>>(fine_keys, ref_table, ckey, c);
  else k_morton_hist<false><<<blocks_for(c.n), kBlock, 0, s>>>(fine_keys, ref_table, ckey, c);
  int l = 1 + launch_exclusive_sum(ref_table, (uint32_t)table_entries, scratch, s);
  if (queued)
    k_reference_rank<true><<<blocks_for((c.n + 1u) / 2u), kBlock, 0, s>>>(fine_keys, ftable, rank_prev, ref_table,
                                                                          rank_new, c);
  else
    k_reference_rank<false><<<blocks_for((c.n + 1u) / 2u), kBlock, 0, s>>>(fine_keys, ftable, rank_prev, ref_table,
                                                                           rank_new, c);
  return l + 1;
}

int launch_reference_debug(const uint32_t* rank, const uint32_t* rank_prev, const uint32_t* ckey,
                           uint32_t n, uint32_t* keys_ref, uint32_t* perm_ref, cudaStream_t s) {
  k_reference_debug<<<blocks_for(n), kBlock, 0, s>>>(rank, rank_prev, ckey, n, keys_ref, perm_ref);
  return 1;
}

int launch_advect(const float4* pos, const float4* vhf, const float4* acc, float4* pos_out,
                  float4* vel_out, float4* vhf_out, uint32_t n, float dt, float restitution,
                  const Face* faces, uint32_t face_count, const float* df, Reduce* red,
                  cudaStream_t s, const SlabAdvect* slab, const FaceBin* bin, const QueueCtl* q) {
  FaceBin b = {};
  if (bin) b = *bin;
  SlabAdvect sa = {};
  if (slab) sa = *slab;
  unsigned grid = blocks_for(n);
  if (grid > (unsigned)sm_count() * 8u) grid = (unsigned)sm_count() * 8u;  // grid-stride: 8 resident CTAs of 256 threads per SM
  if (q)
    k_advect<true><<<grid, kBlock, 0, s>>>(pos, vhf, acc, pos_out, vel_out, vhf_out, n, dt, restitution, faces,
                                           face_count, df, red, sa, b, q);
  else
    k_advect<false><<<grid, kBlock, 0, s>>>(pos, vhf, acc, pos_out, vel_out, vhf_out, n, dt, restitution, faces,
                                            face_count, df, red, sa, b, nullptr);
  return 1;
}

int launch_distance_field(float* df, const Face* faces, const uint32_t* face_offsets,
                          const float* transforms, const float* rvertices,
                          const uint32_t* cand_start, const uint32_t* cand_faces,
                          uint32_t face_count, uint32_t gridcount, cudaStream_t s) {
  if (gridcount == 0 || face_count == 0) return 0;
  k_distance_field<<<blocks_for(gridcount), kBlock, 0, s>>>(df, faces, face_offsets, transforms,
                                                            rvertices, cand_start, cand_faces,
                                                            face_count, gridcount);
  return 1;
}

int launch_aos_to_soa(const sphb_particle* aos, uint32_t n, float4* pos, float4* vel, float4* vhf,
                      float4* acc, float* rho, float* prs, uint32_t* ckey, uint32_t* rank,
                      cudaStream_t s) {
  k_aos_to_soa<<<blocks_for(n), kBlock, 0, s>>>(aos, n, pos, vel, vhf, acc, rho, prs, ckey, rank);
  return 1;
}

int launch_soa_to_aos(sphb_particle* aos, uint32_t n, const float4* pos, const float4* vel,
                      const float4* vhf, const float4* acc, const float* rho, const float* prs,
                      const uint32_t* ckey, const uint32_t* rank, cudaStream_t s) {
  k_soa_to_aos<<<blocks_for(n), kBlock, 0, s>>>(aos, n, pos, vel, vhf, acc, rho, prs, ckey, rank);
  return 1;
}

}  // namespace sphb
