// Shared declarations of the sphb engine's translation units (internal; the public
// boundary is include/sphb.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sphb.h"

namespace sphb {

// Everything a kernel needs from `simulation_parameters` / `precomputed_kernel_values`
// for one step, passed by value (the reference passes the whole 128-byte struct by value
// to every kernel, libclsph/sph_simulation.cpp:205-206,224).
struct StepConsts {
  uint32_t n;                // particles_count
  uint32_t grid_cell_count;  // Morton bound of the grid (cell table length)
  float h, h2, cell;         // support radius, h*h, cell side = h*2
  float minx, miny, minz;    // min_point
  float mass, K, rho0, mu;   // particle_mass, K, fluid_density, dynamic_viscosity
  float st_threshold, st;    // surface_tension_threshold, surface_tension
  float gx, gy, gz;          // constant_acceleration
  float c_poly6, c_poly6_grad, c_poly6_lap, c_spiky, c_visc;
  float spiky_eps;           // -45 / (float)(M_PI * h^6): spiky_gradient's r < 1e-7 branch
  // The engine's own search grid, row-major (x fastest): cells of side h (half a reference
  // cell) in y and z, fine coordinate = 2 * reference coordinate + (upper half ? 1 : 0); along x —
  // the direction a row of cells is walked in — the cell is cut kXSub times finer still, so
  // that the run of candidates walked for a row can be trimmed to the chord of the support sphere
  // at that row instead of to whole cells of side h.
  uint32_t fnx, fny, fnz;    // fine grid dimensions = 2 * {kXSub, 1, 1} * grid_size_{x,y,z}
  uint32_t nfine;            // fnx * fny * fnz
};
#ifndef SPHB_XSUB
#define SPHB_XSUB 4
#endif
constexpr uint32_t kXSub = SPHB_XSUB;         // x sub-cells per cell of side h
constexpr uint32_t kXPerRef = 2u * kXSub;     // x sub-cells per reference cell (side 2h)

// One face of the scene for the collision pass: inflated AABB + its distance-field grid
// (libclsph/common/structures.h:54-65, integers pre-converted to fp32 because the
// reference does its index arithmetic in fp32, kernels/collisionsv2.cl:152-156).
struct Face {
  float maxx, maxy, maxz, minx;  // one float4
  float miny, minz, sx, sz;      // one float4: size_x, size_z as float
  float sxm1, sym1, szm1, off;   // one float4: size-1 as float, offset as float
};

// Uniform-grid triangle bin over the union of the faces' inflated AABBs.  Bin cell c lists, in
// DESCENDING order, every face whose AABB (grown by one bin cell, so that rounding in the
// particle's bin coordinate can never hide a face) touches it: the collision pass walks that
// list and stops at the first face whose AABB really contains the point — which is the
// reference's "last face containing the point" (kernels/collisionsv2.cl:255-261) without
// its loop over ALL faces.
struct FaceBin {
  const uint32_t* start;  // nx*ny*nz + 1 entries (nullptr: bin not built, scan all faces)
  const uint32_t* faces;  // face indices
  float ox, oy, oz;       // origin
  float inv;              // 1 / bin cell side
  uint32_t nx, ny, nz;
};

// Slab mode of the advection pass: ghosts (bit 31 of the id) are skipped, and the owned
// particles are written DENSELY at dst[i] (an exclusive scan of the owned flags), so that the
// step ends with the owned particles compacted without a separate pass.  Besides the state,
// the pass writes the id, the compute-order index (src_of, to find acceleration / density /
// pressure later) and the reference's Morton grid_index under the global grid origin.
struct SlabAdvect {
  const uint32_t* ids;  // nullptr: not in slab mode
  const uint32_t* dst;
  uint32_t* ids_out;
  uint32_t* src_of;
  int write_aux;  // ids_out / src_of do not depend on dt: only a step's first attempt writes them
};

// Reduction cell shared by the bbox / time-step reductions.  Floats are stored in an
// order-preserving unsigned encoding so that atomicMin/atomicMax work on them.
struct Reduce {
  unsigned min_enc[3];
  unsigned max_enc[3];
  unsigned max_v2;  // max |velocity|^2      (kernels/reduction.cl:66-94)
  unsigned max_a2;  // max |acceleration|^2  (kernels/reduction.cl:96-124)
};

__host__ __device__ inline unsigned enc_float(float f) {
#ifdef __CUDA_ARCH__
  unsigned b = __float_as_uint(f);
#else
  unsigned b;
  memcpy(&b, &f, 4);
#endif
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ inline float dec_float(unsigned u) {
  unsigned b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  float f;
  memcpy(&f, &b, 4);
  return f;
#endif
}

// 10-bit-per-axis Morton code (libclsph/common/util.h:41-62) and its inverse (:4-38).
__host__ __device__ inline uint32_t spread10(uint32_t x) {
  x = (x | (x << 16)) & 0x030000FFu;
  x = (x | (x << 8)) & 0x0300F00Fu;
  x = (x | (x << 4)) & 0x030C30C3u;
  x = (x | (x << 2)) & 0x09249249u;
  return x;
}
__host__ __device__ inline uint32_t morton3(uint32_t x, uint32_t y, uint32_t z) {
  return spread10(x) | (spread10(y) << 1) | (spread10(z) << 2);
}
__host__ __device__ inline uint32_t compact10(uint32_t v) {
  v &= 0x09249249u;
  v = (v | (v >> 2)) & 0x030C30C3u;
  v = (v | (v >> 4)) & 0x0300F00Fu;
  v = (v | (v >> 8)) & 0x030000FFu;
  v = (v | (v >> 16)) & 0x000003FFu;
  return v;
}

// ---- host arithmetic of a step, shared by the host path and the device-driven path ---------
// libclsph/sph_simulation.cpp:668-727 — fp32, same expressions.  Returns false when a grid
// dimension reached 1024 (the reference asserts, :722-724).
__host__ __device__ inline bool grid_from_bbox_hd(sphb_params& P, const float mn[3], const float mx[3]) {
  const float side = (P.h * 2);
  int32_t gs[3];
  for (int a = 0; a < 3; ++a) {
    P.min_point.s[a] = mn[a] - 2 * side;
    P.max_point.s[a] = mx[a] + 2 * side;
    gs[a] = (int32_t)static_cast<uint32_t>((P.max_point.s[a] - P.min_point.s[a]) / (side));
  }
  P.grid_size_x = gs[0]; P.grid_size_y = gs[1]; P.grid_size_z = gs[2];
  if (!(gs[0] < 1024) || !(gs[1] < 1024) || !(gs[2] < 1024) || gs[0] < 0 || gs[1] < 0 || gs[2] < 0)
    return false;
  P.grid_cell_count = morton3((uint32_t)gs[0], (uint32_t)gs[1], (uint32_t)gs[2]);
  return true;
}

// libclsph/sph_simulation.cpp:598-602 — fp32 arithmetic; comparisons against the double
// literal 0.00001 as written.
__host__ __device__ inline float dt_from_maxima(const sphb_params& P, float max_v2, float max_a2) {
  float maxVel2 = max_v2;
  float maxVel = sqrtf(max_v2);
  float maxAccel = sqrtf(max_a2);
  float dt = (sqrtf(2 * maxAccel * P.h + maxVel2) - maxVel) / (2 * maxAccel);
  if (dt < 0.00001) dt = 0.00001;
  if (dt > 1.0f / (P.target_fps) * P.simulation_scale) dt = 1.0f / (P.target_fps) * P.simulation_scale;
  return dt;
}

__host__ __device__ inline StepConsts make_consts_hd(const sphb_params& P, const sphb_terms& T,
                                                     float spiky_eps, uint32_t n) {
  StepConsts c;
  c.n = n;
  c.grid_cell_count = P.grid_cell_count;
  c.h = P.h;
  c.h2 = P.h * P.h;
  c.cell = (P.h * 2);
  c.minx = P.min_point.s[0]; c.miny = P.min_point.s[1]; c.minz = P.min_point.s[2];
  c.mass = P.particle_mass; c.K = P.K; c.rho0 = P.fluid_density; c.mu = P.dynamic_viscosity;
  c.st_threshold = P.surface_tension_threshold; c.st = P.surface_tension;
  c.gx = P.constant_acceleration.s[0]; c.gy = P.constant_acceleration.s[1];
  c.gz = P.constant_acceleration.s[2];
  c.c_poly6 = T.poly_6; c.c_poly6_grad = T.poly_6_gradient;
  c.c_poly6_lap = T.poly_6_laplacian; c.c_spiky = T.spiky;
  c.c_visc = T.viscosity;
  c.spiky_eps = spiky_eps;
  c.fnx = kXPerRef * (uint32_t)P.grid_size_x;
  c.fny = 2u * (uint32_t)P.grid_size_y;
  c.fnz = 2u * (uint32_t)P.grid_size_z;
  c.nfine = c.fnx * c.fny * c.fnz;
  return c;
}

// ---- device-driven stepping (sphb_run_frame) ------------------------------------------------
// What the host decides between two steps of the reference's sub-step loop
// (libclsph/sph_simulation.cpp:384-409 and :246-260) — the next grid, whether the advection must
// be repeated with a smaller dt, how much of the frame is left — is decided by one-thread
// kernels instead, so a whole frame's steps are enqueued without a read-back.  Every kernel of
// a queued step has fixed launch arguments: the step's constants reach it through a
// __constant__ copy of QueueHot (refreshed by a device-to-device copy per step), dt and the
// retry decision through the QueueCtl block in global memory.
enum {
  SPHB_Q_RUNNING = 0,   // the batch ended with the frame unfinished: enqueue more
  SPHB_Q_DONE = 1,      // timeleft <= 0
  SPHB_Q_GROW = 2,      // the next step's tables exceed their capacity: the host grows them
  SPHB_Q_GRID = 3,      // a grid dimension reached 1024
  SPHB_Q_ATTEMPTS = 4,  // the queued advection attempts did not settle dt: the host continues
  SPHB_Q_BUDGET = 5     // the step budget (max_steps) is used up
};
struct QueueHot {
  StepConsts c;
  uint32_t active;  // the step's kernels run
  uint32_t pad[3];
};
struct QueueCtl {
  QueueHot hot;
  sphb_params params;     // the reference's `parameters` as of the last step that started
  sphb_terms terms;
  float spiky_eps;
  float dt;               // dt of the next advection attempt; between steps: the loop's `dt`
  float timeleft;
  uint32_t attempt;       // the next advection attempt runs
  uint32_t status;        // SPHB_Q_*
  uint32_t steps;         // steps completed since the block was seeded
  uint32_t attempts;      // advection attempts of the current / last step
  uint32_t budget;        // steps that may still start
  uint32_t fcap, rcap;    // capacities of the two cell tables (entries)
  uint32_t log_cap;       // entries of the timeleft log
};
// kernels_exact.cu / kernels_neighbors.cu each keep a __constant__ QueueHot; these enqueue the
// device-to-device refresh from ctl->hot
int queue_refresh_exact(const QueueCtl* ctl, cudaStream_t s);
int queue_refresh_neighbors(const QueueCtl* ctl, cudaStream_t s);
// one-thread kernels: before a step / after an advection attempt (`last`: no attempt follows)
int launch_queue_plan(QueueCtl* ctl, const Reduce* red, cudaStream_t s);
int launch_queue_attempt(QueueCtl* ctl, const Reduce* red, float* timeleft_log, int last, cudaStream_t s);

// Streaming multiprocessors of the CURRENT device (cudaDevAttrMultiProcessorCount, looked up once
// per device): grid-stride kernels size their grids as a multiple of it.
inline int sm_count() {
  static int cached[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (!cached[dev]) {
    int v = 0;
    cached[dev] = (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) ? v : 148;
  }
  return cached[dev];
}

// ---- launchers (each enqueues on `s` and returns the number of kernels launched) -------
// A launcher does not return CUDA errors itself: an enqueue-time failure of a kernel launch or of a
// cudaMemsetAsync inside it is recorded by the runtime and surfaces in the cudaGetLastError() that
// every caller issues after its launches (CUDA_TRY(e, cudaGetLastError()) in engine.cu / slab.cu).
// kernels_exact.cu  (compiled with -fmad=false: results are those of the literal
// expression trees, bit-identical to the CPU checker)
int launch_reduce_init(Reduce* red, cudaStream_t s, const QueueCtl* q = nullptr);
int launch_bbox(const float4* pos, uint32_t n, Reduce* red, cudaStream_t s);
int launch_locate(const float4* pos, uint32_t* keys, const StepConsts& c, cudaStream_t s);
// Counting sort by search cell (locate fused in): on return keys_sorted[d] = cell key of the
// particle at sorted position d, perm[d] = its index before the sort (stable: ascending inside a
// cell), ftable[f] = first sorted position of cell f (nfine + 1 entries).  keys_tmp, src_tmp:
// n-entry scratch; slot_tmp may alias perm (it is not used any more); src_tmp must not.
int launch_cell_sort(const float4* pos, const StepConsts& c, uint32_t* keys_tmp, uint32_t* slot_tmp,
                     uint32_t* ftable, uint32_t* scan_scratch, uint32_t* keys_sorted, uint32_t* src_tmp,
                     uint32_t* perm, cudaStream_t s, bool queued = false, size_t table_entries = 0,
                     uint8_t* cell_flags = nullptr);  // one byte per cell (table_entries of them), or nullptr
int launch_reorder(const uint32_t* perm, uint32_t n, const float4* pos_in, const float4* vel_in,
                   const float4* vhf_in, const uint32_t* rank_in, float4* pos_out, float4* vel_out,
                   float4* vhf_out, uint32_t* rank_out, cudaStream_t s, bool queued = false,
                   float4* pairs_out = nullptr);
// fine cell table: ftable[f] = index of the first sorted particle whose fine key is >= f,
// for f in [0, nfine]; ftable must hold nfine + 1 entries.  scratch: scan_scratch_entries().
int launch_fine_table(const uint32_t* sorted_keys, uint32_t n, uint32_t* ftable, uint32_t nfine,
                      uint32_t* scratch, cudaStream_t s);
// The reference's order.  Given the compute-ordered fine keys, the fine table and every
// particle's PREVIOUS reference rank (rank_prev, compute order), produces the reference's
// cell table (ref_table[c] = number of particles whose Morton key is < c, ncell entries), the
// Morton key of every particle (ckey) and every particle's NEW reference rank
// (= position in the array the reference would hold after its stable sort by Morton key).
int launch_reference_rank(const uint32_t* fine_keys, const uint32_t* ftable, const uint32_t* rank_prev,
                          uint32_t* ref_table, uint32_t* ckey, uint32_t* rank_new, const StepConsts& c,
                          uint32_t* scratch, cudaStream_t s, bool queued = false, size_t table_entries = 0);
// keys_ref[rank[i]] = ckey[i], perm_ref[rank[i]] = rank_prev[i]
int launch_reference_debug(const uint32_t* rank, const uint32_t* rank_prev, const uint32_t* ckey,
                           uint32_t n, uint32_t* keys_ref, uint32_t* perm_ref, cudaStream_t s);
int launch_advect(const float4* pos, const float4* vhf, const float4* acc, float4* pos_out,
                  float4* vel_out, float4* vhf_out, uint32_t n, float dt, float restitution,
                  const Face* faces, uint32_t face_count, const float* df, Reduce* red,
                  cudaStream_t s, const SlabAdvect* slab = nullptr, const FaceBin* bin = nullptr,
                  const QueueCtl* q = nullptr);
// cand_start / cand_faces: per owning face, the ascending list of faces whose box can hold one
// of its grid points (nullptr: every face is walked, as the reference does)
int launch_distance_field(float* df, const Face* faces, const uint32_t* face_offsets,
                          const float* transforms, const float* rvertices,
                          const uint32_t* cand_start, const uint32_t* cand_faces,
                          uint32_t face_count, uint32_t gridcount, cudaStream_t s);
// upload: particle i of the AoS array gets reference rank i
int launch_aos_to_soa(const sphb_particle* aos, uint32_t n, float4* pos, float4* vel, float4* vhf,
                      float4* acc, float* rho, float* prs, uint32_t* ckey, uint32_t* rank,
                      cudaStream_t s);
// download: particle i is written to aos[rank[i]] (the reference's order)
int launch_soa_to_aos(sphb_particle* aos, uint32_t n, const float4* pos, const float4* vel,
                      const float4* vhf, const float4* acc, const float* rho, const float* prs,
                      const uint32_t* ckey, const uint32_t* rank, cudaStream_t s);

// kernels_sort.cu — stable LSD radix sort of (key, index) pairs, 8 bits per pass.
struct SortBuffers {
  uint32_t* keys[2];     // n each; the unsorted keys are in keys[0]
  uint32_t* vals[2];     // n each
  uint32_t* hist;        // sort_hist_entries(n)
  uint32_t* block_sums;  // sort_block_sum_entries(n)
};
size_t sort_hist_entries(uint32_t n);
size_t sort_block_sum_entries(uint32_t n);
// Sorts over the low `key_bits` bits.  On return *result = r means: sorted keys are in
// keys[r] and vals[r][k] = original index of the element now at k.  Returns launches.
int launch_sort_pairs(const SortBuffers& b, uint32_t n, int key_bits, int* result,
                      cudaStream_t s);
// In-place exclusive scans of `len` u32 entries (sum, or running maximum with identity 0).
// scratch must hold scan_scratch_entries(len) entries.
size_t scan_scratch_entries(size_t len);
void set_scan_variant(int one_pass);  // process-wide: 1 = one-pass (decoupled look-back), 0 = three kernels
int launch_exclusive_sum(uint32_t* data, uint32_t len, uint32_t* scratch, cudaStream_t s);
int launch_exclusive_max(uint32_t* data, uint32_t len, uint32_t* scratch, cudaStream_t s);

// kernels_neighbors.cu (compiled with FMA contraction on; tolerance-checked)
// keys = compute-ordered fine keys, ftable = fine cell table (nfine + 1 entries).
// variant 0: the density pass records ONE neighbour list per pair of consecutive particles in
// `pool` (neighbor_pool_words(n, cap) 32-bit words, cap = neighbor_list_cap(...) entries per
// pair) that the forces pass consumes; variant 1: the round-1 density pass + forces by their own
// search; variant 2: the round-1 kernels (one list per particle); 4 / 5 / 6: the sub-warp-cooperative
// density pass with 4 / 8 / 16 lanes per particle + forces by their own search.  1, 2, 4-6 are kept
// for A/B measurements.
size_t neighbor_pool_words(uint32_t n, uint32_t cap);
uint32_t neighbor_list_cap(float h, float rho0, float mass);
// pairs = the reorder pass's pair records (n + 2 float4), pv = the forces pass's records (2 n float4)
int launch_density(float4* pos, float4* vel, const float4* pairs, float4* pv, const uint32_t* keys,
                   const uint32_t* ftable, float* rho, float* prs, uint32_t* pool, uint32_t cap,
                   const StepConsts& c, int variant, cudaStream_t s, bool queued = false);
int launch_forces(const float4* pos, const float4* vel, const float4* pv, const uint32_t* keys,
                  const uint32_t* ftable, const float* rho, const uint32_t* pool, float4* acc,
                  const StepConsts& c, int variant, cudaStream_t s,
                  const uint32_t* ghost_ids = nullptr, bool queued = false);

int launch_list_stats(const uint32_t* pool, uint32_t n, unsigned long long* out_dev, cudaStream_t s);

// kernels_slab.cu — the slab-decomposed (multi-GPU) step
size_t slab_message_bytes(uint32_t record_capacity);
int launch_slab_pack(const float4* pos, const float4* vel, const float4* vhf, uint32_t* ids,
                     const uint32_t* n_owned_dev, uint32_t n_upper, float xlo, float xhi, float halo,
                     int has_left, int has_right, void* msg_left, void* msg_right, uint32_t cap,
                     uint32_t* counters, cudaStream_t s, uint32_t p2p_seq = 0);
// peer-memory exchange: p2p_seq != 0 makes the pack publish its headers with that sequence number
// last; the receiver's stream waits for it on its own inboxes (nullptr = no such neighbour)
int launch_slab_wait(const void* inbox_left, const void* inbox_right, uint32_t seq, uint32_t* err,
                     unsigned long long timeout_ns, cudaStream_t s);
int launch_slab_unpack(const void* msg, uint32_t count, uint32_t offset, float4* pos, float4* vel,
                       float4* vhf, uint32_t* ids, cudaStream_t s);
// dst[i] = number of owned particles before i (compute order); *n_owned_out = their total
int launch_slab_owned_scan(const uint32_t* ids, uint32_t n, uint32_t* dst, uint32_t* scan_scratch,
                           uint32_t* n_owned_out, cudaStream_t s);
int launch_slab_to_aos(sphb_particle* aos, uint32_t* ids_out, const uint32_t* n_owned_dev, uint32_t n_upper,
                       const float4* pos, const float4* vel, const float4* vhf, const uint32_t* ids,
                       const uint32_t* src_of, const float4* acc, const float* rho, const float* prs,
                       const float4* pos_before, const uint32_t* ckey_uploaded, float gminx, float gminy,
                       float gminz, float cell, cudaStream_t s);
int launch_slab_reduce_floats(const Reduce* red, float* out, cudaStream_t s);
int launch_slab_bbox(const float4* pos, const uint32_t* n_owned_dev, uint32_t n_upper, Reduce* red,
                     cudaStream_t s);

}  // namespace sphb
