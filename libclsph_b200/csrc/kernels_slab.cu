// Kernels of the slab-decomposed (multi-GPU) step: classification of owned particles against
// the slab bounds into the two halo / migration messages, unpacking of received messages,
// and compaction of the owned particles after the step.
//
// The reference is single-device (libclsph/sph_simulation.cpp:273-275); this layer is new.
// One rank owns the particles whose x lies in [xlo, xhi).  Before every step each rank sends
// to each neighbour ONE message holding (a) the particles that crossed into that neighbour's
// slab (migrants: the receiver becomes their owner) and (b) copies of its own particles
// within `halo` (= 2h) of the common face (ghosts: the receiver uses them as neighbours
// only).  With a 2h layer the receiver can also compute the density of the ghosts within h
// of its slab, which its own particles' forces need, so one exchange per step suffices.
//
// Record = 48 bytes = 3 float4: (x, y, z, id) (vx, vy, vz, migrant ? 1 : 0) (v_half xyz, 0).
// Message = 16-byte header {count, migrants, 0, 0} + records.  Bit 31 of a particle's id
// marks a ghost inside the engine's arrays.
//
// Compiled with -fmad=false: the Morton grid_index written at download time must equal the
// reference's (kernels/grid.cl:40-64) bit for bit.
#include "common.cuh"

namespace sphb {

namespace {

constexpr int kBlock = 256;
constexpr uint32_t kGhostBit = 0x80000000u;

__device__ __forceinline__ uint32_t warp_append(uint32_t* counter, bool want) {
  // one atomic per warp; returns this lane's slot (undefined when !want)
  const unsigned mask = __ballot_sync(0xffffffffu, want);
  if (mask == 0) return 0;
  const int lane = threadIdx.x & 31;
  const int leader = __ffs(mask) - 1;
  uint32_t base = 0;
  if (lane == leader) base = atomicAdd(counter, __popc(mask));
  base = __shfl_sync(0xffffffffu, base, leader);
  return base + __popc(mask & ((1u << lane) - 1u));
}

__device__ __forceinline__ void write_record(float4* msg, uint32_t slot, const float4& p, const float4& v,
                                             const float4& w, uint32_t id, bool migrant) {
  float4* r = msg + 1 + 3 * (size_t)slot;  // + 1: the header
  r[0] = make_float4(p.x, p.y, p.z, __uint_as_float(id));
  r[1] = make_float4(v.x, v.y, v.z, migrant ? 1.f : 0.f);
  r[2] = make_float4(w.x, w.y, w.z, 0.f);
}

// counters: [0] records to the left, [1] migrants to the left, [2] records to the right,
// [3] migrants to the right.  A slot >= cap is counted but not written (the host then fails
// with SPHB_ERR_CAPACITY).
__global__ void __launch_bounds__(kBlock)
    k_slab_pack(const float4* __restrict__ pos, const float4* __restrict__ vel,
                const float4* __restrict__ vhf, uint32_t* __restrict__ ids,
                const uint32_t* __restrict__ n_owned, float xlo, float xhi, float halo, int has_left,
                int has_right, float4* __restrict__ msg_left, float4* __restrict__ msg_right,
                uint32_t cap, uint32_t* __restrict__ counters) {
  const uint32_t n = *n_owned;
  const uint32_t base = blockIdx.x * kBlock;
  if (base >= n) return;
  const uint32_t i = base + threadIdx.x;
  const bool live = i < n;
  float4 p = make_float4(0, 0, 0, 0);
  uint32_t id = 0;
  if (live) {
    p = __ldg(pos + i);
    id = ids[i] & ~kGhostBit;
  }
  const bool mig_r = live && has_right && p.x >= xhi;
  const bool mig_l = live && has_left && !mig_r && p.x < xlo;
  const bool to_r = live && has_right && (mig_r || (!mig_l && p.x >= xhi - halo));
  const bool to_l = live && has_left && (mig_l || (!mig_r && p.x < xlo + halo));
  const uint32_t sr = warp_append(counters + 2, to_r);
  const uint32_t sl = warp_append(counters + 0, to_l);
  warp_append(counters + 3, mig_r);
  warp_append(counters + 1, mig_l);
  if (to_r || to_l) {  // few particles are near a face: only they read the rest of their state
    const float4 v = __ldg(vel + i), w = __ldg(vhf + i);
    if (to_r && sr < cap) write_record(msg_right, sr, p, v, w, id, mig_r);
    if (to_l && sl < cap) write_record(msg_left, sl, p, v, w, id, mig_l);
  }
  // a particle that left stays here as a ghost of its new owner
  if (mig_r || mig_l) ids[i] = id | kGhostBit;
}

__global__ void k_slab_headers(const uint32_t* __restrict__ counters, float4* msg_left, float4* msg_right) {
  if (threadIdx.x == 0) {
    if (msg_left) *reinterpret_cast<uint4*>(msg_left) = make_uint4(counters[0], counters[1], 0, 0);
    if (msg_right) *reinterpret_cast<uint4*>(msg_right) = make_uint4(counters[2], counters[3], 0, 0);
  }
}

// ---- the exchange over peer memory -------------------------------------------------------
// With the neighbours' inboxes mapped into this GPU's address space (slab_nccl.cu: same-process
// peer access or CUDA IPC), k_slab_pack writes its records straight into them over NVLink and
// there is no send / receive at all: the header goes last, its third word — the step's
// sequence number — after a system-scope fence, and the receiver's stream waits for that number
// (k_slab_wait) before it reads the counts.
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__global__ void k_slab_headers_p2p(const uint32_t* __restrict__ counters, uint32_t* msg_left,
                                   uint32_t* msg_right, uint32_t seq) {
  if (threadIdx.x == 0) {
    if (msg_left) {
      volatile uint32_t* m = msg_left;
      m[0] = counters[0]; m[1] = counters[1]; m[3] = 0;
    }
    if (msg_right) {
      volatile uint32_t* m = msg_right;
      m[0] = counters[2]; m[1] = counters[3]; m[3] = 0;
    }
    __threadfence_system();  // the records (previous kernel) and the counts before the number
    if (msg_left) st_release_sys(msg_left + 2, seq);
    if (msg_right) st_release_sys(msg_right + 2, seq);
  }
}

// One thread polls the two local inboxes until both carry this step's number; gives up after
// timeout_ns (a neighbour that failed must not hang the device) and says so in *err.
__global__ void k_slab_wait(const uint32_t* inbox_left, const uint32_t* inbox_right, uint32_t seq,
                            uint32_t* err, unsigned long long timeout_ns) {
  if (threadIdx.x) return;
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  const uint32_t* boxes[2] = {inbox_left, inbox_right};
  for (int k = 0; k < 2; ++k) {
    if (!boxes[k]) continue;
    while (ld_acquire_sys(boxes[k] + 2) != seq) {
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      if (t - t0 > timeout_ns) {
        *err = 1u + (uint32_t)k;
        return;
      }
      __nanosleep(100);
    }
  }
}

// Appends the records of one received message at [offset, offset + count).
__global__ void __launch_bounds__(kBlock)
    k_slab_unpack(const float4* __restrict__ msg, uint32_t count, uint32_t offset, float4* __restrict__ pos,
                  float4* __restrict__ vel, float4* __restrict__ vhf, uint32_t* __restrict__ ids) {
  const uint32_t k = blockIdx.x * kBlock + threadIdx.x;
  if (k >= count) return;
  const float4* r = msg + 1 + 3 * (size_t)k;
  const float4 a = __ldg(r), b = __ldg(r + 1), c = __ldg(r + 2);
  const uint32_t id = __float_as_uint(a.w) & ~kGhostBit;
  pos[offset + k] = make_float4(a.x, a.y, a.z, 0.f);
  vel[offset + k] = make_float4(b.x, b.y, b.z, 0.f);
  vhf[offset + k] = make_float4(c.x, c.y, c.z, 0.f);
  ids[offset + k] = b.w != 0.f ? id : (id | kGhostBit);
}

// owned flags (1 = not a ghost) for the exclusive scan that gives every owned particle its
// dense output slot
__global__ void __launch_bounds__(kBlock)
    k_slab_owned_flags(const uint32_t* __restrict__ ids, uint32_t n, uint32_t* __restrict__ flags) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i < n) flags[i] = (__ldg(ids + i) & kGhostBit) ? 0u : 1u;
}
__global__ void k_slab_owned_count(const uint32_t* __restrict__ ids, const uint32_t* __restrict__ dst,
                                   uint32_t n, uint32_t* __restrict__ n_owned_out) {
  if (threadIdx.x == 0) *n_owned_out = n ? dst[n - 1] + ((ids[n - 1] & kGhostBit) ? 0u : 1u) : 0u;
}

// Owned particles -> 80-byte AoS (local order) + their ids.  grid_index is the reference's
// Morton key (grid.cl:40-64) of the position the last step STARTED from (`pos_before`, still
// intact in the other half of the ping-pong) under the GLOBAL grid origin of that step.
__global__ void __launch_bounds__(kBlock)
    k_slab_to_aos(sphb_particle* __restrict__ aos, uint32_t* __restrict__ ids_out,
                  const uint32_t* __restrict__ n_owned, const float4* __restrict__ pos,
                  const float4* __restrict__ vel, const float4* __restrict__ vhf,
                  const uint32_t* __restrict__ ids, const uint32_t* __restrict__ src_of,
                  const float4* __restrict__ acc, const float* __restrict__ rho,
                  const float* __restrict__ prs, const float4* __restrict__ pos_before,
                  const uint32_t* __restrict__ ckey_uploaded, float gminx, float gminy, float gminz,
                  float cell) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= *n_owned) return;
  float4 p = pos[i], v = vel[i], w = vhf[i];
  p.w = 0.f; v.w = 0.f; w.w = 0.f;
  float4 a = make_float4(0, 0, 0, 0);
  float d = 0.f, pr = 0.f;
  uint32_t key = 0;
  if (src_of) {
    const uint32_t s = src_of[i];
    a = acc[s]; a.w = 0.f;
    d = rho[s]; pr = prs[s];
    const float4 x0 = pos_before[s];
    key = morton3((uint32_t)((x0.x - gminx) / cell), (uint32_t)((x0.y - gminy) / cell),
                  (uint32_t)((x0.z - gminz) / cell));
  } else {  // no step since the upload: what was uploaded comes back
    a = acc[i]; a.w = 0.f;
    d = rho[i]; pr = prs[i];
    key = ckey_uploaded[i];
  }
  float4* dst = reinterpret_cast<float4*>(aos + i);
  dst[0] = p; dst[1] = v; dst[2] = w; dst[3] = a;
  dst[4] = make_float4(d, pr, __uint_as_float(key), 0.f);
  ids_out[i] = ids[i] & ~kGhostBit;
}

// local reduction cell (order-preserving encodings) -> floats the ranks can all-reduce with
// MAX: (-min x, -min y, -min z, max x, max y, max z, max |v|^2, max |a|^2)
__global__ void k_slab_reduce_floats(const Reduce* __restrict__ red, float* __restrict__ out) {
  if (threadIdx.x == 0) {
    for (int a = 0; a < 3; ++a) {
      out[a] = -dec_float(red->min_enc[a]);
      out[3 + a] = dec_float(red->max_enc[a]);
    }
    out[6] = dec_float(red->max_v2);
    out[7] = dec_float(red->max_a2);
  }
}

__global__ void __launch_bounds__(kBlock)
    k_slab_bbox(const float4* __restrict__ pos, const uint32_t* __restrict__ n_owned, Reduce* red) {
  const uint32_t n = *n_owned;
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (uint32_t i = blockIdx.x * kBlock + threadIdx.x; i < n; i += gridDim.x * kBlock) {
    const float4 p = __ldg(pos + i);
    mn[0] = fminf(mn[0], p.x); mn[1] = fminf(mn[1], p.y); mn[2] = fminf(mn[2], p.z);
    mx[0] = fmaxf(mx[0], p.x); mx[1] = fmaxf(mx[1], p.y); mx[2] = fmaxf(mx[2], p.z);
  }
  for (int a = 0; a < 3; ++a) {
    for (int o = 16; o > 0; o >>= 1) {
      mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
    if ((threadIdx.x & 31) == 0) {
      atomicMin(&red->min_enc[a], enc_float(mn[a]));
      atomicMax(&red->max_enc[a], enc_float(mx[a]));
    }
  }
}

inline unsigned blocks_for(uint32_t n) { return (n + kBlock - 1) / kBlock; }

}  // namespace

size_t slab_message_bytes(uint32_t record_capacity) { return 16 + (size_t)48 * record_capacity; }

int launch_slab_pack(const float4* pos, const float4* vel, const float4* vhf, uint32_t* ids,
                     const uint32_t* n_owned_dev, uint32_t n_upper, float xlo, float xhi, float halo,
                     int has_left, int has_right, void* msg_left, void* msg_right, uint32_t cap,
                     uint32_t* counters, cudaStream_t s, uint32_t p2p_seq) {
  cudaMemsetAsync(counters, 0, 4 * sizeof(uint32_t), s);
  if (n_upper)
    k_slab_pack<<<blocks_for(n_upper), kBlock, 0, s>>>(pos, vel, vhf, ids, n_owned_dev, xlo, xhi, halo,
                                                      has_left, has_right, (float4*)msg_left,
                                                      (float4*)msg_right, cap, counters);
  if (p2p_seq) k_slab_headers_p2p<<<1, 32, 0, s>>>(counters, (uint32_t*)msg_left, (uint32_t*)msg_right, p2p_seq);
  else k_slab_headers<<<1, 32, 0, s>>>(counters, (float4*)msg_left, (float4*)msg_right);
  return 2;
}

int launch_slab_wait(const void* inbox_left, const void* inbox_right, uint32_t seq, uint32_t* err,
                     unsigned long long timeout_ns, cudaStream_t s) {
  k_slab_wait<<<1, 32, 0, s>>>((const uint32_t*)inbox_left, (const uint32_t*)inbox_right, seq, err, timeout_ns);
  return 1;
}

int launch_slab_unpack(const void* msg, uint32_t count, uint32_t offset, float4* pos, float4* vel,
                       float4* vhf, uint32_t* ids, cudaStream_t s) {
  if (count == 0) return 0;
  k_slab_unpack<<<blocks_for(count), kBlock, 0, s>>>((const float4*)msg, count, offset, pos, vel, vhf, ids);
  return 1;
}

int launch_slab_owned_scan(const uint32_t* ids, uint32_t n, uint32_t* dst, uint32_t* scan_scratch,
                           uint32_t* n_owned_out, cudaStream_t s) {
  if (n == 0) {
    cudaMemsetAsync(n_owned_out, 0, sizeof(uint32_t), s);
    return 0;
  }
  k_slab_owned_flags<<<blocks_for(n), kBlock, 0, s>>>(ids, n, dst);
  int l = 1 + launch_exclusive_sum(dst, n, scan_scratch, s);
  k_slab_owned_count<<<1, 32, 0, s>>>(ids, dst, n, n_owned_out);
  return l + 1;
}

int launch_slab_to_aos(sphb_particle* aos, uint32_t* ids_out, const uint32_t* n_owned_dev, uint32_t n_upper,
                       const float4* pos, const float4* vel, const float4* vhf, const uint32_t* ids,
                       const uint32_t* src_of, const float4* acc, const float* rho, const float* prs,
                       const float4* pos_before, const uint32_t* ckey_uploaded, float gminx, float gminy,
                       float gminz, float cell, cudaStream_t s) {
  if (n_upper == 0) return 0;
  k_slab_to_aos<<<blocks_for(n_upper), kBlock, 0, s>>>(aos, ids_out, n_owned_dev, pos, vel, vhf, ids, src_of,
                                                      acc, rho, prs, pos_before, ckey_uploaded, gminx, gminy,
                                                      gminz, cell);
  return 1;
}

int launch_slab_reduce_floats(const Reduce* red, float* out, cudaStream_t s) {
  k_slab_reduce_floats<<<1, 32, 0, s>>>(red, out);
  return 1;
}

int launch_slab_bbox(const float4* pos, const uint32_t* n_owned_dev, uint32_t n_upper, Reduce* red,
                     cudaStream_t s) {
  unsigned grid = blocks_for(n_upper ? n_upper : 1);
  if (grid > (unsigned)sm_count() * 8u) grid = (unsigned)sm_count() * 8u;  // grid-stride: 8 resident CTAs of 256 threads per SM
  k_slab_bbox<<<grid, kBlock, 0, s>>>(pos, n_owned_dev, red);
  return 1;
}

}  // namespace sphb
