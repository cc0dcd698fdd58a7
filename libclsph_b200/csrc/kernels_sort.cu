// Stable least-significant-digit radix sort of (Morton key, particle index) pairs.
//
// Replaces the reference's sort_particles (libclsph/sph_simulation.cpp:110-198 with
// kernels/sort.cl:23-59 and the Blelloch scans :108-200), which moves whole 80-byte
// particles through 4 passes with at most 1024 work-items.  The observable contract is
// "output = stable sort of the input order by grid_index"; it is independent of the
// reference's thread count, so any stable sort reproduces its permutation bit for bit.
//
// Per 8-bit pass: (1) per-tile digit histogram, (2) exclusive scan of the
// [digit][tile] table (two kernels; the third — adding the scanned block sums — is folded
// into the scatter), (3) stable scatter: every warp owns a contiguous chunk of the tile,
// ranks its elements with __match_any_sync, a 256-thread scan across warps turns per-warp
// digit counts into offsets.  Only as many passes as the key has significant bits are run
// (3 for grids up to 128 cells per axis, not 4).
#include "common.cuh"

namespace sphb {

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kRounds = 16;                       // 32-element rounds per warp
constexpr int kWarpChunk = 32 * kRounds;          // 512
constexpr int kTile = kWarps * kWarpChunk;        // 4096 elements per CTA
constexpr int kScanTile = 2048;                   // entries per CTA in the table scan

__global__ void __launch_bounds__(kThreads)
    k_radix_hist(const uint32_t* __restrict__ keys, uint32_t n, int shift,
                 uint32_t* __restrict__ hist, uint32_t ntiles) {
  __shared__ uint32_t cnt[256];
  cnt[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t base = blockIdx.x * kTile;
  const int lane = threadIdx.x & 31;
#pragma unroll 4
  for (int r = 0; r < kTile / kThreads; ++r) {
    const uint32_t i = base + r * kThreads + threadIdx.x;
    const bool valid = i < n;
    const uint32_t d = valid ? ((__ldg(keys + i) >> shift) & 0xFFu) : 0x100u;
    const unsigned peers = __match_any_sync(0xffffffffu, d);
    if (valid && lane == __ffs(peers) - 1) atomicAdd(&cnt[d], __popc(peers));
  }
  __syncthreads();
  hist[threadIdx.x * ntiles + blockIdx.x] = cnt[threadIdx.x];
}

// In-place exclusive scan of kScanTile entries per CTA; CTA totals go to sums[].
__global__ void __launch_bounds__(kThreads)
    k_scan_tiles(uint32_t* __restrict__ data, uint32_t len, uint32_t* __restrict__ sums) {
  constexpr int kPer = kScanTile / kThreads;  // 8
  __shared__ uint32_t warp_tot[kWarps];
  const uint32_t base = blockIdx.x * kScanTile + threadIdx.x * kPer;
  uint32_t v[kPer];
  uint32_t t = 0;
#pragma unroll
  for (int k = 0; k < kPer; ++k) {
    v[k] = (base + k < len) ? data[base + k] : 0u;
    t += v[k];
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = t;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += u;
  }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  uint32_t woff = 0;
#pragma unroll
  for (int w = 0; w < kWarps; ++w)
    if (w < warp) woff += warp_tot[w];
  uint32_t run = woff + incl - t;
#pragma unroll
  for (int k = 0; k < kPer; ++k) {
    if (base + k < len) data[base + k] = run;
    run += v[k];
  }
  if (threadIdx.x == kThreads - 1) sums[blockIdx.x] = run;
}

// Single-CTA exclusive scan of the per-tile totals (any length).
__global__ void __launch_bounds__(kThreads) k_scan_sums(uint32_t* __restrict__ sums, uint32_t len) {
  __shared__ uint32_t warp_tot[kWarps];
  __shared__ uint32_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t start = 0; start < len; start += kThreads) {
    const uint32_t i = start + threadIdx.x;
    const uint32_t v = i < len ? sums[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += u;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    uint32_t woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      if (w < warp) woff += warp_tot[w];
      total += warp_tot[w];
    }
    const uint32_t carry = carry_s;
    if (i < len) sums[i] = carry + woff + incl - v;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + total;
    __syncthreads();
  }
}

// Stable scatter of one tile.  vals_in == nullptr means "identity" (first pass).
__global__ void __launch_bounds__(kThreads)
    k_radix_scatter(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                    uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n,
                    int shift, const uint32_t* __restrict__ hist,
                    const uint32_t* __restrict__ sums, uint32_t ntiles) {
  __shared__ uint32_t warp_cnt[kWarps][256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int w = 0; w < kWarps; ++w) warp_cnt[w][threadIdx.x] = 0;
  __syncthreads();

  const uint32_t chunk = blockIdx.x * kTile + warp * kWarpChunk;
  uint32_t key[kRounds];
  uint32_t rank[kRounds];  // rank of the element among equal digits inside this warp's chunk
#pragma unroll
  for (int r = 0; r < kRounds; ++r) {
    const uint32_t i = chunk + r * 32 + lane;
    const bool valid = i < n;
    key[r] = valid ? __ldg(keys_in + i) : 0u;
    const uint32_t d = valid ? ((key[r] >> shift) & 0xFFu) : 0x100u;
    const unsigned peers = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(peers) - 1;
    uint32_t before = 0;
    if (valid && lane == leader) {
      before = warp_cnt[warp][d];
      warp_cnt[warp][d] = before + __popc(peers);
    }
    before = __shfl_sync(0xffffffffu, before, leader);
    rank[r] = before + __popc(peers & ((1u << lane) - 1u));
    __syncwarp();
  }
  __syncthreads();

  {  // digit `threadIdx.x`: global base of this tile, then offsets of each warp's run
    const uint32_t d = threadIdx.x;
    const uint32_t e = d * ntiles + blockIdx.x;
    uint32_t run = __ldg(hist + e) + __ldg(sums + e / kScanTile);
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const uint32_t c = warp_cnt[w][d];
      warp_cnt[w][d] = run;
      run += c;
    }
  }
  __syncthreads();

#pragma unroll
  for (int r = 0; r < kRounds; ++r) {
    const uint32_t i = chunk + r * 32 + lane;
    if (i < n) {
      const uint32_t d = (key[r] >> shift) & 0xFFu;
      const uint32_t dst = warp_cnt[warp][d] + rank[r];
      keys_out[dst] = key[r];
      vals_out[dst] = vals_in ? __ldg(vals_in + i) : i;
    }
  }
}

// ---- generic in-place exclusive scans (sum / running maximum) ---------------------------
struct OpSum {
  __device__ static uint32_t apply(uint32_t a, uint32_t b) { return a + b; }
};
struct OpMax {
  __device__ static uint32_t apply(uint32_t a, uint32_t b) { return a > b ? a : b; }
};

template <typename Op>
__device__ __forceinline__ uint32_t warp_inclusive(uint32_t v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v = Op::apply(v, u);
  }
  return v;
}

// exclusive scan of kScanTile entries per CTA; CTA totals go to sums[]
template <typename Op>
__global__ void __launch_bounds__(kThreads)
    k_xscan_tiles(uint32_t* __restrict__ data, uint32_t len, uint32_t* __restrict__ sums) {
  constexpr int kPer = kScanTile / kThreads;
  __shared__ uint32_t warp_tot[kWarps];
  const uint32_t base = blockIdx.x * kScanTile + threadIdx.x * kPer;
  uint32_t v[kPer];
  uint32_t t = 0;
#pragma unroll
  for (int k = 0; k < kPer; ++k) {
    v[k] = (base + k < len) ? data[base + k] : 0u;
    t = Op::apply(t, v[k]);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t incl = warp_inclusive<Op>(t, lane);
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  uint32_t run = 0;
#pragma unroll
  for (int w = 0; w < kWarps; ++w)
    if (w < warp) run = Op::apply(run, warp_tot[w]);
  const uint32_t prev_lane = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane > 0) run = Op::apply(run, prev_lane);
#pragma unroll
  for (int k = 0; k < kPer; ++k) {
    if (base + k < len) data[base + k] = run;
    run = Op::apply(run, v[k]);
  }
  if (threadIdx.x == kThreads - 1) sums[blockIdx.x] = run;
}

// single-CTA exclusive scan of the CTA totals (any length)
template <typename Op>
__global__ void __launch_bounds__(kThreads) k_xscan_sums(uint32_t* __restrict__ sums, uint32_t len) {
  __shared__ uint32_t warp_tot[kWarps];
  __shared__ uint32_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t start = 0; start < len; start += kThreads) {
    const uint32_t i = start + threadIdx.x;
    const uint32_t v = i < len ? sums[i] : 0u;
    const uint32_t incl = warp_inclusive<Op>(v, lane);
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    uint32_t run = carry_s, total = carry_s;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      if (w < warp) run = Op::apply(run, warp_tot[w]);
      total = Op::apply(total, warp_tot[w]);
    }
    const uint32_t prev_lane = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane > 0) run = Op::apply(run, prev_lane);
    if (i < len) sums[i] = run;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = total;
    __syncthreads();
  }
}

template <typename Op>
__global__ void __launch_bounds__(kThreads)
    k_xscan_apply(uint32_t* __restrict__ data, uint32_t len, const uint32_t* __restrict__ sums) {
  const uint32_t off = __ldg(sums + blockIdx.x);
  const uint32_t base = blockIdx.x * kScanTile;
#pragma unroll
  for (int k = 0; k < kScanTile / kThreads; ++k) {
    const uint32_t i = base + k * kThreads + threadIdx.x;
    if (i < len) data[i] = Op::apply(data[i], off);
  }
}

template <typename Op>
int exclusive_scan(uint32_t* data, uint32_t len, uint32_t* scratch, cudaStream_t s) {
  if (len == 0) return 0;
  const uint32_t tiles = (len + kScanTile - 1) / kScanTile;
  k_xscan_tiles<Op><<<tiles, kThreads, 0, s>>>(data, len, scratch);
  if (tiles == 1) return 1;
  k_xscan_sums<Op><<<1, kThreads, 0, s>>>(scratch, tiles);
  k_xscan_apply<Op><<<tiles, kThreads, 0, s>>>(data, len, scratch);
  return 3;
}

// ---- the same scans in ONE pass over the data (decoupled look-back) ------------------------
// A CTA takes the next tile by ticket (so every tile before it has been started), scans its
// kLookTile entries in registers, publishes the tile's aggregate as one 64-bit word (flag, value),
// and its first warp walks back over the preceding tiles' words — 32 at a time — until it meets
// one that already carries an inclusive prefix.  One read and one write of the table instead of
// two reads and two writes, one launch instead of three.  status[] and the ticket are zeroed by
// the launcher.  A predecessor that does not answer within 60 s is a bug: the kernel traps (the
// next CUDA call fails) rather than spinning for ever.
constexpr int kLookPer = 16;                       // entries per thread: 4 x uint4, warp-striped
constexpr int kLookTile = kThreads * kLookPer;     // 4096 entries per CTA
constexpr uint32_t kFlagAggregate = 1u, kFlagInclusive = 2u;

__device__ __forceinline__ unsigned long long look_pack(uint32_t flag, uint32_t value) {
  return ((unsigned long long)flag << 32) | value;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

template <typename Op>
__global__ void __launch_bounds__(kThreads)
    k_scan_lookback(uint32_t* __restrict__ data, uint32_t len, volatile unsigned long long* status,
                    uint32_t* ticket) {
  __shared__ uint32_t tile_s, prefix_s;
  __shared__ uint32_t warp_tot[kWarps];
  if (threadIdx.x == 0) tile_s = atomicAdd(ticket, 1u);
  __syncthreads();
  const uint32_t tile = tile_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // warp w owns 512 consecutive entries: 4 chunks of 128, lane l holds entries 4l .. 4l+3 of each
  const uint32_t wbase = tile * (uint32_t)kLookTile + (uint32_t)warp * (32u * kLookPer);
  uint4 v[4];
  uint32_t excl[4], chunk_tot[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const uint32_t i = wbase + r * 128u + lane * 4u;
    if (i + 3u < len) {
      v[r] = *reinterpret_cast<const uint4*>(data + i);
    } else {
      v[r].x = i < len ? data[i] : 0u;
      v[r].y = i + 1u < len ? data[i + 1u] : 0u;
      v[r].z = i + 2u < len ? data[i + 2u] : 0u;
      v[r].w = 0u;
    }
    const uint32_t t = Op::apply(Op::apply(v[r].x, v[r].y), Op::apply(v[r].z, v[r].w));
    const uint32_t incl = warp_inclusive<Op>(t, lane);
    const uint32_t up = __shfl_up_sync(0xffffffffu, incl, 1);
    excl[r] = lane ? up : 0u;
    chunk_tot[r] = __shfl_sync(0xffffffffu, incl, 31);
  }
  uint32_t chunk_off[4];
  chunk_off[0] = 0u;
#pragma unroll
  for (int r = 1; r < 4; ++r) chunk_off[r] = Op::apply(chunk_off[r - 1], chunk_tot[r - 1]);
  if (lane == 0) warp_tot[warp] = Op::apply(chunk_off[3], chunk_tot[3]);
  __syncthreads();
  uint32_t warp_off = 0u, aggregate = 0u;
#pragma unroll
  for (int w = 0; w < kWarps; ++w) {
    if (w < warp) warp_off = Op::apply(warp_off, warp_tot[w]);
    aggregate = Op::apply(aggregate, warp_tot[w]);
  }
  if (warp == 0) {
    uint32_t prefix = 0u;
    if (tile == 0u) {
      if (lane == 0) status[0] = look_pack(kFlagInclusive, aggregate);
    } else {
      if (lane == 0) status[tile] = look_pack(kFlagAggregate, aggregate);
      const unsigned long long t0 = globaltimer_ns();
      long long first = (long long)tile - 1;  // nearest tile of the window looked at
      for (;;) {
        const long long mine = first - lane;
        unsigned long long word = look_pack(kFlagInclusive, 0u);  // before tile 0: nothing
        if (mine >= 0) {
          word = status[mine];
          while ((uint32_t)(word >> 32) == 0u) {
            // %globaltimer is wall clock: it keeps running while the context is time-sliced out
            // (other processes on the GPU, a profiler's replay), so the bound is generous — a
            // predecessor silent for a whole minute is a bug, and a trap beats a hung device
            if (globaltimer_ns() - t0 > 60000000000ull) __trap();
            __nanosleep(40);
            word = status[mine];
          }
        }
        const uint32_t has_incl = __ballot_sync(0xffffffffu, (uint32_t)(word >> 32) == kFlagInclusive);
        // lanes up to and including the nearest inclusive one contribute
        const int stop = has_incl ? __ffs((int)has_incl) - 1 : 31;
        uint32_t part = lane <= stop ? (uint32_t)word : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) part = Op::apply(part, __shfl_xor_sync(0xffffffffu, part, o));
        prefix = Op::apply(prefix, part);
        if (has_incl) break;
        first -= 32;
      }
      if (lane == 0) status[tile] = look_pack(kFlagInclusive, Op::apply(prefix, aggregate));
    }
    if (lane == 0) prefix_s = prefix;
  }
  __syncthreads();
  const uint32_t base_off = Op::apply(prefix_s, warp_off);
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const uint32_t i = wbase + r * 128u + lane * 4u;
    uint4 o;
    o.x = Op::apply(base_off, Op::apply(chunk_off[r], excl[r]));
    o.y = Op::apply(o.x, v[r].x);
    o.z = Op::apply(o.y, v[r].y);
    o.w = Op::apply(o.z, v[r].z);
    if (i + 3u < len) {
      *reinterpret_cast<uint4*>(data + i) = o;
    } else {
      if (i < len) data[i] = o.x;
      if (i + 1u < len) data[i + 1u] = o.y;
      if (i + 2u < len) data[i + 2u] = o.z;
    }
  }
}

int g_scan_one_pass = 1;  // process-wide: sphb_set_option("scan", 0) selects the three-kernel scan

template <typename Op>
int scan_in_place(uint32_t* data, uint32_t len, uint32_t* scratch, cudaStream_t s) {
  // (the one-pass kernel wants 16-byte aligned data and 8-byte aligned scratch: cudaMalloc'ed bases are)
  if (!g_scan_one_pass || len <= (uint32_t)kLookTile || ((uintptr_t)data & 15u) || ((uintptr_t)scratch & 7u))
    return exclusive_scan<Op>(data, len, scratch, s);
  const uint32_t tiles = (len + kLookTile - 1) / kLookTile;
  cudaMemsetAsync(scratch, 0, sizeof(unsigned long long) * ((size_t)tiles + 1), s);
  unsigned long long* status = reinterpret_cast<unsigned long long*>(scratch);
  k_scan_lookback<Op><<<tiles, kThreads, 0, s>>>(data, len, status, reinterpret_cast<uint32_t*>(status + tiles));
  return 1;
}

inline uint32_t tiles_for(uint32_t n) { return (n + kTile - 1) / kTile; }

}  // namespace

void set_scan_variant(int one_pass) { g_scan_one_pass = one_pass != 0; }

size_t sort_hist_entries(uint32_t n) { return (size_t)256 * (tiles_for(n) ? tiles_for(n) : 1); }
size_t sort_block_sum_entries(uint32_t n) {
  return (sort_hist_entries(n) + kScanTile - 1) / kScanTile;
}

// the three-kernel scan's CTA totals, or the one-pass scan's 64-bit status words + ticket
size_t scan_scratch_entries(size_t len) {
  const size_t three = (len + kScanTile - 1) / kScanTile + 1;
  const size_t one = 2 * ((len + kLookTile - 1) / kLookTile + 2);
  return three > one ? three : one;
}
int launch_exclusive_sum(uint32_t* data, uint32_t len, uint32_t* scratch, cudaStream_t s) {
  return scan_in_place<OpSum>(data, len, scratch, s);
}
int launch_exclusive_max(uint32_t* data, uint32_t len, uint32_t* scratch, cudaStream_t s) {
  return scan_in_place<OpMax>(data, len, scratch, s);
}

int launch_sort_pairs(const SortBuffers& b, uint32_t n, int key_bits, int* result, cudaStream_t s) {
  *result = 0;
  if (n == 0) return 0;
  if (key_bits < 1) key_bits = 1;
  if (key_bits > 32) key_bits = 32;
  const int passes = (key_bits + 7) / 8;
  const uint32_t ntiles = tiles_for(n);
  const uint32_t hist_len = 256u * ntiles;
  const uint32_t scan_tiles = (hist_len + kScanTile - 1) / kScanTile;
  int launches = 0;
  for (int p = 0; p < passes; ++p) {
    const int in = p & 1, out = in ^ 1, shift = 8 * p;
    k_radix_hist<<<ntiles, kThreads, 0, s>>>(b.keys[in], n, shift, b.hist, ntiles);
    k_scan_tiles<<<scan_tiles, kThreads, 0, s>>>(b.hist, hist_len, b.block_sums);
    k_scan_sums<<<1, kThreads, 0, s>>>(b.block_sums, scan_tiles);
    k_radix_scatter<<<ntiles, kThreads, 0, s>>>(b.keys[in], p == 0 ? nullptr : b.vals[in],
                                                b.keys[out], b.vals[out], n, shift, b.hist,
                                                b.block_sums, ntiles);
    launches += 4;
  }
  *result = passes & 1;
  return launches;
}

}  // namespace sphb
