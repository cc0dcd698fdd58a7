// sphb_cell_table_noalloc — the reference's cell table from device-resident positions, WITHOUT any
// allocation (include/sphb.h "entry points that allocate nothing").  No counterpart in the
// reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// What it computes is what the engine's own path computes for sphb_download_debug's `cell_table`
// (kernels/grid.cl:40-64 locate, common/util.h:41-62 Morton key, kernels/sort.cl:61-78 cell table):
// cell_table[c] = number of particles whose Morton key is < c, for c in [0, grid_cell_count) — by a
// histogram into the caller's output array followed by an in-place exclusive scan.  The scan is the
// kernel that needs scratch memory (one partial sum per tile of 2048 entries); the engine's scan
// takes it from a cudaMalloc'ed buffer (kernels_sort.cu: exclusive_scan), this variant from the
// file-scope array below.
//
// The array is shared by every call in the process (one copy per GPU that loads the library, held
// until exit): calls that overlap on one GPU — two streams, two threads — race on it.  Calls on one
// stream are ordered and safe.  Every entry of the array that a call reads is written by that call
// first (k_na_tiles writes all `tiles` sums), so nothing needs clearing; the caller's output array
// is the histogram and IS cleared here.
//
// Compiled with -fmad=false like kernels_exact.cu: (p - min) / (2h) must not be contracted.
#include <mutex>

#include "engine_internal.cuh"

namespace {

constexpr int kNaThreads = 256;
constexpr int kNaTile = 2048;                              // entries per CTA of the scan
constexpr uint32_t kNaMaxCells = SPHB_NOALLOC_MAX_CELLS;   // 4 194 304 entries (include/sphb.h)
constexpr uint32_t kNaMaxTiles = kNaMaxCells / kNaTile;    // 2 048 partial sums = 8 KB

// The scratch: tile sums of the scan.  In HOST code this name is a stand-in, not a device address:
// the address is looked up per device with cudaGetSymbolAddress (na_scratch_address below).
static __device__ __align__(256) uint32_t g_na_tile_sums[kNaMaxTiles];

__global__ void __launch_bounds__(kNaThreads)
    k_na_histogram(const float4* __restrict__ pos, uint32_t n, float minx, float miny, float minz,
                   float cell, uint32_t gx, uint32_t gy, uint32_t gz, uint32_t ncell,
                   uint32_t* __restrict__ table) {
  const uint32_t i = blockIdx.x * kNaThreads + threadIdx.x;
  if (i >= n) return;
  const float4 p = __ldg(pos + i);
  // grid.cl:47-52: (uint)((p - min) / (h * 2)); NaN / runaway positions are parked in the last
  // cell, as the engine's own locate does (kernels_exact.cu: fine_key_of)
  const uint32_t cx = min((uint32_t)((p.x - minx) / cell), gx - 1u);
  const uint32_t cy = min((uint32_t)((p.y - miny) / cell), gy - 1u);
  const uint32_t cz = min((uint32_t)((p.z - minz) / cell), gz - 1u);
  const uint32_t key = morton3(cx, cy, cz);
  if (key + 1u < ncell) atomicAdd(table + key + 1u, 1u);  // table[c] counts keys == c - 1: its
                                                          // INCLUSIVE scan is the cell table
}

// Inclusive scan of kNaTile entries per CTA, in place; the CTA's total goes to sums[blockIdx.x].
__global__ void __launch_bounds__(kNaThreads)
    k_na_tiles(uint32_t* __restrict__ data, uint32_t len, uint32_t* __restrict__ sums) {
  constexpr int kPer = kNaTile / kNaThreads;  // 8 consecutive entries per thread
  __shared__ uint32_t warp_tot[kNaThreads / 32];
  const uint32_t base = blockIdx.x * kNaTile + threadIdx.x * kPer;
  uint32_t v[kPer], run = 0;
#pragma unroll
  for (int k = 0; k < kPer; ++k) {
    v[k] = base + k < len ? data[base + k] : 0u;
    run += v[k];
    v[k] = run;
  }
  // exclusive prefix of the threads' totals: warp shuffle scan, then the warps' totals
  uint32_t incl = run;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= (uint32_t)o) incl += up;
  }
  if (lane == 31u) warp_tot[warp] = incl;
  __syncthreads();
  uint32_t before = incl - run;
  for (uint32_t w = 0; w < warp; ++w) before += warp_tot[w];
#pragma unroll
  for (int k = 0; k < kPer; ++k)
    if (base + k < len) data[base + k] = v[k] + before;
  if (threadIdx.x == kNaThreads - 1) sums[blockIdx.x] = before + run;
}

// Exclusive scan of the tile sums by ONE CTA (tiles <= kNaMaxTiles = 2048 = 8 per thread).
__global__ void __launch_bounds__(kNaThreads) k_na_sums(uint32_t* __restrict__ sums, uint32_t tiles) {
  constexpr int kPer = kNaMaxTiles / kNaThreads;
  __shared__ uint32_t warp_tot[kNaThreads / 32];
  const uint32_t base = threadIdx.x * kPer;
  uint32_t v[kPer], run = 0;
#pragma unroll
  for (int k = 0; k < kPer; ++k) {
    const uint32_t x = base + k < tiles ? sums[base + k] : 0u;
    v[k] = run;  // exclusive inside the thread
    run += x;
  }
  uint32_t incl = run;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= (uint32_t)o) incl += up;
  }
  if (lane == 31u) warp_tot[warp] = incl;
  __syncthreads();
  uint32_t before = incl - run;
  for (uint32_t w = 0; w < warp; ++w) before += warp_tot[w];
#pragma unroll
  for (int k = 0; k < kPer; ++k)
    if (base + k < tiles) sums[base + k] = v[k] + before;
}

__global__ void __launch_bounds__(kNaThreads)
    k_na_apply(uint32_t* __restrict__ data, uint32_t len, const uint32_t* __restrict__ sums) {
  const uint32_t off = sums[blockIdx.x];
#pragma unroll
  for (int k = 0; k < kNaTile / kNaThreads; ++k) {
    const uint32_t i = blockIdx.x * kNaTile + k * kNaThreads + threadIdx.x;
    if (i < len) data[i] += off;
  }
}

// Device address of g_na_tile_sums on the CURRENT device, looked up once per device.
cudaError_t na_scratch_address(int dev, uint32_t** out) {
  static std::mutex mu;
  static uint32_t* addr[64] = {};
  std::lock_guard<std::mutex> lock(mu);
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  if (!addr[dev]) {
    void* p = nullptr;
    const cudaError_t err = cudaGetSymbolAddress(&p, g_na_tile_sums);
    if (err != cudaSuccess) return err;
    addr[dev] = static_cast<uint32_t*>(p);
  }
  *out = addr[dev];
  return cudaSuccess;
}

}  // namespace

extern "C" int sphb_cell_table_noalloc(const float* positions_dev, uint32_t n, const sphb_params* params,
                                       uint32_t* cell_table_dev, uint32_t cell_table_entries,
                                       void* stream) {
  sphb_engine* const none = nullptr;
  if (!positions_dev || !params || !cell_table_dev)
    SPHB_FAIL(none, SPHB_ERR_INVALID, "sphb_cell_table_noalloc: null argument");
  if (((uintptr_t)positions_dev & 15u) || ((uintptr_t)cell_table_dev & 3u))
    SPHB_FAIL(none, SPHB_ERR_INVALID, "sphb_cell_table_noalloc: positions must be 16-byte aligned (x, y, z, - per "
              "particle), the table 4-byte aligned");
  const uint32_t ncell = params->grid_cell_count;
  if (params->grid_size_x <= 0 || params->grid_size_y <= 0 || params->grid_size_z <= 0 ||
      params->grid_size_x >= 1024 || params->grid_size_y >= 1024 || params->grid_size_z >= 1024 || ncell == 0)
    SPHB_FAIL(none, SPHB_ERR_GRID, "sphb_cell_table_noalloc: grid %d x %d x %d, %u cells: not a grid the reference "
              "accepts (sph_simulation.cpp:722-724)", params->grid_size_x, params->grid_size_y, params->grid_size_z,
              ncell);
  if (ncell > kNaMaxCells)
    SPHB_FAIL(none, SPHB_ERR_CAPACITY, "sphb_cell_table_noalloc: grid_cell_count %u exceeds SPHB_NOALLOC_MAX_CELLS = "
              "%u (the static scratch holds one partial sum per 2048 cells); use an engine (sphb_create) instead",
              ncell, kNaMaxCells);
  if (cell_table_entries < ncell)
    SPHB_FAIL(none, SPHB_ERR_CAPACITY, "sphb_cell_table_noalloc: the output holds %u entries, grid_cell_count is %u",
              cell_table_entries, ncell);
  int dev = 0;
  CUDA_TRY(none, cudaGetDevice(&dev));
  uint32_t* sums = nullptr;
  CUDA_TRY(none, na_scratch_address(dev, &sums));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const float cell = params->h * 2;
  const uint32_t tiles = (ncell + kNaTile - 1) / kNaTile;
  CUDA_TRY(none, cudaMemsetAsync(cell_table_dev, 0, sizeof(uint32_t) * ncell, s));
  if (n)
    k_na_histogram<<<(n + kNaThreads - 1) / kNaThreads, kNaThreads, 0, s>>>(
        reinterpret_cast<const float4*>(positions_dev), n, params->min_point.s[0], params->min_point.s[1],
        params->min_point.s[2], cell, (uint32_t)params->grid_size_x, (uint32_t)params->grid_size_y,
        (uint32_t)params->grid_size_z, ncell, cell_table_dev);
  k_na_tiles<<<tiles, kNaThreads, 0, s>>>(cell_table_dev, ncell, sums);
  if (tiles > 1) {
    k_na_sums<<<1, kNaThreads, 0, s>>>(sums, tiles);
    k_na_apply<<<tiles, kNaThreads, 0, s>>>(cell_table_dev, ncell, sums);
  }
  CUDA_TRY(none, cudaGetLastError());
  return 0;
}
