// sphb engine: host orchestration of one SPH step on one B200 + the extern "C" boundary
// declared in include/sphb.h.  Restates libclsph/sph_simulation.cpp:200-263
// (simulate_single_frame), :110-198 (sort_particles), :634-728 (findMinMaxPosition),
// :536-605 (computeTimeStep) and :607-632 (computeDistanceField) as a sequence of CUDA
// launches on one stream with ONE host read-back per step (32 bytes: max|v|^2, max|a|^2 and
// the bbox of the advected positions, which is the next step's grid), instead of the
// reference's four blocking reads and five buffer allocations per step.
//
// There is no CPU path in this file: every pass is a kernel from kernels_*.cu.



#include <atomic>

#include "engine_internal.cuh"

namespace {
constexpr uint32_t kQueueLog = 4096;  // timeleft log entries = most steps one batch may hold
// Device-driven stepping (option "device_loop" = 2: automatic) is used up to this many particles.
// Measured on a B200 (tools/ab_options.py, ms per step, host round trip per step -> device loop):
// 0.355 -> 0.324 at 1 M particles, 1.169 -> 1.147 at 4.1 M, and with the round-2 kernels 3.86 -> 3.80
// at 16.7 M (round 1 measured a small loss there, when the whole table capacity was scanned in three passes).
constexpr uint32_t kQueueAutoMax = 1u << 26;

// Device-driven steps read their per-step constants from ONE __constant__ copy per translation
// unit (g_hot in kernels_exact.cu / kernels_neighbors.cu), refreshed on the engine's stream: two
// engines of one process stepping device-driven on the same GPU at the same time would overwrite
// each other's grid.  So one engine at a time per device holds this lease for the length of a
// batch (which ends in a stream synchronisation); an engine that finds it taken steps from the host
// for that batch (sphb_step passes its constants as kernel arguments and is not affected).
std::atomic<sphb_engine*> g_queue_owner[64];
struct QueueLease {
  std::atomic<sphb_engine*>* slot = nullptr;
  bool held = false;
  QueueLease(sphb_engine* e, bool want) {
    if (!want || e->device < 0 || e->device >= 64) return;
    slot = &g_queue_owner[e->device];
    sphb_engine* nobody = nullptr;
    held = slot->compare_exchange_strong(nobody, e);
  }
  ~QueueLease() {
    if (held) slot->store(nullptr);
  }
};
}  // namespace

std::string& sphb_create_error_slot() {
  static thread_local std::string slot;
  return slot;
}

extern "C" {

const char* sphb_error_string(int code) {
  if (!g_create_error.empty()) return g_create_error.c_str();
  switch (code) {
    case 0: return "ok";
    case SPHB_ERR_INVALID: return "invalid argument or call order";
    case SPHB_ERR_GRID: return "grid dimension reached 1024";
    case SPHB_ERR_NO_DEVICE: return "no usable CUDA device";
    case SPHB_ERR_NO_SCENE: return "no scene set";
    case SPHB_ERR_NCCL_MISSING: return "libnccl could not be loaded";
    default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "unknown sphb error";
  }
}

const char* sphb_last_error(sphb_engine* e) { return e ? e->err.c_str() : g_create_error.c_str(); }

int sphb_destroy(sphb_engine* e) {
  if (!e) return 0;
  sphb_slab_comm_release(e);
  cudaSetDevice(e->device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  for (int k = 0; k < 2; ++k) {
    cudaFree(e->pos[k]); cudaFree(e->vel[k]); cudaFree(e->vhf[k]);
    cudaFree(e->sortb.keys[k]); cudaFree(e->sortb.vals[k]);
  }
  cudaFree(e->acc); cudaFree(e->rho); cudaFree(e->prs);
  cudaFree(e->sortb.hist); cudaFree(e->sortb.block_sums);
  cudaFree(e->rank_cur); cudaFree(e->rank_prev); cudaFree(e->ckey);
  cudaFree(e->npool); cudaFree(e->pairs); cudaFree(e->pv);
  cudaFree(e->n_owned_dev); cudaFree(e->counters); cudaFree(e->src_of);
  if (e->host_counts) cudaFreeHost(e->host_counts);
  cudaFree(e->ftable); cudaFree(e->ref_table); cudaFree(e->scan_scratch); cudaFree(e->aos_stage);
  cudaFree(e->faces); cudaFree(e->face_offsets); cudaFree(e->df);
  cudaFree(e->bin_start); cudaFree(e->bin_faces);
  cudaFree(e->red);
  if (e->red_host) cudaFreeHost(e->red_host);
  if (e->qgraph) cudaGraphExecDestroy(e->qgraph);
  cudaFree(e->qctl); cudaFree(e->qlog);
  if (e->qctl_host) cudaFreeHost(e->qctl_host);
  if (e->qlog_host) cudaFreeHost(e->qlog_host);
  if (e->seg_created)
    for (int k = 0; k < 9; ++k) cudaEventDestroy(e->seg_ev[k]);
  if (e->ev_created)
    for (int p = 0; p <= SPHB_PASS_COUNT; ++p) {
      cudaEventDestroy(e->ev[p][0]);
      cudaEventDestroy(e->ev[p][1]);
      if (p < 4) cudaEventDestroy(e->user_ev[p]);
    }
  if (e->ev_fork) cudaEventDestroy(e->ev_fork);
  if (e->ev_join) cudaEventDestroy(e->ev_join);
  if (e->side) cudaStreamDestroy(e->side);
  if (e->ev_staged) cudaEventDestroy(e->ev_staged);
  if (e->ev_copied) cudaEventDestroy(e->ev_copied);
  if (e->copy) cudaStreamDestroy(e->copy);
  if (e->ev_up_done) cudaEventDestroy(e->ev_up_done);
  if (e->ev_up_consumed) cudaEventDestroy(e->ev_up_consumed);
  if (e->copy_up) cudaStreamDestroy(e->copy_up);
  cudaFree(e->aos_up);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
  return 0;
}

int sphb_create(const sphb_params* params, const sphb_terms* terms, int device, sphb_engine** out) {
  return sphb_create_with_capacity(params, terms, device, params ? params->particles_count : 0, out);
}

int sphb_create_with_capacity(const sphb_params* params, const sphb_terms* terms, int device,
                              uint32_t capacity, sphb_engine** out) {
  g_create_error.clear();
  if (!params || !terms || !out) SPHB_FAIL(nullptr, SPHB_ERR_INVALID, "null argument");
  *out = nullptr;
  if (params->particles_count == 0) SPHB_FAIL(nullptr, SPHB_ERR_INVALID, "particles_count == 0");
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
    cudaGetLastError();
    SPHB_FAIL(nullptr, SPHB_ERR_NO_DEVICE, "no CUDA device (this engine has no CPU path)");
  }
  if (device < 0) {
    if (cudaGetDevice(&device) != cudaSuccess) device = 0;
  }
  if (device >= count) SPHB_FAIL(nullptr, SPHB_ERR_NO_DEVICE, "device %d of %d", device, count);

  sphb_engine* e = new sphb_engine();
  e->device = device;
  e->params = *params;
  e->terms = *terms;
  e->n = params->particles_count;
  e->cap = capacity ? capacity : e->n;
  {  // smoothing.cl:24: -45.f / convert_float(M_PI * pown(h, 6))
    float h6 = params->h;
    for (int k = 1; k < 6; ++k) h6 = h6 * params->h;
    e->spiky_eps = -45.f / (float)(M_PI * h6);
  }
  const uint32_t n = e->cap;
#define TRY_OR_DESTROY(call)                                                           \
  do {                                                                                 \
    cudaError_t err__ = (call);                                                        \
    if (err__ != cudaSuccess) {                                                        \
      int rc = fail(nullptr, (int)err__, __FILE__, __LINE__, "%s -> %s", #call,        \
                    cudaGetErrorString(err__));                                        \
      sphb_destroy(e);                                                                 \
      return rc;                                                                       \
    }                                                                                  \
  } while (0)
  TRY_OR_DESTROY(cudaSetDevice(device));
  TRY_OR_DESTROY(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
  {  // highest priority: its small kernels must slot in between the neighbour passes' CTAs
    int lo = 0, hi = 0;
    TRY_OR_DESTROY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    TRY_OR_DESTROY(cudaStreamCreateWithPriority(&e->side, cudaStreamNonBlocking, hi));
  }
  TRY_OR_DESTROY(cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
  TRY_OR_DESTROY(cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming));
  TRY_OR_DESTROY(cudaStreamCreateWithFlags(&e->copy, cudaStreamNonBlocking));
  TRY_OR_DESTROY(cudaEventCreateWithFlags(&e->ev_staged, cudaEventDisableTiming));
  TRY_OR_DESTROY(cudaEventCreateWithFlags(&e->ev_copied, cudaEventDisableTiming));
  for (int k = 0; k < 2; ++k) {
    TRY_OR_DESTROY(dev_alloc(&e->pos[k], n));
    TRY_OR_DESTROY(dev_alloc(&e->vel[k], n));
    TRY_OR_DESTROY(dev_alloc(&e->vhf[k], n));
    TRY_OR_DESTROY(dev_alloc(&e->sortb.keys[k], n));
    TRY_OR_DESTROY(dev_alloc(&e->sortb.vals[k], n));
  }
  TRY_OR_DESTROY(dev_alloc(&e->rank_cur, n));
  TRY_OR_DESTROY(dev_alloc(&e->rank_prev, n));
  TRY_OR_DESTROY(dev_alloc(&e->ckey, n));
  e->list_cap = neighbor_list_cap(params->h, params->fluid_density, params->particle_mass);
  TRY_OR_DESTROY(dev_alloc(&e->npool, neighbor_pool_words(n, e->list_cap)));
  TRY_OR_DESTROY(dev_alloc(&e->pairs, (size_t)n + 4));
  TRY_OR_DESTROY(dev_alloc(&e->pv, 2 * (size_t)n + 2));
  TRY_OR_DESTROY(dev_alloc(&e->acc, n));
  TRY_OR_DESTROY(dev_alloc(&e->rho, n));
  TRY_OR_DESTROY(dev_alloc(&e->prs, n));
  TRY_OR_DESTROY(dev_alloc(&e->sortb.hist, sort_hist_entries(n)));
  TRY_OR_DESTROY(dev_alloc(&e->sortb.block_sums, sort_block_sum_entries(n)));
  TRY_OR_DESTROY(dev_alloc(&e->red, 1));
  TRY_OR_DESTROY(cudaMallocHost(reinterpret_cast<void**>(&e->red_host), sizeof(Reduce)));
  TRY_OR_DESTROY(dev_alloc(&e->qctl, 1));
  TRY_OR_DESTROY(dev_alloc(&e->qlog, kQueueLog));
  TRY_OR_DESTROY(cudaMallocHost(reinterpret_cast<void**>(&e->qctl_host), sizeof(QueueCtl)));
  TRY_OR_DESTROY(cudaMallocHost(reinterpret_cast<void**>(&e->qlog_host), sizeof(float) * kQueueLog));
  for (int p = 0; p <= SPHB_PASS_COUNT; ++p) {
    TRY_OR_DESTROY(cudaEventCreate(&e->ev[p][0]));
    TRY_OR_DESTROY(cudaEventCreate(&e->ev[p][1]));
  }
  for (int k = 0; k < 4; ++k) TRY_OR_DESTROY(cudaEventCreate(&e->user_ev[k]));
  e->ev_created = true;
  // zero everything that a download-before-step may read
  TRY_OR_DESTROY(cudaMemsetAsync(e->ckey, 0, sizeof(uint32_t) * n, e->stream));
  TRY_OR_DESTROY(cudaMemsetAsync(e->rank_prev, 0, sizeof(uint32_t) * n, e->stream));
  TRY_OR_DESTROY(cudaStreamSynchronize(e->stream));
#undef TRY_OR_DESTROY
  *out = e;
  return 0;
}

namespace {
// Uniform-grid triangle bin (common.cuh: FaceBin), built on the host once per scene: a few
// thousand faces at most.  Bin cell side = the largest of (scene extent / 96) per axis, so the
// grid has at most ~96^3 cells; each face is entered into every bin cell its inflated AABB
// touches, grown by one cell on every side.
int build_face_bin(sphb_engine* e, const sphb_bb* bbs, uint32_t face_count) {
  cudaFree(e->bin_start); cudaFree(e->bin_faces);
  e->bin_start = e->bin_faces = nullptr;
  e->bin = FaceBin();
  if ((int)face_count < e->bin_min_faces || face_count == 0) return 0;
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (uint32_t f = 0; f < face_count; ++f) {
    const float mn[3] = {bbs[f].minx, bbs[f].miny, bbs[f].minz}, mx[3] = {bbs[f].maxx, bbs[f].maxy, bbs[f].maxz};
    for (int a = 0; a < 3; ++a) {
      if (mn[a] < lo[a]) lo[a] = mn[a];
      if (mx[a] > hi[a]) hi[a] = mx[a];
    }
  }
  float side = 0.f;
  for (int a = 0; a < 3; ++a) side = fmaxf(side, (hi[a] - lo[a]) / 96.f);
  if (!(side > 0.f) || !std::isfinite(side)) return 0;
  FaceBin b = FaceBin();
  b.inv = 1.f / side;
  b.ox = lo[0] - side; b.oy = lo[1] - side; b.oz = lo[2] - side;  // one spare cell below
  uint32_t dims[3];
  for (int a = 0; a < 3; ++a) dims[a] = (uint32_t)((hi[a] - lo[a]) * b.inv) + 4u;
  b.nx = dims[0]; b.ny = dims[1]; b.nz = dims[2];
  const size_t ncell = (size_t)b.nx * b.ny * b.nz;
  const float org[3] = {b.ox, b.oy, b.oz};
  auto cell_range = [&](const sphb_bb& bb, int a, uint32_t& c0, uint32_t& c1) {
    const float mn = a == 0 ? bb.minx : a == 1 ? bb.miny : bb.minz;
    const float mx = a == 0 ? bb.maxx : a == 1 ? bb.maxy : bb.maxz;
    const long first = (long)floorf((mn - org[a]) * b.inv) - 1, last = (long)floorf((mx - org[a]) * b.inv) + 1;
    c0 = (uint32_t)(first < 0 ? 0 : first);
    c1 = (uint32_t)(last >= (long)dims[a] ? (long)dims[a] - 1 : last);
  };
  std::vector<uint32_t> start(ncell + 1, 0);
  for (int pass = 0; pass < 2; ++pass) {
    std::vector<uint32_t> cursor;
    std::vector<uint32_t> ids;
    if (pass == 1) {
      uint32_t run = 0;
      for (size_t c = 0; c <= ncell; ++c) {
        const uint32_t cnt = start[c];
        start[c] = run;
        run += cnt;
      }
      cursor.assign(start.begin(), start.end() - 1);
      ids.resize(start[ncell]);
    }
    // descending face order: the first containing face found is the reference's LAST one
    for (uint32_t k = 0; k < face_count; ++k) {
      const uint32_t f = face_count - 1 - k;
      uint32_t x0, x1, y0, y1, z0, z1;
      cell_range(bbs[f], 0, x0, x1); cell_range(bbs[f], 1, y0, y1); cell_range(bbs[f], 2, z0, z1);
      for (uint32_t z = z0; z <= z1; ++z)
        for (uint32_t y = y0; y <= y1; ++y)
          for (uint32_t x = x0; x <= x1; ++x) {
            const size_t c = ((size_t)z * b.ny + y) * b.nx + x;
            if (pass == 0) ++start[c]; else ids[cursor[c]++] = f;
          }
    }
    if (pass == 1) {
      CUDA_TRY(e, dev_alloc(&e->bin_start, ncell + 1));
      CUDA_TRY(e, dev_alloc(&e->bin_faces, ids.size()));
      CUDA_TRY(e, cudaMemcpy(e->bin_start, start.data(), sizeof(uint32_t) * (ncell + 1), cudaMemcpyHostToDevice));
      if (!ids.empty())
        CUDA_TRY(e, cudaMemcpy(e->bin_faces, ids.data(), sizeof(uint32_t) * ids.size(), cudaMemcpyHostToDevice));
    }
  }
  b.start = e->bin_start;
  b.faces = e->bin_faces;
  e->bin = b;
  return 0;
}
}  // namespace

int sphb_set_scene(sphb_engine* e, const sphb_bb* bbs, uint32_t face_count, const float* transforms,
                   const float* rvertices, uint32_t total_grid_points) {
  if (!e) return SPHB_ERR_INVALID;
  if (face_count && (!bbs || !transforms || !rvertices))
    SPHB_FAIL(e, SPHB_ERR_INVALID, "null scene arrays");
  CUDA_TRY(e, cudaSetDevice(e->device));
  CUDA_TRY(e, cudaStreamSynchronize(e->stream));
  cudaFree(e->faces); cudaFree(e->face_offsets); cudaFree(e->df);
  e->faces = nullptr; e->face_offsets = nullptr; e->df = nullptr;
  e->face_count = face_count;
  e->total_grid_points = total_grid_points;

  std::vector<Face> faces(face_count);
  std::vector<uint32_t> offsets(face_count);
  size_t pad = 2;
  for (uint32_t i = 0; i < face_count; ++i) {
    const sphb_bb& b = bbs[i];
    Face& f = faces[i];
    f.maxx = b.maxx; f.maxy = b.maxy; f.maxz = b.maxz;
    f.minx = b.minx; f.miny = b.miny; f.minz = b.minz;
    f.sx = (float)b.size_x; f.sz = (float)b.size_z;
    f.sxm1 = (float)(b.size_x - 1); f.sym1 = (float)(b.size_y - 1); f.szm1 = (float)(b.size_z - 1);
    f.off = (float)b.offset;
    offsets[i] = (uint32_t)b.offset;
    if (i && offsets[i] <= offsets[i - 1])
      SPHB_FAIL(e, SPHB_ERR_INVALID, "BB offsets must be strictly increasing (face %u)", i);
    const size_t reach = b.size_x * b.size_z + b.size_x + 2;
    if (reach > pad) pad = reach;
  }
  // The reference samples df at (x+1, y+1, z+1) of the containing grid cell without a bound
  // check (collisionsv2.cl:279-292): a position on a grid's upper face reads past that
  // grid.  The field is padded with zeros so such reads stay inside the allocation.
  const size_t df_len = (size_t)total_grid_points + pad;
  CUDA_TRY(e, dev_alloc(&e->faces, face_count));
  CUDA_TRY(e, dev_alloc(&e->face_offsets, face_count));
  CUDA_TRY(e, dev_alloc(&e->df, df_len));
  CUDA_TRY(e, cudaMemsetAsync(e->df, 0, sizeof(float) * df_len, e->stream));
  float* d_tr = nullptr;
  float* d_rv = nullptr;
  uint32_t* d_cs = nullptr;
  uint32_t* d_cf = nullptr;
  cudaEvent_t t0 = nullptr, t1 = nullptr;
  struct Temporaries {  // released on every way out, the error returns of CUDA_TRY included
    float*& tr; float*& rv; uint32_t*& cs; uint32_t*& cf; cudaEvent_t& a; cudaEvent_t& b;
    ~Temporaries() {
      cudaFree(tr); cudaFree(rv); cudaFree(cs); cudaFree(cf);
      if (a) cudaEventDestroy(a);
      if (b) cudaEventDestroy(b);
    }
  } temporaries = {d_tr, d_rv, d_cs, d_cf, t0, t1};
  e->df_build_ms = 0.f;
  if (face_count) {
    CUDA_TRY(e, cudaMemcpyAsync(e->faces, faces.data(), sizeof(Face) * face_count,
                                cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(e->face_offsets, offsets.data(), sizeof(uint32_t) * face_count,
                                cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, dev_alloc(&d_tr, (size_t)face_count * 12));
    CUDA_TRY(e, dev_alloc(&d_rv, (size_t)face_count * 4));
    CUDA_TRY(e, cudaMemcpyAsync(d_tr, transforms, sizeof(float) * 12 * face_count,
                                cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(d_rv, rvertices, sizeof(float) * 4 * face_count,
                                cudaMemcpyHostToDevice, e->stream));
    // Which faces can hold a grid point of face o's grid: those whose box meets o's box.  The
    // grid points are min + k * (max - min) / (size - 1) in fp32 (collisionsv2.cl:84-86), which
    // can land a few ulps outside [min, max]; the margin covers that, and a face let in by it
    // merely fails the per-point test the kernel still makes.  Ascending order = the
    // reference's loop order (:70), so ties resolve as they do there.
    // (built on the host by comparing every pair of boxes: 1.6 M pairs for the reference's
    // largest scene; past 32 768 faces — 10^9 pairs — the kernel walks all faces instead)
    if (e->df_face_lists && face_count > 1 && face_count <= 32768u) {
      std::vector<uint32_t> cstart(face_count + 1, 0), cfaces;
      for (uint32_t o = 0; o < face_count; ++o) {
        const sphb_bb& a = bbs[o];
        const float mx = 1e-4f * (fabsf(a.maxx) + fabsf(a.minx)) + 1e-6f;
        const float my = 1e-4f * (fabsf(a.maxy) + fabsf(a.miny)) + 1e-6f;
        const float mz = 1e-4f * (fabsf(a.maxz) + fabsf(a.minz)) + 1e-6f;
        for (uint32_t i = 0; i < face_count; ++i) {
          const sphb_bb& b = bbs[i];
          const bool apart = b.minx > a.maxx + mx || b.maxx < a.minx - mx || b.miny > a.maxy + my ||
                             b.maxy < a.miny - my || b.minz > a.maxz + mz || b.maxz < a.minz - mz;
          if (!apart) cfaces.push_back(i);  // (NaN boxes compare false everywhere: kept)
        }
        cstart[o + 1] = (uint32_t)cfaces.size();
      }
      CUDA_TRY(e, dev_alloc(&d_cs, cstart.size()));
      CUDA_TRY(e, dev_alloc(&d_cf, cfaces.size()));
      CUDA_TRY(e, cudaMemcpyAsync(d_cs, cstart.data(), sizeof(uint32_t) * cstart.size(),
                                  cudaMemcpyHostToDevice, e->stream));
      CUDA_TRY(e, cudaMemcpyAsync(d_cf, cfaces.data(), sizeof(uint32_t) * cfaces.size(),
                                  cudaMemcpyHostToDevice, e->stream));
      // (pageable sources: both copies have been staged when the calls return)
    }
    CUDA_TRY(e, cudaEventCreate(&t0));
    CUDA_TRY(e, cudaEventCreate(&t1));
    CUDA_TRY(e, cudaEventRecord(t0, e->stream));
    e->launches += launch_distance_field(e->df, e->faces, e->face_offsets, d_tr, d_rv, d_cs, d_cf,
                                         face_count, total_grid_points, e->stream);
    CUDA_TRY(e, cudaGetLastError());
    CUDA_TRY(e, cudaEventRecord(t1, e->stream));
    CUDA_TRY(e, cudaEventSynchronize(t1));
    CUDA_TRY(e, cudaEventElapsedTime(&e->df_build_ms, t0, t1));
  }
  CUDA_TRY(e, cudaStreamSynchronize(e->stream));
  if (int rc = build_face_bin(e, bbs, face_count)) return rc;
  e->have_scene = true;
  return 0;
}

int sphb_get_distance_field(sphb_engine* e, float* out) {
  if (!e || !out) return SPHB_ERR_INVALID;
  if (!e->have_scene) SPHB_FAIL(e, SPHB_ERR_NO_SCENE, "no scene");
  CUDA_TRY(e, cudaSetDevice(e->device));
  CUDA_TRY(e, cudaMemcpyAsync(out, e->df, sizeof(float) * e->total_grid_points,
                              cudaMemcpyDeviceToHost, e->stream));
  CUDA_TRY(e, cudaStreamSynchronize(e->stream));
  return 0;
}

// sphb_upload in two halves.  _begin starts the host -> device copy on a stream of its own and
// returns; _commit makes the engine's stream wait for it and converts the stage into the live
// state.  Between the two the engine keeps stepping (and reading back) its current state.
int sphb_upload_begin(sphb_engine* e, const sphb_particle* aos, uint32_t n) {
  if (!e || !aos) return SPHB_ERR_INVALID;
  if (e->slab) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_upload_begin on a slab engine");
  if (n != e->n) SPHB_FAIL(e, SPHB_ERR_INVALID, "upload of %u particles into an engine of %u", n, e->n);
  if (e->upload_pending) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_upload_begin twice without sphb_upload_commit");
  CUDA_TRY(e, cudaSetDevice(e->device));
  if (!e->copy_up) {
    CUDA_TRY(e, cudaStreamCreateWithFlags(&e->copy_up, cudaStreamNonBlocking));
    CUDA_TRY(e, cudaEventCreateWithFlags(&e->ev_up_done, cudaEventDisableTiming));
    CUDA_TRY(e, cudaEventCreateWithFlags(&e->ev_up_consumed, cudaEventDisableTiming));
  }
  if (!e->aos_up) CUDA_TRY(e, dev_alloc(&e->aos_up, n));
  // the stage may still be being converted by the previous commit
  if (e->up_consumed_valid) CUDA_TRY(e, cudaStreamWaitEvent(e->copy_up, e->ev_up_consumed, 0));
  CUDA_TRY(e, cudaMemcpyAsync(e->aos_up, aos, sizeof(sphb_particle) * n, cudaMemcpyHostToDevice, e->copy_up));
  CUDA_TRY(e, cudaEventRecord(e->ev_up_done, e->copy_up));
  e->upload_pending = true;
  return 0;
}

int sphb_upload_commit(sphb_engine* e) {
  if (!e) return SPHB_ERR_INVALID;
  if (!e->upload_pending) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_upload_commit without sphb_upload_begin");
  CUDA_TRY(e, cudaSetDevice(e->device));
  CUDA_TRY(e, cudaStreamWaitEvent(e->stream, e->ev_up_done, 0));
  e->launches += launch_aos_to_soa(e->aos_up, e->n, e->pos[e->cur], e->vel[e->cur], e->vhf[e->cur], e->acc,
                                   e->rho, e->prs, e->ckey, e->rank_cur, e->stream);
  CUDA_TRY(e, cudaGetLastError());
  CUDA_TRY(e, cudaEventRecord(e->ev_up_consumed, e->stream));
  e->up_consumed_valid = true;
  e->upload_pending = false;
  e->bbox_valid = false;
  e->have_state = true;
  return 0;
}

float sphb_scene_build_ms(sphb_engine* e) { return e ? e->df_build_ms : 0.f; }

int sphb_upload(sphb_engine* e, const sphb_particle* aos, uint32_t n) {
  if (!e || !aos) return SPHB_ERR_INVALID;
  if (n != e->n) SPHB_FAIL(e, SPHB_ERR_INVALID, "upload of %u particles into an engine of %u", n, e->n);
  CUDA_TRY(e, cudaSetDevice(e->device));
  if (int rc = sphb_download_wait(e)) return rc;
  if (!e->aos_stage) CUDA_TRY(e, dev_alloc(&e->aos_stage, n));
  CUDA_TRY(e, cudaMemcpyAsync(e->aos_stage, aos, sizeof(sphb_particle) * n, cudaMemcpyHostToDevice,
                              e->stream));
  e->launches += launch_aos_to_soa(e->aos_stage, n, e->pos[e->cur], e->vel[e->cur], e->vhf[e->cur],
                                   e->acc, e->rho, e->prs, e->ckey, e->rank_cur, e->stream);
  CUDA_TRY(e, cudaGetLastError());
  CUDA_TRY(e, cudaStreamSynchronize(e->stream));
  e->bbox_valid = false;
  e->have_state = true;
  return 0;
}


}  // extern "C"

namespace {
// do { advection_collision(dt); computed = computeTimeStep } while (dt - computed > 1e-5)
// (sph_simulation.cpp:246-260), from the post-forces state in the live half: one 32-byte
// read-back per attempt.  `attempts_before`: attempts the device-driven path already made.
int advect_until_settled(sphb_engine* e, float dt_first, int attempts_before, float* computed_out) {
  cudaStream_t s = e->stream;
  int attempts = attempts_before;
  float dt = dt_first, computed = dt_first;
  const int in = e->cur, out = in ^ 1;
  do {
    dt = computed;
    {
      PassTimer t(e, SPHB_PASS_ADVECT);
      int l = launch_reduce_init(e->red, s);
      l += launch_advect(e->pos[in], e->vhf[in], e->acc, e->pos[out], e->vel[out], e->vhf[out], e->n,
                         dt, e->params.restitution, e->faces, e->face_count, e->df, e->red, s, nullptr, &e->bin);
      t.done(l);
    }
    CUDA_TRY(e, cudaGetLastError());
    if (int rc = fetch_reduce(e)) return rc;
    if (e->profile) accumulate_pass(e, SPHB_PASS_ADVECT);
    ++attempts;
    computed = dt_from_maxima(e->params, dec_float(e->red_host->max_v2), dec_float(e->red_host->max_a2));
  } while (dt - computed > 0.00001);
  e->cur = out;
  e->bbox_valid = true;
  e->last_attempts = attempts;
  e->stepped = true;
  *computed_out = computed;
  return 0;
}

// One device-driven step (common.cuh: QueueCtl): the launches of sphb_step with every argument
// fixed — the grid, dt and the retry decision come from e->qctl.  Leaves e->cur as it was (the
// reorder and the advection flip it once each).
int enqueue_queued_step(sphb_engine* e) {
  cudaStream_t s = e->stream;
  const StepConsts c = make_consts(e);  // only c.n shapes the launches; the kernels read g_hot
  SortBuffers& b = e->sortb;
  const int in = e->cur, mid = in ^ 1;
  int l = launch_queue_plan(e->qctl, e->red, s);
  CUDA_TRY(e, (cudaError_t)queue_refresh_exact(e->qctl, s));
  CUDA_TRY(e, (cudaError_t)queue_refresh_neighbors(e->qctl, s));
  l += launch_cell_sort(e->pos[in], c, b.keys[0], b.vals[0], e->ftable, e->scan_scratch, b.keys[1], b.vals[1],
                        b.vals[0], s, true, e->ftable_cap, e->cell_flags);
  const uint32_t* skeys = b.keys[1];
  const uint32_t* perm = b.vals[0];
  l += launch_reorder(perm, e->n, e->pos[in], e->vel[in], e->vhf[in], e->rank_cur, e->pos[mid], e->vel[mid],
                      e->vhf[mid], e->rank_prev, s, true, e->pairs);
  cudaStream_t rs = e->use_side_stream ? e->side : s;
  if (e->use_side_stream) {
    CUDA_TRY(e, cudaEventRecord(e->ev_fork, s));
    CUDA_TRY(e, cudaStreamWaitEvent(e->side, e->ev_fork, 0));
  }
  l += launch_reference_rank(skeys, e->ftable, e->rank_prev, e->ref_table, e->ckey, e->rank_cur, c,
                             e->scan_scratch, rs, true, e->ref_table_cap);
  if (e->use_side_stream) CUDA_TRY(e, cudaEventRecord(e->ev_join, e->side));
  l += launch_density(e->pos[mid], e->vel[mid], e->pairs, e->pv, skeys, e->ftable, e->rho, e->prs, e->npool, e->list_cap, c,
                      e->neighbor_variant, s, true);
  l += launch_forces(e->pos[mid], e->vel[mid], e->pv, skeys, e->ftable, e->rho, e->npool, e->acc, c,
                     e->neighbor_variant, s, nullptr, true);
  if (e->use_side_stream) CUDA_TRY(e, cudaStreamWaitEvent(s, e->ev_join, 0));
  for (int a = 0; a < e->queued_attempts; ++a) {
    l += launch_reduce_init(e->red, s, e->qctl);
    l += launch_advect(e->pos[mid], e->vhf[mid], e->acc, e->pos[in], e->vel[in], e->vhf[in], e->n, 0.f,
                       e->params.restitution, e->faces, e->face_count, e->df, e->red, s, nullptr, &e->bin,
                       e->qctl);
    l += launch_queue_attempt(e->qctl, e->red, e->qlog, a + 1 == e->queued_attempts, s);
  }
  CUDA_TRY(e, cudaGetLastError());
  e->launches += l;
  return 0;
}

// The same step as one CUDA graph launch; re-captured when an argument it froze has changed.
int launch_queued_graph(sphb_engine* e) {
  const uint64_t key[8] = {(uint64_t)(uintptr_t)e->ftable, (uint64_t)(uintptr_t)e->ref_table,
                           (uint64_t)e->ftable_cap, (uint64_t)e->ref_table_cap,
                           (uint64_t)(uintptr_t)e->scan_scratch, (uint64_t)(uintptr_t)e->df,
                           (uint64_t)(uintptr_t)e->bin.start,
                           ((uint64_t)e->face_count << 12) | ((uint64_t)e->neighbor_variant << 8) |
                               ((uint64_t)e->queued_attempts << 4) |
                               ((uint64_t)e->cur << 1) | (e->use_side_stream ? 1u : 0u)};
  if (!e->qgraph || memcmp(key, e->qgraph_key, sizeof key) != 0) {
    if (e->qgraph) cudaGraphExecDestroy(e->qgraph);
    e->qgraph = nullptr;
    cudaGraph_t g = nullptr;
    const uint64_t before = e->launches;
    CUDA_TRY(e, cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
    const int rc = enqueue_queued_step(e);
    const cudaError_t end = cudaStreamEndCapture(e->stream, &g);
    e->qgraph_launches = (uint32_t)(e->launches - before);
    e->launches = before;
    if (rc) {
      if (g) cudaGraphDestroy(g);
      return rc;
    }
    CUDA_TRY(e, end);
    const cudaError_t inst = cudaGraphInstantiate(&e->qgraph, g, 0);
    cudaGraphDestroy(g);
    CUDA_TRY(e, inst);
    memcpy(e->qgraph_key, key, sizeof key);
  }
  CUDA_TRY(e, cudaGraphLaunch(e->qgraph, e->stream));
  e->launches += e->qgraph_launches;
  return 0;
}
}  // namespace

extern "C" {

// The sub-step loop of one frame (libclsph/sph_simulation.cpp:384-409):
//   while (timeleft > 0) { dt = simulate_single_frame(dt); timeleft -= dt; if (timeleft < dt) dt = timeleft; }
// The steps are enqueued in batches without a host round trip between them: the grid of the
// next step, dt and the advection retry are decided on the device (common.cuh: QueueCtl).
int sphb_run_frame(sphb_engine* e, float dt_in, float frame_time, uint64_t max_steps, float* dt_out,
                   uint32_t* steps_out, float* timeleft_log, uint32_t log_cap) {
  if (!e) return SPHB_ERR_INVALID;
  if (!e->have_state) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_run_frame before sphb_upload");
  if (!e->have_scene) SPHB_FAIL(e, SPHB_ERR_NO_SCENE, "sphb_run_frame before sphb_set_scene");
  if (e->slab) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_run_frame on a slab engine (use sphb_slab_step)");
  CUDA_TRY(e, cudaSetDevice(e->device));
  float dt = dt_in, timeleft = frame_time;
  uint64_t steps = 0;
  auto finish_step = [&](float computed) {  // the caller's half of the loop
    dt = computed;
    timeleft -= dt;
    if (timeleft < dt) dt = timeleft;
    if (timeleft_log && steps < log_cap) timeleft_log[steps] = timeleft;
    ++steps;
  };
  while (timeleft > 0.0) {
    if (max_steps && steps >= max_steps) break;
    const bool want_queue = e->use_queue == 1 || (e->use_queue == 2 && e->n <= kQueueAutoMax);
    const bool may_queue = want_queue && e->bbox_valid && e->sort_variant == 0 && (e->neighbor_variant == 0 || e->neighbor_variant == 2) &&
                           !e->profile && e->ftable_cap && e->ref_table_cap && e->scan_scratch;
    const QueueLease lease(e, may_queue);  // (released at the end of this iteration, after the batch's sync)
    const bool can_queue = may_queue && lease.held;
    if (!can_queue) {  // the first step after an upload (its bbox needs a pass of its own), or an option
      float computed = dt;
      if (int rc = sphb_step(e, dt, &computed)) return rc;
      finish_step(computed);
      continue;
    }
    // dt never exceeds simulation_scale / target_fps, so the frame needs at least this many steps
    const float dt_cap = 1.0f / e->params.target_fps * e->params.simulation_scale;
    double want = dt_cap > 0.f ? ceil((double)timeleft / (double)dt_cap) : 1.0;
    if (!(want >= 1.0)) want = 1.0;
    if (want > 64.0) want = 64.0;
    uint32_t batch = (uint32_t)want;
    if (max_steps && (uint64_t)batch > max_steps - steps) batch = (uint32_t)(max_steps - steps);
    QueueCtl& h = *e->qctl_host;
    memset(&h, 0, sizeof h);
    h.hot.c = make_consts(e);
    h.params = e->params;
    h.terms = e->terms;
    h.spiky_eps = e->spiky_eps;
    h.dt = dt;
    h.timeleft = timeleft;
    h.status = SPHB_Q_RUNNING;
    h.budget = batch;
    h.fcap = (uint32_t)(e->ftable_cap > 0xFFFFFFFFull ? 0xFFFFFFFFull : e->ftable_cap);
    h.rcap = (uint32_t)(e->ref_table_cap > 0xFFFFFFFFull ? 0xFFFFFFFFull : e->ref_table_cap);
    h.log_cap = kQueueLog;
    cudaStream_t s = e->stream;
    CUDA_TRY(e, cudaMemcpyAsync(e->qctl, &h, sizeof h, cudaMemcpyHostToDevice, s));
    for (uint32_t k = 0; k < batch; ++k)
      if (int rc = e->use_graph ? launch_queued_graph(e) : enqueue_queued_step(e)) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(&h, e->qctl, sizeof h, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(e, cudaMemcpyAsync(e->qlog_host, e->qlog, sizeof(float) * batch, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(e, cudaMemcpyAsync(e->red_host, e->red, sizeof(Reduce), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(e, cudaStreamSynchronize(s));
    // bookkeeping sphb_step would have done
    e->sorted = 1;
    e->free_keys = e->sortb.keys[0];
    e->free_vals = e->sortb.vals[1];
    e->params = h.params;
    for (uint32_t k = 0; k < h.steps; ++k) {
      if (timeleft_log && steps < log_cap) timeleft_log[steps] = e->qlog_host[k];
      ++steps;
    }
    if (h.steps) {
      e->stepped = true;
      e->last_attempts = (int)h.attempts;
      e->queued_steps += h.steps;
    }
    dt = h.dt;
    timeleft = h.timeleft;
    switch (h.status) {
      case SPHB_Q_GRID:
        SPHB_FAIL(e, SPHB_ERR_GRID, "grid %d x %d x %d: a dimension reached 1024 (the reference asserts, "
                  "sph_simulation.cpp:722-724)", e->params.grid_size_x, e->params.grid_size_y, e->params.grid_size_z);
      case SPHB_Q_GROW: {  // one host-driven step: it grows the tables
        float computed = dt;
        if (int rc = sphb_step(e, dt, &computed)) return rc;
        finish_step(computed);
        break;
      }
      case SPHB_Q_ATTEMPTS: {  // forces are done; the advection goes on from the host
        float computed = dt;
        e->cur ^= 1;  // the reordered half is live, as in sphb_step after the reorder
        if (int rc = advect_until_settled(e, dt, (int)h.attempts, &computed)) return rc;
        finish_step(computed);
        break;
      }
      default:
        break;
    }
  }
  if (dt_out) *dt_out = dt;
  if (steps_out) *steps_out = (uint32_t)steps;
  return 0;
}

uint64_t sphb_queued_steps(sphb_engine* e) { return e ? e->queued_steps : 0; }

int sphb_step(sphb_engine* e, float dt_in, float* dt_out) {
  if (!e) return SPHB_ERR_INVALID;
  if (!e->have_state) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_step before sphb_upload");
  if (!e->have_scene) SPHB_FAIL(e, SPHB_ERR_NO_SCENE, "sphb_step before sphb_set_scene");
  CUDA_TRY(e, cudaSetDevice(e->device));
  cudaStream_t s = e->stream;
  const uint32_t n = e->n;
  bool ran_bbox = false;

  // 1. findMinMaxPosition (sph_simulation.cpp:634-728).  After the first step the bbox of
  //    the live positions was already reduced by the advection kernel that wrote them.
  if (!e->bbox_valid) {
    PassTimer t(e, SPHB_PASS_BBOX);
    int l = launch_reduce_init(e->red, s);
    l += launch_bbox(e->pos[e->cur], n, e->red, s);
    t.done(l);
    CUDA_TRY(e, cudaGetLastError());
    if (int rc = fetch_reduce(e)) return rc;
    ran_bbox = true;
  }
  {
    float mn[3], mx[3];
    for (int a = 0; a < 3; ++a) {
      mn[a] = dec_float(e->red_host->min_enc[a]);
      mx[a] = dec_float(e->red_host->max_enc[a]);
    }
    if (int rc = grid_from_bbox(e, mn, mx)) return rc;
  }
  const uint32_t ncell = e->params.grid_cell_count;
  const StepConsts c = make_consts(e);
  if (int rc = check_search_grid(e, c)) return rc;
  if (int rc = ensure_tables(e, (size_t)c.nfine + 1, (size_t)ncell + 1)) return rc;

  // 2. locate_in_grid (kernels/grid.cl:40-64) + 3. sort_particles (sph_simulation.cpp:110-198):
  //    keys only, no struct copies; then ONE gather of the carried state
  const uint32_t* skeys = nullptr;
  const uint32_t* perm = nullptr;
  if (int rc = order_by_cell(e, c, &skeys, &perm)) return rc;
  {
    PassTimer t(e, SPHB_PASS_REORDER);
    const int in = e->cur, out = in ^ 1;
    t.done(launch_reorder(perm, n, e->pos[in], e->vel[in], e->vhf[in], e->rank_cur,
                          e->pos[out], e->vel[out], e->vhf[out], e->rank_prev, s, false, e->pairs));
    e->cur = out;
  }
  // the reference's cell table (kernels/sort.cl:61-78) and every particle's place in the
  // reference's sorted array: nothing in the step depends on it, so it runs on the side
  // stream while the (compute-bound) neighbour passes run, and is joined before the step ends
  // (option "side_stream" = 0 keeps it on the main stream: per-pass times are then each
  // kernel's own)
  cudaStream_t rs = e->use_side_stream ? e->side : s;
  if (e->use_side_stream) {
    CUDA_TRY(e, cudaEventRecord(e->ev_fork, s));
    CUDA_TRY(e, cudaStreamWaitEvent(e->side, e->ev_fork, 0));
  }
  {
    PassTimer t(e, SPHB_PASS_REFRANK, rs);
    t.done(launch_reference_rank(skeys, e->ftable, e->rank_prev, e->ref_table, e->ckey,
                                 e->rank_cur, c, e->scan_scratch, rs));
  }
  if (e->use_side_stream) CUDA_TRY(e, cudaEventRecord(e->ev_join, e->side));
  // 4. density_pressure (kernels/sphb.cl:3-35), 5. forces (:37-126)
  {
    PassTimer t(e, SPHB_PASS_DENSITY);
    t.done(launch_density(e->pos[e->cur], e->vel[e->cur], e->pairs, e->pv, skeys, e->ftable, e->rho,
                          e->prs, e->npool, e->list_cap, c, e->neighbor_variant, s));
  }
  {
    PassTimer t(e, SPHB_PASS_FORCES);
    t.done(launch_forces(e->pos[e->cur], e->vel[e->cur], e->pv, skeys, e->ftable, e->rho,
                         e->npool, e->acc, c, e->neighbor_variant, s));
  }
  if (e->use_side_stream) CUDA_TRY(e, cudaStreamWaitEvent(s, e->ev_join, 0));
  CUDA_TRY(e, cudaGetLastError());

  // 6. do { advection_collision(dt); computed = computeTimeStep } while (dt - computed > 1e-5)
  float computed = dt_in;
  if (int rc = advect_until_settled(e, dt_in, 0, &computed)) return rc;
  if (e->profile) {
    if (ran_bbox) accumulate_pass(e, SPHB_PASS_BBOX);
    for (int p = SPHB_PASS_LOCATE; p < SPHB_PASS_COUNT; ++p)
      if (p != SPHB_PASS_ADVECT) accumulate_pass(e, p);
    e->prof_steps++;
  }
  if (dt_out) *dt_out = computed;
  return 0;
}

int sphb_neighbor_list_stats(sphb_engine* e, uint64_t out[4]) {
  if (!e || !out) return SPHB_ERR_INVALID;
  if (!e->stepped) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_neighbor_list_stats before the first step");
  if (e->neighbor_variant != 0) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_neighbor_list_stats: option neighbor_kernel != 0");
  CUDA_TRY(e, cudaSetDevice(e->device));
  // (the four counters are staged in the free half of the sort buffers, as sphb_download_debug does)
  unsigned long long* dev = reinterpret_cast<unsigned long long*>(e->free_keys);
  if (!dev || e->cap < 8) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_neighbor_list_stats: no scratch");
  e->launches += launch_list_stats(e->npool, e->n, dev, e->stream);
  unsigned long long host[4];
  CUDA_TRY(e, cudaMemcpyAsync(host, dev, sizeof host, cudaMemcpyDeviceToHost, e->stream));
  CUDA_TRY(e, cudaStreamSynchronize(e->stream));
  for (int k = 0; k < 4; ++k) out[k] = host[k];
  return 0;
}

int sphb_get_parameters(sphb_engine* e, sphb_params* out) {
  if (!e || !out) return SPHB_ERR_INVALID;
  *out = e->params;
  return 0;
}

int sphb_download_begin(sphb_engine* e, sphb_particle* aos) {
  if (!e || !aos) return SPHB_ERR_INVALID;
  if (!e->have_state) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_download before sphb_upload");
  CUDA_TRY(e, cudaSetDevice(e->device));
  if (int rc = sphb_download_wait(e)) return rc;  // the stage is reused
  if (!e->aos_stage) CUDA_TRY(e, dev_alloc(&e->aos_stage, e->n));
  e->launches += launch_soa_to_aos(e->aos_stage, e->n, e->pos[e->cur], e->vel[e->cur], e->vhf[e->cur],
                                   e->acc, e->rho, e->prs, e->ckey, e->rank_cur, e->stream);
  CUDA_TRY(e, cudaGetLastError());
  // the snapshot is complete once the stage is written; the copy then runs beside later steps
  CUDA_TRY(e, cudaEventRecord(e->ev_staged, e->stream));
  CUDA_TRY(e, cudaStreamWaitEvent(e->copy, e->ev_staged, 0));
  CUDA_TRY(e, cudaMemcpyAsync(aos, e->aos_stage, sizeof(sphb_particle) * e->n, cudaMemcpyDeviceToHost,
                              e->copy));
  CUDA_TRY(e, cudaEventRecord(e->ev_copied, e->copy));
  e->download_pending = true;
  return 0;
}

// May be called from another host thread than the one that steps the engine (it touches only
// the event), and any number of times.
int sphb_download_wait(sphb_engine* e) {
  if (!e) return SPHB_ERR_INVALID;
  if (!e->download_pending) return 0;
  CUDA_TRY(e, cudaSetDevice(e->device));
  CUDA_TRY(e, cudaEventSynchronize(e->ev_copied));
  return 0;
}

int sphb_download(sphb_engine* e, sphb_particle* aos) {
  if (int rc = sphb_download_begin(e, aos)) return rc;
  if (int rc = sphb_download_wait(e)) return rc;
  e->download_pending = false;
  return 0;
}

int sphb_download_debug(sphb_engine* e, uint32_t* keys, uint32_t* perm, uint32_t* cell_table,
                        uint32_t cell_table_cap, int* attempts) {
  if (!e) return SPHB_ERR_INVALID;
  if (!e->stepped) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_download_debug before the first sphb_step");
  CUDA_TRY(e, cudaSetDevice(e->device));
  // the sort buffers are free between steps: use the half that does not hold the fine keys
  uint32_t* const scratch_k = e->free_keys;
  uint32_t* const scratch_v = e->free_vals;
  e->launches += launch_reference_debug(e->rank_cur, e->rank_prev, e->ckey, e->n, scratch_k, scratch_v,
                                        e->stream);
  CUDA_TRY(e, cudaGetLastError());
  if (keys)
    CUDA_TRY(e, cudaMemcpyAsync(keys, scratch_k, sizeof(uint32_t) * e->n,
                                cudaMemcpyDeviceToHost, e->stream));
  if (perm)
    CUDA_TRY(e, cudaMemcpyAsync(perm, scratch_v, sizeof(uint32_t) * e->n,
                                cudaMemcpyDeviceToHost, e->stream));
  uint32_t cnt = 0;
  uint32_t first_key = 0;
  if (cell_table) {
    cnt = e->params.grid_cell_count < cell_table_cap ? e->params.grid_cell_count : cell_table_cap;
    if (cnt)
      CUDA_TRY(e, cudaMemcpyAsync(cell_table, e->ref_table, sizeof(uint32_t) * cnt,
                                  cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(&first_key, scratch_k, sizeof(uint32_t),
                                cudaMemcpyDeviceToHost, e->stream));
  }
  CUDA_TRY(e, cudaStreamSynchronize(e->stream));
  // kernels/sort.cl:61-78 quirk: the table is pre-filled with N, and entries c <= key[0] are
  // zeroed by work-item c — which exists only when c < N.
  for (uint32_t cidx = e->n; cell_table && cidx <= first_key && cidx < cnt; ++cidx) cell_table[cidx] = e->n;
  if (attempts) *attempts = e->last_attempts;
  return 0;
}

int sphb_debug_advect(sphb_engine* e, float dt) {
  if (!e) return SPHB_ERR_INVALID;
  if (!e->have_state) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_debug_advect before sphb_upload");
  if (!e->have_scene) SPHB_FAIL(e, SPHB_ERR_NO_SCENE, "sphb_debug_advect before sphb_set_scene");
  CUDA_TRY(e, cudaSetDevice(e->device));
  const int in = e->cur, out = in ^ 1;
  e->launches += launch_reduce_init(e->red, e->stream);
  e->launches += launch_advect(e->pos[in], e->vhf[in], e->acc, e->pos[out], e->vel[out], e->vhf[out],
                               e->n, dt, e->params.restitution, e->faces, e->face_count, e->df, e->red,
                               e->stream, nullptr, &e->bin);
  CUDA_TRY(e, cudaGetLastError());
  if (int rc = fetch_reduce(e)) return rc;
  e->cur = out;
  e->bbox_valid = true;
  return 0;
}

int sphb_profile_enable(sphb_engine* e, int on) {
  if (!e) return SPHB_ERR_INVALID;
  e->profile = on != 0;
  return 0;
}

int sphb_profile_reset(sphb_engine* e) {
  if (!e) return SPHB_ERR_INVALID;
  for (int p = 0; p < SPHB_PASS_COUNT; ++p) {
    e->pass_ms[p] = 0.f;
    e->pass_launches[p] = 0;
  }
  e->prof_steps = 0;
  return 0;
}

int sphb_get_pass_times(sphb_engine* e, float ms[SPHB_PASS_COUNT], uint32_t launches[SPHB_PASS_COUNT],
                        uint32_t* steps) {
  if (!e) return SPHB_ERR_INVALID;
  CUDA_TRY(e, cudaSetDevice(e->device));
  CUDA_TRY(e, cudaStreamSynchronize(e->stream));
  for (int p = 0; p < SPHB_PASS_COUNT; ++p) {
    if (ms) ms[p] = e->pass_ms[p];
    if (launches) launches[p] = e->pass_launches[p];
  }
  if (steps) *steps = e->prof_steps;
  return 0;
}

int sphb_host_alloc(void** out, size_t bytes) {
  if (!out) return SPHB_ERR_INVALID;
  *out = nullptr;
  g_create_error.clear();
  cudaError_t err = cudaMallocHost(out, bytes ? bytes : 1);
  if (err != cudaSuccess) {
    cudaGetLastError();
    return fail(nullptr, (int)err, __FILE__, __LINE__, "cudaMallocHost(%zu) -> %s", bytes,
                cudaGetErrorString(err));
  }
  return 0;
}

int sphb_host_free(void* p) {
  if (!p) return 0;
  return (int)cudaFreeHost(p);
}

int sphb_event_record(sphb_engine* e, int slot) {
  if (!e || slot < 0 || slot >= 4) return SPHB_ERR_INVALID;
  CUDA_TRY(e, cudaSetDevice(e->device));
  CUDA_TRY(e, cudaEventRecord(e->user_ev[slot], e->stream));
  return 0;
}

int sphb_event_elapsed_ms(sphb_engine* e, int from_slot, int to_slot, float* ms) {
  if (!e || !ms || from_slot < 0 || from_slot >= 4 || to_slot < 0 || to_slot >= 4) return SPHB_ERR_INVALID;
  CUDA_TRY(e, cudaSetDevice(e->device));
  CUDA_TRY(e, cudaEventSynchronize(e->user_ev[to_slot]));
  CUDA_TRY(e, cudaEventElapsedTime(ms, e->user_ev[from_slot], e->user_ev[to_slot]));
  return 0;
}

uint64_t sphb_launch_count(sphb_engine* e) { return e ? e->launches : 0; }

void* sphb_stream(sphb_engine* e) { return e ? (void*)e->stream : nullptr; }

int sphb_set_option(sphb_engine* e, const char* name, int value) {
  if (!e || !name) return SPHB_ERR_INVALID;
  if (!strcmp(name, "triangle_bin")) {  // takes effect at the next sphb_set_scene
    e->bin_min_faces = value ? 17 : 0x7fffffff;
    return 0;
  }
  if (!strcmp(name, "df_face_lists")) {  // takes effect at the next sphb_set_scene
    e->df_face_lists = value != 0;
    return 0;
  }
  if (!strcmp(name, "sort")) {
    e->sort_variant = value ? 1 : 0;
    return 0;
  }
  if (!strcmp(name, "scan")) {  // the cell tables' prefix scans: 1 = one pass, 0 = three kernels
    CUDA_TRY(e, cudaSetDevice(e->device));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    if (e->qgraph) {  // a captured step holds the other variant's launches
      cudaGraphExecDestroy(e->qgraph);
      e->qgraph = nullptr;
    }
    set_scan_variant(value);  // (process-wide: every engine of this process)
    return 0;
  }
  if (!strcmp(name, "neighbor_kernel")) {
    e->neighbor_variant = value;
    return 0;
  }
  if (!strcmp(name, "device_loop")) {  // sphb_run_frame: 0 = one host round trip per step
    e->use_queue = value < 0 || value > 2 ? 2 : value;
    return 0;
  }
  if (!strcmp(name, "queued_attempts")) {  // advection attempts enqueued per device-driven step
    e->queued_attempts = value < 1 ? 1 : (value > 4 ? 4 : value);
    return 0;
  }
  if (!strcmp(name, "table_headroom")) {  // 0: cell tables are allocated without spare entries
    e->table_headroom = value != 0;
    return 0;
  }
  if (!strcmp(name, "graph")) {  // sphb_run_frame: a device-driven step is one CUDA graph launch
    e->use_graph = value != 0;
    return 0;
  }
  if (!strcmp(name, "side_stream")) {
    e->use_side_stream = value != 0;
    return 0;
  }
  if (!strcmp(name, "p2p_halo")) {  // before sphb_slab_comm_init: 0 = always exchange through NCCL
    e->p2p_want = value != 0;
    return 0;
  }
  if (!strcmp(name, "halo_margin")) {  // sphb_slab_step: spare records per halo message
    e->halo_margin = value < 0 ? 0u : (uint32_t)value;
    return 0;
  }
  SPHB_FAIL(e, SPHB_ERR_INVALID, "unknown option '%s'", name);
}

}  // extern "C"
