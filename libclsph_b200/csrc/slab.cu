// sphb slab API: one engine = one x-slab of the domain on one GPU (include/sphb.h, "slab"
// section).  The caller (one process per GPU) moves the messages between neighbouring ranks
// and all-reduces 8 floats per advection attempt; everything else of the step is here.
//
// Step protocol, per rank (the same on every rank):
//   sphb_slab_pack      classify owned particles -> message to the left / right neighbour
//   [caller: send / recv the two messages]
//   sphb_slab_unpack    append received migrants (owned) and ghosts; host learns the counts
//   sphb_slab_forces    global bbox -> grid; locate, sort, reorder, cell table, density over
//                       owned + ghosts, forces over owned
//   repeat { sphb_slab_advect(dt) ; [caller: all-reduce MAX of the 8 floats] ; dt' = sphb_compute_dt }
//     until dt - dt' <= 1e-5                     (sph_simulation.cpp:246-260)
//   sphb_slab_commit    drop the ghosts: the advected owned particles become the live state;
//                       the all-reduced floats hold the next step's global bbox
#include "engine_internal.cuh"

namespace {
const float kHaloCells = 2.f;  // ghost layer = 2 h: ghosts within h of the slab get a correct density

int require_slab(sphb_engine* e) {
  if (!e) return SPHB_ERR_INVALID;
  if (!e->slab) SPHB_FAIL(e, SPHB_ERR_INVALID, "not a slab engine (use sphb_slab_create)");
  return 0;
}
}  // namespace

extern "C" {

int sphb_slab_create(const sphb_params* params, const sphb_terms* terms, int device, uint32_t capacity,
                     uint32_t exchange_capacity, sphb_engine** out) {
  if (!params || !out) return SPHB_ERR_INVALID;
  sphb_params p = *params;
  if (p.particles_count == 0) p.particles_count = 1;
  if (capacity == 0) capacity = p.particles_count;
  // particles_count stays the GLOBAL count in the engine's copy of `parameters`
  int rc = sphb_create_with_capacity(&p, terms, device, capacity, out);
  if (rc != 0) return rc;
  sphb_engine* e = *out;
  e->params = *params;
  e->slab = true;
  e->n = 0;
  e->exch_cap = exchange_capacity;
  cudaError_t err = cudaSuccess;
  if (err == cudaSuccess) err = dev_alloc(&e->n_owned_dev, 1);
  if (err == cudaSuccess) err = dev_alloc(&e->counters, 8);  // 4 pack counters + [4] exchange error
  if (err == cudaSuccess) err = cudaMemsetAsync(e->counters, 0, 8 * sizeof(uint32_t), e->stream);
  if (err == cudaSuccess) err = dev_alloc(&e->src_of, capacity);
  if (err == cudaSuccess) err = dev_alloc(&e->aos_stage, capacity);
  if (err == cudaSuccess) err = cudaMallocHost(reinterpret_cast<void**>(&e->host_counts), 16 * sizeof(uint32_t));
  if (err == cudaSuccess) err = cudaMemsetAsync(e->n_owned_dev, 0, sizeof(uint32_t), e->stream);
  if (err == cudaSuccess) err = cudaStreamSynchronize(e->stream);
  if (err != cudaSuccess) {
    rc = fail(nullptr, (int)err, __FILE__, __LINE__, "slab allocation -> %s", cudaGetErrorString(err));
    sphb_destroy(e);
    *out = nullptr;
    return rc;
  }
  return 0;
}

size_t sphb_slab_message_bytes(sphb_engine* e) { return e ? slab_message_bytes(e->exch_cap) : 0; }

int sphb_slab_configure(sphb_engine* e, float xlo, float xhi, int has_left, int has_right) {
  if (int rc = require_slab(e)) return rc;
  if (!(xlo < xhi)) SPHB_FAIL(e, SPHB_ERR_INVALID, "slab bounds [%g, %g)", xlo, xhi);
  // A slab with a neighbour on BOTH sides must be wider than the ghost layer (2h) plus what a
  // particle can move in one step (< h: the time step keeps v dt below h): a rank exchanges only
  // with its two neighbours and sends only particles it OWNS, so across a thinner slab the
  // particles its neighbours need — and a migrant's new owner — would lie two ranks away and
  // ghost densities would be silently wrong.
  const float min_width = (kHaloCells + 1.f) * e->params.h;
  if (has_left && has_right && !(xhi - xlo >= min_width))
    SPHB_FAIL(e, SPHB_ERR_INVALID, "slab [%g, %g) is %g wide: an inner slab must be at least %g (3 h) — use fewer "
              "GPUs for this fluid, or re-cut", xlo, xhi, xhi - xlo, min_width);
  e->xlo = xlo; e->xhi = xhi;
  e->has_left = has_left ? 1 : 0;
  e->has_right = has_right ? 1 : 0;
  return 0;
}

int sphb_slab_upload(sphb_engine* e, const sphb_particle* aos, const uint32_t* ids, uint32_t n_owned) {
  if (int rc = require_slab(e)) return rc;
  if (n_owned > e->cap) SPHB_FAIL(e, SPHB_ERR_CAPACITY, "%u particles into a slab of capacity %u", n_owned, e->cap);
  if (n_owned && (!aos || !ids)) return SPHB_ERR_INVALID;
  CUDA_TRY(e, cudaSetDevice(e->device));
  cudaStream_t s = e->stream;
  if (n_owned) {
    CUDA_TRY(e, cudaMemcpyAsync(e->aos_stage, aos, sizeof(sphb_particle) * n_owned, cudaMemcpyHostToDevice, s));
    e->launches += launch_aos_to_soa(e->aos_stage, n_owned, e->pos[e->cur], e->vel[e->cur], e->vhf[e->cur],
                                     e->acc, e->rho, e->prs, e->ckey, e->rank_cur, s);
    CUDA_TRY(e, cudaMemcpyAsync(e->rank_cur, ids, sizeof(uint32_t) * n_owned, cudaMemcpyHostToDevice, s));
  }
  e->host_counts[0] = n_owned;
  CUDA_TRY(e, cudaMemcpyAsync(e->n_owned_dev, e->host_counts, sizeof(uint32_t), cudaMemcpyHostToDevice, s));
  CUDA_TRY(e, cudaGetLastError());
  CUDA_TRY(e, cudaStreamSynchronize(s));
  e->n = e->n_owned = e->n_upper = n_owned;
  e->have_state = true;
  e->committed = false;
  return 0;
}

/* out8_dev (device, 8 floats): (-min xyz, max xyz, -inf, -inf) of the owned positions */
int sphb_slab_bbox(sphb_engine* e, float* out8_dev) {
  if (int rc = require_slab(e)) return rc;
  if (!out8_dev) return SPHB_ERR_INVALID;
  CUDA_TRY(e, cudaSetDevice(e->device));
  e->launches += launch_reduce_init(e->red, e->stream);
  e->launches += launch_slab_bbox(e->pos[e->cur], e->n_owned_dev, e->n_upper, e->red, e->stream);
  e->launches += launch_slab_reduce_floats(e->red, out8_dev, e->stream);
  CUDA_TRY(e, cudaGetLastError());
  return 0;
}

int sphb_slab_pack(sphb_engine* e, void* msg_left_dev, void* msg_right_dev) {
  return slab_pack_to(e, msg_left_dev, msg_right_dev, 0);
}

}  // extern "C"

int slab_pack_to(sphb_engine* e, void* msg_left_dev, void* msg_right_dev, uint32_t p2p_seq) {
  if (int rc = require_slab(e)) return rc;
  if (!e->have_state) SPHB_FAIL(e, SPHB_ERR_INVALID, "sphb_slab_pack before sphb_slab_upload");
  if ((e->has_left && !msg_left_dev) || (e->has_right && !msg_right_dev)) return SPHB_ERR_INVALID;
  CUDA_TRY(e, cudaSetDevice(e->device));
  const int cur = e->cur;
  e->launches += launch_slab_pack(e->pos[cur], e->vel[cur], e->vhf[cur], e->rank_cur, e->n_owned_dev,
                                  e->n_upper, e->xlo, e->xhi, kHaloCells * e->params.h, e->has_left,
                                  e->has_right, e->has_left ? msg_left_dev : nullptr,
                                  e->has_right ? msg_right_dev : nullptr, e->exch_cap, e->counters, e->stream,
                                  p2p_seq);
  CUDA_TRY(e, cudaGetLastError());
  return 0;
}

// The step's first host synchronisation: the headers of the two received messages, this
// rank's own pack counters and its particle count, into e->host_counts
// ([0..3] recv left header, [4..7] recv right header, [8..11] pack counters, [12] owned before).
int slab_read_counts(sphb_engine* e, const void* recv_left_dev, const void* recv_right_dev) {
  if (int rc = require_slab(e)) return rc;
  if ((e->has_left && !recv_left_dev) || (e->has_right && !recv_right_dev)) return SPHB_ERR_INVALID;
  CUDA_TRY(e, cudaSetDevice(e->device));
  cudaStream_t s = e->stream;
  uint32_t* hc = e->host_counts;
  memset(hc, 0, 16 * sizeof(uint32_t));
  if (e->has_left) CUDA_TRY(e, cudaMemcpyAsync(hc + 0, recv_left_dev, 16, cudaMemcpyDeviceToHost, s));
  if (e->has_right) CUDA_TRY(e, cudaMemcpyAsync(hc + 4, recv_right_dev, 16, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(e, cudaMemcpyAsync(hc + 8, e->counters, 16, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(e, cudaMemcpyAsync(hc + 12, e->n_owned_dev, 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(e, cudaMemcpyAsync(hc + 13, e->counters + 4, 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(e, cudaStreamSynchronize(s));
  return 0;
}

// Appends the received records (counts already in e->host_counts).
int slab_apply_unpack(sphb_engine* e, const void* recv_left_dev, const void* recv_right_dev,
                      uint32_t* n_owned_out, uint32_t* n_ghost_out) {
  cudaStream_t s = e->stream;
  const uint32_t* hc = e->host_counts;
  const uint32_t in_l = hc[0], mig_in_l = hc[1], in_r = hc[4], mig_in_r = hc[5];
  const uint32_t out_l = hc[8], mig_out_l = hc[9], out_r = hc[10], mig_out_r = hc[11];
  const uint32_t before = hc[12];
  if (out_l > e->exch_cap || out_r > e->exch_cap || in_l > e->exch_cap || in_r > e->exch_cap)
    SPHB_FAIL(e, SPHB_ERR_CAPACITY, "halo message of %u / %u (sent) %u / %u (received) records exceeds the "
              "exchange capacity %u", out_l, out_r, in_l, in_r, e->exch_cap);
  if ((uint64_t)before + in_l + in_r > e->cap)
    SPHB_FAIL(e, SPHB_ERR_CAPACITY, "%u local + %u + %u received particles exceed the slab capacity %u",
              before, in_l, in_r, e->cap);
  const int cur = e->cur;
  if (e->has_left)
    e->launches += launch_slab_unpack(recv_left_dev, in_l, before, e->pos[cur], e->vel[cur], e->vhf[cur],
                                      e->rank_cur, s);
  if (e->has_right)
    e->launches += launch_slab_unpack(recv_right_dev, in_r, before + in_l, e->pos[cur], e->vel[cur],
                                      e->vhf[cur], e->rank_cur, s);
  CUDA_TRY(e, cudaGetLastError());
  e->n = before + in_l + in_r;
  e->n_owned = before - mig_out_l - mig_out_r + mig_in_l + mig_in_r;
  e->committed = false;
  if (n_owned_out) *n_owned_out = e->n_owned;
  if (n_ghost_out) *n_ghost_out = e->n - e->n_owned;
  return 0;
}

extern "C" {

int sphb_slab_unpack(sphb_engine* e, const void* recv_left_dev, const void* recv_right_dev,
                     uint32_t* n_owned_out, uint32_t* n_ghost_out) {
  if (int rc = slab_read_counts(e, recv_left_dev, recv_right_dev)) return rc;
  return slab_apply_unpack(e, recv_left_dev, recv_right_dev, n_owned_out, n_ghost_out);
}

int sphb_slab_forces(sphb_engine* e, const float gmin[3], const float gmax[3]) {
  if (int rc = require_slab(e)) return rc;
  if (!e->have_scene) SPHB_FAIL(e, SPHB_ERR_NO_SCENE, "sphb_slab_forces before sphb_set_scene");
  if (!gmin || !gmax) return SPHB_ERR_INVALID;
  CUDA_TRY(e, cudaSetDevice(e->device));
  cudaStream_t s = e->stream;
  // the reference's grid from the GLOBAL bbox: identical `parameters` on every rank
  if (int rc = grid_from_bbox(e, gmin, gmax)) return rc;
  const sphb_params& P = e->params;
  StepConsts c = make_consts(e);
  const uint32_t n = e->n;
  c.n = n;
  // the search grid covers only this slab (+ ghost layer + the usual 2-cell padding) in x
  const float halo = kHaloCells * P.h;
  float lo = P.min_point.s[0], hi = P.max_point.s[0];
  if (e->has_left) lo = fmaxf(lo, e->xlo - halo - 2.f * c.cell);
  if (e->has_right) hi = fminf(hi, e->xhi + halo + 2.f * c.cell);
  if (!(hi > lo)) hi = lo + c.cell;
  c.minx = lo;
  c.fnx = kXPerRef * ((uint32_t)((hi - lo) / c.cell) + 1u);
  c.nfine = c.fnx * c.fny * c.fnz;
  if (int rc = check_search_grid(e, c)) return rc;
  if (int rc = ensure_tables(e, (size_t)c.nfine + 1, 1)) return rc;
  if (n == 0) {
    e->slab_in = e->cur;
    return 0;
  }
  const uint32_t* skeys = nullptr;
  const uint32_t* perm = nullptr;
  if (int rc = order_by_cell(e, c, &skeys, &perm)) return rc;
  {
    PassTimer t(e, SPHB_PASS_REORDER);
    const int in = e->cur, out = in ^ 1;
    int l = launch_reorder(perm, n, e->pos[in], e->vel[in], e->vhf[in], e->rank_cur, e->pos[out],
                           e->vel[out], e->vhf[out], e->rank_prev, s, false, e->pairs);
    std::swap(e->rank_cur, e->rank_prev);  // ids (+ ghost bit) now in compute order
    e->cur = out;
    t.done(l);
  }
  {
    PassTimer t(e, SPHB_PASS_DENSITY);
    t.done(launch_density(e->pos[e->cur], e->vel[e->cur], e->pairs, e->pv, skeys, e->ftable, e->rho, e->prs,
                          e->npool, e->list_cap, c, 0, s));
  }
  {
    PassTimer t(e, SPHB_PASS_FORCES);
    t.done(launch_forces(e->pos[e->cur], e->vel[e->cur], e->pv, skeys, e->ftable, e->rho, e->npool,
                         e->acc, c, 0, s, e->rank_cur));
  }
  // every owned particle's dense slot in the advected arrays (ghosts are dropped there)
  e->launches += launch_slab_owned_scan(e->rank_cur, n, e->free_keys, e->scan_scratch, e->n_owned_dev, s);
  CUDA_TRY(e, cudaGetLastError());
  e->slab_in = e->cur;
  e->slab_attempt = 0;
  e->stepped = true;
  if (e->profile) {
    CUDA_TRY(e, cudaStreamSynchronize(s));
    for (int p = SPHB_PASS_LOCATE; p <= SPHB_PASS_FORCES; ++p) accumulate_pass(e, p);
    e->prof_steps++;
  }
  return 0;
}

int sphb_slab_advect(sphb_engine* e, float dt, float* out8_dev) {
  if (int rc = require_slab(e)) return rc;
  if (!out8_dev) return SPHB_ERR_INVALID;
  CUDA_TRY(e, cudaSetDevice(e->device));
  cudaStream_t s = e->stream;
  const int in = e->slab_in, out = in ^ 1;
  {
    PassTimer t(e, SPHB_PASS_ADVECT);
    int l = launch_reduce_init(e->red, s);
    if (e->n) {
      SlabAdvect sa;
      sa.ids = e->rank_cur;
      sa.dst = e->free_keys;
      sa.ids_out = e->rank_prev;
      sa.src_of = e->src_of;
      sa.write_aux = e->slab_attempt == 0;
      ++e->slab_attempt;
      l += launch_advect(e->pos[in], e->vhf[in], e->acc, e->pos[out], e->vel[out], e->vhf[out], e->n, dt,
                         e->params.restitution, e->faces, e->face_count, e->df, e->red, s, &sa, &e->bin);
    }
    l += launch_slab_reduce_floats(e->red, out8_dev, s);
    t.done(l);
  }
  CUDA_TRY(e, cudaGetLastError());
  return 0;
}

int sphb_slab_commit(sphb_engine* e) {
  if (int rc = require_slab(e)) return rc;
  CUDA_TRY(e, cudaSetDevice(e->device));
  cudaStream_t s = e->stream;
  const int in = e->slab_in, out = in ^ 1;
  // the advection pass already wrote the owned particles densely into the `out` half (ids into
  // rank_prev): committing is a pointer swap
  (void)in;
  std::swap(e->rank_cur, e->rank_prev);
  e->cur = out;
  if (e->profile) {  // the accepted advection attempt
    CUDA_TRY(e, cudaStreamSynchronize(s));
    accumulate_pass(e, SPHB_PASS_ADVECT);
  }
  e->n_upper = e->n;  // the exact count is on the device; the host reads it at the next unpack
  e->committed = true;
  return 0;
}

int sphb_slab_download(sphb_engine* e, sphb_particle* aos, uint32_t* ids, uint32_t capacity,
                       uint32_t* n_owned_out) {
  if (int rc = require_slab(e)) return rc;
  if (!n_owned_out) return SPHB_ERR_INVALID;
  CUDA_TRY(e, cudaSetDevice(e->device));
  cudaStream_t s = e->stream;
  CUDA_TRY(e, cudaMemcpyAsync(e->host_counts + 12, e->n_owned_dev, 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(e, cudaStreamSynchronize(s));
  const uint32_t n = e->host_counts[12];
  *n_owned_out = n;
  if (!aos || !ids) return 0;
  if (n > capacity) SPHB_FAIL(e, SPHB_ERR_CAPACITY, "download of %u particles into %u slots", n, capacity);
  if (n == 0) return 0;
  uint32_t* ids_dev = e->free_vals ? e->free_vals : e->sortb.vals[1];  // free between steps
  const int cur = e->cur;
  e->launches += launch_slab_to_aos(e->aos_stage, ids_dev, e->n_owned_dev, n, e->pos[cur], e->vel[cur],
                                    e->vhf[cur], e->rank_cur, e->committed ? e->src_of : nullptr, e->acc,
                                    e->rho, e->prs, e->pos[cur ^ 1], e->ckey, e->params.min_point.s[0],
                                    e->params.min_point.s[1], e->params.min_point.s[2], e->params.h * 2, s);
  CUDA_TRY(e, cudaGetLastError());
  CUDA_TRY(e, cudaMemcpyAsync(aos, e->aos_stage, sizeof(sphb_particle) * n, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(e, cudaMemcpyAsync(ids, ids_dev, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(e, cudaStreamSynchronize(s));
  return 0;
}

float sphb_compute_dt(const sphb_params* params, float max_v2, float max_a2) {
  return params ? dt_from_maxima(*params, max_v2, max_a2) : 0.f;
}

}  // extern "C"
