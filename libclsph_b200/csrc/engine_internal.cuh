// Internal definitions shared by the engine's host translation units (engine.cu, slab.cu):
// the engine object and its small host helpers.  Not part of the public boundary.
#pragma once
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "common.cuh"

using namespace sphb;

struct sphb_engine {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t side = nullptr;  // the reference-order bookkeeping runs beside the neighbour passes
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool use_side_stream = true;  // option "side_stream"
  // asynchronous read-back (sphb_download_begin / _wait): device -> host copy on its own stream
  cudaStream_t copy = nullptr;
  cudaEvent_t ev_staged = nullptr, ev_copied = nullptr;
  bool download_pending = false;
  // asynchronous upload (sphb_upload_begin / _commit): host -> device copy on its own stream into
  // its own stage, so that it runs beside a step and beside a read-back (PCIe is full duplex)
  cudaStream_t copy_up = nullptr;
  cudaEvent_t ev_up_done = nullptr, ev_up_consumed = nullptr;
  sphb_particle* aos_up = nullptr;
  bool upload_pending = false, up_consumed_valid = false;
  int slab_attempt = 0;  // advection attempts of the current slab step
  // native multi-GPU driver (slab_nccl.cu): communicator, the four halo messages
  // (send left, recv left, send right, recv right), the 8-float reduction cell, the global bbox
  void* comm = nullptr;
  int comm_rank = 0, comm_world = 1;
  void* msg[4] = {nullptr, nullptr, nullptr, nullptr};
  float* red8_dev = nullptr;
  float* red8_host = nullptr;
  float gmin[3] = {0.f, 0.f, 0.f}, gmax[3] = {0.f, 0.f, 0.f};
  uint64_t bytes_sent = 0;
  // records moved per message last step ([0] to / [1] from the left, [2] to / [3] from the right):
  // the next step's messages carry that many + a margin instead of the full capacity
  uint32_t last_records[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
  uint64_t resends = 0;
  uint32_t halo_margin = 4096;  // option "halo_margin"
  // exchange over peer memory (option "p2p_halo"): the neighbours' inboxes as seen from this GPU
  // ([0] the left neighbour's inbox from its right, [1] the right neighbour's inbox from its left)
  int p2p_want = 1;
  bool p2p_active = false;
  void* p2p_remote[2] = {nullptr, nullptr};
  bool p2p_ipc[2] = {false, false};  // opened with cudaIpcOpenMemHandle (to be closed)
  uint32_t p2p_seq = 0;
  double phase_ms[3] = {0.0, 0.0, 0.0};  // pack..unpack | sort..first reduction | retries + commit
  // GPU-side segments of a slab step (sphb_slab_gpu_segments; recorded while profiling is on):
  // pack | wait for the halos | count read-back (host) + unpack | sort..forces + owned scan |
  // first advection | reduction + all-reduce | later attempts | (idle between steps)
  cudaEvent_t seg_ev[9] = {};
  bool seg_created = false;
  double seg_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  uint64_t seg_steps = 0;
  uint32_t n = 0;    // particles the kernels run over (slab mode: owned + ghosts of this step)
  uint32_t cap = 0;  // allocated particle slots (== n outside slab mode)
  // slab mode (multi-GPU, slab.cu): this engine owns the particles with x in [xlo, xhi)
  bool slab = false;
  float xlo = 0.f, xhi = 0.f;
  int has_left = 0, has_right = 0;
  uint32_t exch_cap = 0;           // records per halo / migration message
  uint32_t n_owned = 0;            // host copy (valid after unpack / upload)
  uint32_t n_upper = 0;            // host upper bound of *n_owned_dev between syncs
  uint32_t* n_owned_dev = nullptr; // device count of owned particles (written by compaction)
  uint32_t* counters = nullptr;    // 4 pack counters (device)
  uint32_t* src_of = nullptr;      // compaction map of the last commit
  uint32_t* host_counts = nullptr; // pinned, 16 words
  bool committed = false;          // src_of / acc / rho / prs describe the live owned particles
  int slab_in = 0;                 // ping-pong half holding the pre-advect state of this step
  sphb_params params;   // the reference's `parameters`, mutated every step
  sphb_terms terms;
  float spiky_eps = 0.f;

  // sorted SoA state; pos/vel/vhf are ping-pong pairs, cur[] says which half is live
  float4* pos[2] = {nullptr, nullptr};
  float4* vel[2] = {nullptr, nullptr};
  float4* vhf[2] = {nullptr, nullptr};
  int cur = 0;
  float4* acc = nullptr;
  float* rho = nullptr;
  float* prs = nullptr;
  SortBuffers sortb = {};
  int sorted = 0;            // which half of sortb holds the last step's sorted keys
  int sort_variant = 0;      // 0 = counting sort by search cell, 1 = LSD radix sort
  uint32_t* free_keys = nullptr;  // sort buffers that are free between steps (download scratch)
  uint32_t* free_vals = nullptr;
  // The compute order is the engine's own (fine row-major cells); rank_cur[i] is the position
  // of particle i in the REFERENCE's order, rank_prev[i] its position one step earlier
  // (= the permutation the reference's sort applied), ckey[i] its Morton key (grid_index).
  uint32_t* rank_cur = nullptr;
  uint32_t* rank_prev = nullptr;
  uint32_t* ckey = nullptr;
  uint32_t* npool = nullptr;  // neighbour lists recorded by the density pass (kernels_neighbors.cu)
  uint32_t list_cap = 64;     // entries a pair's list has room for
  float4* pairs = nullptr;    // positions as one record per pair of particles (reorder -> density)
  float4* pv = nullptr;       // position + velocity + pair factors per particle (density -> forces)
  uint32_t* ftable = nullptr;     // fine cell table, nfine + 1 entries
  uint8_t* cell_flags = nullptr;  // one byte per fine cell, inside ftable's allocation (behind ftable_cap entries)
  size_t ftable_cap = 0;
  uint32_t* ref_table = nullptr;  // the reference's cell table, grid_cell_count (+1) entries
  size_t ref_table_cap = 0;
  uint32_t* scan_scratch = nullptr;
  size_t scan_scratch_cap = 0;
  sphb_particle* aos_stage = nullptr;  // device staging for upload / download

  // scene
  Face* faces = nullptr;
  uint32_t* face_offsets = nullptr;
  float* df = nullptr;
  uint32_t face_count = 0, total_grid_points = 0;
  bool have_scene = false;
  FaceBin bin = {};               // uniform-grid triangle bin (start == nullptr: not built)
  uint32_t* bin_start = nullptr;  // device storage behind `bin`
  uint32_t* bin_faces = nullptr;
  int bin_min_faces = 17;         // scenes with fewer faces scan them all (option "triangle_bin")
  bool df_face_lists = true;      // distance-field build walks per-face candidate lists (option "df_face_lists")
  float df_build_ms = 0.f;        // device time of the last distance-field build

  // reductions
  Reduce* red = nullptr;       // device
  Reduce* red_host = nullptr;  // pinned
  bool bbox_valid = false;     // red_host holds the bbox of the live positions
  bool have_state = false;
  int last_attempts = 0;
  bool stepped = false;

  // device-driven stepping (sphb_run_frame; common.cuh: QueueCtl)
  QueueCtl* qctl = nullptr;       // device
  QueueCtl* qctl_host = nullptr;  // pinned
  float* qlog = nullptr;          // device: timeleft after every step of a batch
  float* qlog_host = nullptr;     // pinned
  int use_queue = 2;              // option "device_loop": 0 off, 1 on, 2 = on below kQueueAutoMax particles
  bool use_graph = true;          // option "graph": a queued step is one CUDA graph launch
  int queued_attempts = 2;        // option "queued_attempts": advection attempts enqueued per step
  bool table_headroom = true;     // option "table_headroom": cell tables keep 1.5x spare entries
  cudaGraphExec_t qgraph = nullptr;
  uint32_t qgraph_launches = 0;   // kernels inside the captured step
  uint64_t qgraph_key[8] = {};    // what the captured step's launch arguments depend on
  uint64_t queued_steps = 0;      // steps that ran without a host round trip

  // options / measurement
  int neighbor_variant = 0;
  bool profile = false;
  uint32_t recorded = 0;  // bit p: pass p's event pair was recorded and not yet accumulated
  cudaEvent_t ev[SPHB_PASS_COUNT + 1][2];
  cudaEvent_t user_ev[4];
  bool ev_created = false;
  float pass_ms[SPHB_PASS_COUNT] = {};
  uint32_t pass_launches[SPHB_PASS_COUNT] = {};
  uint32_t prof_steps = 0;
  uint64_t launches = 0;

  std::string err;
};

std::string& sphb_create_error_slot();
void sphb_slab_comm_release(sphb_engine* e);  // slab_nccl.cu
int slab_read_counts(sphb_engine* e, const void* recv_left_dev, const void* recv_right_dev);  // slab.cu
int slab_pack_to(sphb_engine* e, void* msg_left_dev, void* msg_right_dev, uint32_t p2p_seq);  // slab.cu
int slab_apply_unpack(sphb_engine* e, const void* recv_left_dev, const void* recv_right_dev,
                      uint32_t* n_owned_out, uint32_t* n_ghost_out);
extern "C" int sphb_create_with_capacity(const sphb_params* params, const sphb_terms* terms, int device,
                                         uint32_t capacity, sphb_engine** out);

namespace {

// text of the last failure that happened without an engine (one slot per thread, engine.cu)
#define g_create_error (::sphb_create_error_slot())

int fail(sphb_engine* e, int code, const char* file, int line, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  char full[640];
  snprintf(full, sizeof full, "sphb error %d (%s:%d): %s", code, file, line, buf);
  if (e) e->err = full; else g_create_error = full;
  return code;
}

#define CUDA_TRY(e, call)                                                                   \
  do {                                                                                      \
    cudaError_t err__ = (call);                                                             \
    if (err__ != cudaSuccess)                                                               \
      return fail((e), (int)err__, __FILE__, __LINE__, "%s -> %s", #call,                   \
                  cudaGetErrorString(err__));                                               \
  } while (0)

#define SPHB_FAIL(e, code, ...) return fail((e), (code), __FILE__, __LINE__, __VA_ARGS__)

template <typename T>
cudaError_t dev_alloc(T** p, size_t count) {
  return cudaMalloc(reinterpret_cast<void**>(p), sizeof(T) * (count ? count : 1));
}

// libclsph/sph_simulation.cpp:668-727 (arithmetic in common.cuh: grid_from_bbox_hd)
int grid_from_bbox(sphb_engine* e, const float mn[3], const float mx[3]) {
  sphb_params& P = e->params;
  if (!grid_from_bbox_hd(P, mn, mx))
    SPHB_FAIL(e, SPHB_ERR_GRID, "grid %d x %d x %d: a dimension reached 1024 (the reference asserts, "
              "sph_simulation.cpp:722-724)", P.grid_size_x, P.grid_size_y, P.grid_size_z);
  return 0;
}

StepConsts make_consts(const sphb_engine* e) { return make_consts_hd(e->params, e->terms, e->spiky_eps, e->n); }

struct PassTimer {
  sphb_engine* e;
  int pass;
  cudaStream_t st;
  PassTimer(sphb_engine* e_, int pass_, cudaStream_t st_ = nullptr)
      : e(e_), pass(pass_), st(st_ ? st_ : e_->stream) {
    if (e->profile) cudaEventRecord(e->ev[pass][0], st);
  }
  void done(int launches) {
    e->launches += launches;
    if (e->profile) {
      cudaEventRecord(e->ev[pass][1], st);
      e->recorded |= 1u << pass;
      e->pass_launches[pass] += launches;
    }
  }
};

}  // namespace

namespace {

// Reads the device reduction cell back (32 bytes) and waits for it: the one host
// synchronisation of a step.
int fetch_reduce(sphb_engine* e) {
  CUDA_TRY(e, cudaMemcpyAsync(e->red_host, e->red, sizeof(Reduce), cudaMemcpyDeviceToHost, e->stream));
  CUDA_TRY(e, cudaStreamSynchronize(e->stream));
  return 0;
}

void accumulate_pass(sphb_engine* e, int pass) {
  if (!(e->recorded & (1u << pass))) return;
  e->recorded &= ~(1u << pass);
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, e->ev[pass][0], e->ev[pass][1]) == cudaSuccess) e->pass_ms[pass] += ms;
}

// The search grid's cell count is kept in 32 bits (keys, table indices): fnx * fny * fnz is
// computed in 64 bits here and refused when it does not fit, BEFORE any table is sized from the
// wrapped product (the reference allows up to 1023 cells per axis; 32 x 512^3 search cells already
// exceed 2^32).
int check_search_grid(sphb_engine* e, const StepConsts& c) {
  const unsigned long long cells = (unsigned long long)c.fnx * c.fny * c.fnz;
  if (cells + 1ull > 0xFFFFFFFFull || cells != (unsigned long long)c.nfine)
    SPHB_FAIL(e, SPHB_ERR_GRID, "search grid %u x %u x %u = %llu cells does not fit 32-bit cell keys (grid %d x %d x %d)",
              c.fnx, c.fny, c.fnz, cells, e->params.grid_size_x, e->params.grid_size_y, e->params.grid_size_z);
  return 0;
}

int ensure_tables(sphb_engine* e, size_t fine_entries, size_t ref_entries) {
  // grow rarely: capacities are kept with 1.5x headroom
  const bool grow_f = fine_entries > e->ftable_cap, grow_r = ref_entries > e->ref_table_cap;
  if (!grow_f && !grow_r && e->scan_scratch) return 0;
  CUDA_TRY(e, cudaStreamSynchronize(e->stream));
  if (grow_f) {
    cudaFree(e->ftable);
    e->ftable = nullptr;
    e->ftable_cap = e->table_headroom ? fine_entries + fine_entries / 2 + 1024 : fine_entries;
    e->ftable_cap = (e->ftable_cap + 3) & ~(size_t)3;  // the flags behind the table start 16-byte aligned
    // (+ one byte per cell behind the table, padded to 16: the counting sort's "this cell needs ordering" flags)
    CUDA_TRY(e, dev_alloc(&e->ftable, e->ftable_cap + e->ftable_cap / 4 + 8));
    e->cell_flags = reinterpret_cast<uint8_t*>(e->ftable + e->ftable_cap);
  }
  if (grow_r) {
    cudaFree(e->ref_table);
    e->ref_table = nullptr;
    e->ref_table_cap = e->table_headroom ? ref_entries + ref_entries / 2 + 1024 : ref_entries;
    CUDA_TRY(e, dev_alloc(&e->ref_table, e->ref_table_cap));
  }
  size_t longest = e->ftable_cap > e->ref_table_cap ? e->ftable_cap : e->ref_table_cap;
  if (e->cap > longest) longest = e->cap;  // slab mode scans one flag per particle
  const size_t need = scan_scratch_entries(longest);
  if (need > e->scan_scratch_cap) {
    cudaFree(e->scan_scratch);
    e->scan_scratch = nullptr;
    e->scan_scratch_cap = need;
    CUDA_TRY(e, dev_alloc(&e->scan_scratch, need));
  }
  return 0;
}

int key_bits_for(uint32_t grid_cell_count);

// locate + sort + search-cell table, by either variant.  On return: *skeys = sorted search keys,
// *perm = previous index of every sorted particle, e->ftable built.
int order_by_cell(sphb_engine* e, const StepConsts& c, const uint32_t** skeys, const uint32_t** perm) {
  cudaStream_t s = e->stream;
  SortBuffers& b = e->sortb;
  if (e->sort_variant == 0) {
    PassTimer t(e, SPHB_PASS_SORT);
    t.done(launch_cell_sort(e->pos[e->cur], c, b.keys[0], b.vals[0], e->ftable, e->scan_scratch, b.keys[1],
                            b.vals[1], b.vals[0], s, false, 0, e->cell_flags));
    e->sorted = 1;
    *skeys = b.keys[1];
    *perm = b.vals[0];
    e->free_keys = b.keys[0];
    e->free_vals = b.vals[1];
    return 0;
  }
  {
    PassTimer t(e, SPHB_PASS_LOCATE);
    t.done(launch_locate(e->pos[e->cur], b.keys[0], c, s));
  }
  int r = 0;
  {
    PassTimer t(e, SPHB_PASS_SORT);
    t.done(launch_sort_pairs(b, c.n, key_bits_for(c.nfine), &r, s));
  }
  e->sorted = r;
  {
    PassTimer t(e, SPHB_PASS_CELLTABLE);
    t.done(launch_fine_table(b.keys[r], c.n, e->ftable, c.nfine, e->scan_scratch, s));
  }
  *skeys = b.keys[r];
  *perm = b.vals[r];
  e->free_keys = b.keys[r ^ 1];
  e->free_vals = b.vals[r ^ 1];
  return 0;
}

int key_bits_for(uint32_t grid_cell_count) {
  int bits = 0;
  while (bits < 32 && (grid_cell_count >> bits) != 0) ++bits;
  return bits < 1 ? 1 : bits;
}

}  // namespace
