// The slab step driven natively: one C call per step and rank, with the halo + migration
// exchange (ncclSend / ncclRecv with the two x-neighbours) and the 8-float all-reduce issued
// on the engine's own stream.  The same protocol as libclsph_b200/slab.py::SlabDriver.step —
// that Python driver remains the reference implementation of the protocol (and is what the
// gloo tests run) — minus the interpreter between the kernels: at small slabs the step is
// bound by launch and host overhead, not by the GPU.
//
// NCCL is resolved at run time (dlopen of the libnccl already in the process, else by name),
// so that libsphb.so itself carries no link-time dependency on it and single-GPU users need
// no NCCL at all.  Only plain-C entry points with ABI-stable enum values are used.
#include <dlfcn.h>
#include <unistd.h>

#include <chrono>

#include "engine_internal.cuh"

namespace {

// the subset of nccl.h this file needs (values are part of NCCL's stable ABI)
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { kNcclSuccess = 0, kNcclUint8 = 1, kNcclFloat32 = 7, kNcclMax = 2 };

struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(ncclUniqueId*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool ok = false;
};

NcclApi load_nccl() {
  NcclApi api;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    api.handle = dlopen(nm, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);  // the copy already loaded (torch's)
    if (api.handle) break;
  }
  for (const char* nm : names) {
    if (api.handle) break;
    api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
  }
  if (!api.handle) return api;
#define SPHB_NCCL_SYM(field, name) \
  *reinterpret_cast<void**>(&api.field) = dlsym(api.handle, name); \
  if (!api.field) return api;
  SPHB_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
  SPHB_NCCL_SYM(CommInitRank, "ncclCommInitRank")
  SPHB_NCCL_SYM(CommDestroy, "ncclCommDestroy")
  SPHB_NCCL_SYM(GroupStart, "ncclGroupStart")
  SPHB_NCCL_SYM(GroupEnd, "ncclGroupEnd")
  SPHB_NCCL_SYM(Send, "ncclSend")
  SPHB_NCCL_SYM(Recv, "ncclRecv")
  SPHB_NCCL_SYM(AllReduce, "ncclAllReduce")
  SPHB_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef SPHB_NCCL_SYM
  api.ok = true;
  return api;
}

// resolved once, whichever thread asks first (several slab engines of one process are driven by
// one host thread each: libclsph_b200/host/sph_simulation_slabs.cpp)
NcclApi& nccl() {
  static NcclApi api = load_nccl();
  return api;
}

int nccl_fail(sphb_engine* e, int rc, const char* what, const char* file, int line) {
  NcclApi& a = nccl();
  return fail(e, SPHB_ERR_NCCL_BASE + rc, file, line, "%s -> NCCL error %d (%s)", what, rc,
              a.GetErrorString ? a.GetErrorString(rc) : "?");
}
#define NCCL_TRY(e, call)                                                        \
  do {                                                                           \
    const int nccl_rc_ = (call);                                                 \
    if (nccl_rc_ != kNcclSuccess) return nccl_fail((e), nccl_rc_, #call, __FILE__, __LINE__); \
  } while (0)

int require_comm(sphb_engine* e) {
  if (!e) return SPHB_ERR_INVALID;
  if (!e->slab) SPHB_FAIL(e, SPHB_ERR_INVALID, "not a slab engine (use sphb_slab_create)");
  if (!e->comm) SPHB_FAIL(e, SPHB_ERR_INVALID, "no communicator (call sphb_slab_comm_init)");
  return 0;
}

// all-reduce MAX of the 8 floats at e->red8_dev, then bring them to the host
int reduce8(sphb_engine* e, cudaEvent_t after_allreduce = nullptr) {
  NcclApi& a = nccl();
  if (e->comm_world > 1)
    NCCL_TRY(e, a.AllReduce(e->red8_dev, e->red8_dev, 8, kNcclFloat32, kNcclMax,
                            static_cast<ncclComm_t>(e->comm), e->stream));
  if (after_allreduce) cudaEventRecord(after_allreduce, e->stream);
  CUDA_TRY(e, cudaMemcpyAsync(e->red8_host, e->red8_dev, 8 * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
  CUDA_TRY(e, cudaStreamSynchronize(e->stream));
  return 0;
}

void bbox_from_red8(sphb_engine* e) {
  for (int a = 0; a < 3; ++a) {
    e->gmin[a] = -e->red8_host[a];
    e->gmax[a] = e->red8_host[3 + a];
  }
}

}  // namespace

namespace {
void p2p_release(sphb_engine* e) {
  for (int k = 0; k < 2; ++k) {
    if (e->p2p_remote[k] && e->p2p_ipc[k]) cudaIpcCloseMemHandle(e->p2p_remote[k]);
    e->p2p_remote[k] = nullptr;
    e->p2p_ipc[k] = false;
  }
  e->p2p_active = false;
}

// What a rank tells a neighbour about the inbox that neighbour is to write into.
struct InboxCard {
  uint64_t pid;
  int32_t device;
  uint32_t valid;
  uint64_t ptr;
  cudaIpcMemHandle_t handle;
};
static_assert(sizeof(InboxCard) <= 128, "an inbox card travels in 128 bytes");

// Maps the neighbours' inboxes into this GPU's address space and agrees with every other rank on
// whether the exchange runs over peer memory: all of them, or none.  Collective.
int p2p_setup(sphb_engine* e) {
  NcclApi& a = nccl();
  ncclComm_t comm = static_cast<ncclComm_t>(e->comm);
  p2p_release(e);
  const int rank = e->comm_rank, world = e->comm_world;
  const bool left = rank > 0, right = rank + 1 < world;
  bool ok = e->p2p_want != 0 && world > 1;
  // cards of my two inboxes: msg[1] (from the left) goes to the left neighbour, msg[3] to the right
  InboxCard mine[2];
  memset(mine, 0, sizeof mine);
  for (int k = 0; k < 2; ++k) {
    void* box = e->msg[k == 0 ? 1 : 3];
    mine[k].pid = (uint64_t)getpid();
    mine[k].device = e->device;
    mine[k].ptr = (uint64_t)(uintptr_t)box;
    mine[k].valid = 0;
    if (ok && cudaIpcGetMemHandle(&mine[k].handle, box) == cudaSuccess) mine[k].valid = 1;
    else cudaGetLastError();
  }
  // staged through the (still unused) send / receive messages: 128 bytes each way
  unsigned char stage[2][128];
  memset(stage, 0, sizeof stage);
  memcpy(stage[0], &mine[0], sizeof(InboxCard));
  memcpy(stage[1], &mine[1], sizeof(InboxCard));
  CUDA_TRY(e, cudaMemcpyAsync(e->msg[0], stage[0], 128, cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(e, cudaMemcpyAsync(e->msg[2], stage[1], 128, cudaMemcpyHostToDevice, e->stream));
  NCCL_TRY(e, a.GroupStart());
  if (left) {
    NCCL_TRY(e, a.Send(e->msg[0], 128, kNcclUint8, rank - 1, comm, e->stream));
    NCCL_TRY(e, a.Recv(e->msg[1], 128, kNcclUint8, rank - 1, comm, e->stream));
  }
  if (right) {
    NCCL_TRY(e, a.Send(e->msg[2], 128, kNcclUint8, rank + 1, comm, e->stream));
    NCCL_TRY(e, a.Recv(e->msg[3], 128, kNcclUint8, rank + 1, comm, e->stream));
  }
  NCCL_TRY(e, a.GroupEnd());
  InboxCard theirs[2];  // [0] the left neighbour's inbox from its right, [1] the right one's from its left
  memset(theirs, 0, sizeof theirs);
  if (left) CUDA_TRY(e, cudaMemcpyAsync(&theirs[0], e->msg[1], sizeof(InboxCard), cudaMemcpyDeviceToHost, e->stream));
  if (right) CUDA_TRY(e, cudaMemcpyAsync(&theirs[1], e->msg[3], sizeof(InboxCard), cudaMemcpyDeviceToHost, e->stream));
  CUDA_TRY(e, cudaStreamSynchronize(e->stream));
  for (int k = 0; k < 2 && ok; ++k) {
    if (!(k == 0 ? left : right)) continue;
    const InboxCard& c = theirs[k];
    if (c.pid == (uint64_t)getpid()) {  // another engine of this process: plain peer access
      int can = 0;
      if (c.device == e->device) can = 1;
      else if (cudaDeviceCanAccessPeer(&can, e->device, c.device) != cudaSuccess) can = 0;
      if (can && c.device != e->device) {
        const cudaError_t err = cudaDeviceEnablePeerAccess(c.device, 0);
        if (err != cudaSuccess && err != cudaErrorPeerAccessAlreadyEnabled) can = 0;
      }
      cudaGetLastError();
      if (can) e->p2p_remote[k] = (void*)(uintptr_t)c.ptr;
      else ok = false;
    } else if (c.valid) {
      void* p = nullptr;
      if (cudaIpcOpenMemHandle(&p, c.handle, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess && p) {
        e->p2p_remote[k] = p;
        e->p2p_ipc[k] = true;
      } else {
        cudaGetLastError();
        ok = false;
      }
    } else {
      ok = false;
    }
  }
  // the inboxes start without a sequence number; then: all ranks, or none
  for (int k = 1; k < 4; k += 2) CUDA_TRY(e, cudaMemsetAsync(e->msg[k], 0, 128, e->stream));
  float flag[8] = {ok ? 0.f : 1.f, 0, 0, 0, 0, 0, 0, 0};
  CUDA_TRY(e, cudaMemcpyAsync(e->red8_dev, flag, sizeof flag, cudaMemcpyHostToDevice, e->stream));
  if (int rc = reduce8(e)) return rc;
  if (e->red8_host[0] != 0.f || !ok) {
    p2p_release(e);
    return 0;
  }
  e->p2p_active = true;
  e->p2p_seq = 0;
  return 0;
}
}  // namespace

void sphb_slab_comm_release(sphb_engine* e) {
  if (!e) return;
  p2p_release(e);
  if (e->comm) {
    NcclApi& a = nccl();
    if (a.ok) a.CommDestroy(static_cast<ncclComm_t>(e->comm));
    e->comm = nullptr;
  }
  for (int k = 0; k < 4; ++k) {
    if (e->msg[k]) cudaFree(e->msg[k]);
    e->msg[k] = nullptr;
  }
  if (e->red8_dev) cudaFree(e->red8_dev);
  if (e->red8_host) cudaFreeHost(e->red8_host);
  e->red8_dev = nullptr;
  e->red8_host = nullptr;
}

extern "C" {

int sphb_nccl_unique_id(unsigned char id_out[128]) {
  if (!id_out) return SPHB_ERR_INVALID;
  NcclApi& a = nccl();
  if (!a.ok) return fail(nullptr, SPHB_ERR_NCCL_MISSING, __FILE__, __LINE__, "libnccl.so.2 could not be loaded");
  ncclUniqueId id;
  const int rc = a.GetUniqueId(&id);
  if (rc != kNcclSuccess) return nccl_fail(nullptr, rc, "ncclGetUniqueId", __FILE__, __LINE__);
  memcpy(id_out, id.internal, 128);
  return 0;
}

int sphb_slab_comm_init(sphb_engine* e, const unsigned char id_in[128], int rank, int world) {
  if (!e || !id_in || world < 1 || rank < 0 || rank >= world) return SPHB_ERR_INVALID;
  if (!e->slab) SPHB_FAIL(e, SPHB_ERR_INVALID, "not a slab engine (use sphb_slab_create)");
  NcclApi& a = nccl();
  if (!a.ok) SPHB_FAIL(e, SPHB_ERR_NCCL_MISSING, "libnccl.so.2 could not be loaded");
  CUDA_TRY(e, cudaSetDevice(e->device));
  sphb_slab_comm_release(e);
  ncclUniqueId id;
  memcpy(id.internal, id_in, 128);
  ncclComm_t comm = nullptr;
  NCCL_TRY(e, a.CommInitRank(&comm, world, id, rank));
  e->comm = comm;
  e->comm_rank = rank;
  e->comm_world = world;
  const size_t bytes = slab_message_bytes(e->exch_cap);
  for (int k = 0; k < 4; ++k) {
    CUDA_TRY(e, cudaMalloc(&e->msg[k], bytes));
    CUDA_TRY(e, cudaMemsetAsync(e->msg[k], 0, 16, e->stream));
  }
  CUDA_TRY(e, cudaMalloc(reinterpret_cast<void**>(&e->red8_dev), 8 * sizeof(float)));
  CUDA_TRY(e, cudaMallocHost(reinterpret_cast<void**>(&e->red8_host), 8 * sizeof(float)));
  CUDA_TRY(e, cudaStreamSynchronize(e->stream));
  return p2p_setup(e);
}

// After sphb_slab_configure + sphb_slab_upload on every rank: the global bbox of the initial
// state (collective).
int sphb_slab_start(sphb_engine* e) {
  if (int rc = require_comm(e)) return rc;
  if (int rc = sphb_slab_bbox(e, e->red8_dev)) return rc;
  if (int rc = reduce8(e)) return rc;
  bbox_from_red8(e);
  e->bytes_sent = 0;
  e->resends = 0;
  for (int k = 0; k < 4; ++k) e->last_records[k] = 0xFFFFFFFFu;
  e->phase_ms[0] = e->phase_ms[1] = e->phase_ms[2] = 0.0;
  return 0;
}

// One simulate_single_frame over all slabs (collective; the same dt_in on every rank gives the
// same dt_out on every rank).  sph_simulation.cpp:200-263 with the exchange in front.
int sphb_slab_step(sphb_engine* e, float dt_in, float* dt_out, int* attempts_out, uint32_t* n_owned_out,
                   uint32_t* n_ghost_out) {
  if (int rc = require_comm(e)) return rc;
  NcclApi& a = nccl();
  ncclComm_t comm = static_cast<ncclComm_t>(e->comm);
  typedef std::chrono::steady_clock clk;
  const clk::time_point t0 = clk::now();
  clk::time_point t1 = t0, t2 = t0;
  void* send_l = e->msg[0]; void* recv_l = e->msg[1]; void* send_r = e->msg[2]; void* recv_r = e->msg[3];
  uint32_t n_owned = 0, n_ghost = 0;
  // GPU-side segment marks (only while profiling: an event record is cheap but not free)
  const bool seg = e->profile;
  if (seg && !e->seg_created) {
    for (int k = 0; k < 9; ++k) CUDA_TRY(e, cudaEventCreate(&e->seg_ev[k]));
    e->seg_created = true;
  }
  bool seg_prev = seg && e->seg_steps > 0;  // seg_ev[8]: the end of the previous profiled step
  auto mark = [&](int k) { if (seg) cudaEventRecord(e->seg_ev[k], e->stream); };
  mark(0);
  if (e->p2p_active) {
    // The exchange over peer memory: the pack kernel writes its records into the neighbours'
    // inboxes over NVLink and publishes this step's number behind them; this stream waits for
    // the same number in its own inboxes.  No send, no receive, no message sizing.
    const uint32_t seq = ++e->p2p_seq;
    if (int rc = slab_pack_to(e, e->p2p_remote[0], e->p2p_remote[1], seq)) return rc;
    mark(1);
    e->launches += launch_slab_wait(e->has_left ? recv_l : nullptr, e->has_right ? recv_r : nullptr, seq,
                                    e->counters + 4, 20ull * 1000ull * 1000ull * 1000ull, e->stream);
    CUDA_TRY(e, cudaGetLastError());
    mark(2);
    if (int rc = slab_read_counts(e, recv_l, recv_r)) return rc;  // host sync
    const uint32_t* hc = e->host_counts;
    if (hc[13]) SPHB_FAIL(e, SPHB_ERR_INVALID, "the %s neighbour's halo did not arrive within 20 s",
                          hc[13] == 1 ? "left" : "right");
    e->bytes_sent += (e->has_left ? slab_message_bytes(hc[8] > e->exch_cap ? e->exch_cap : hc[8]) : 0) +
                     (e->has_right ? slab_message_bytes(hc[10] > e->exch_cap ? e->exch_cap : hc[10]) : 0);
    if (int rc = slab_apply_unpack(e, recv_l, recv_r, &n_owned, &n_ghost)) return rc;
  } else {
  if (int rc = sphb_slab_pack(e, send_l, send_r)) return rc;
  mark(1);
  // A message carries the records its predecessor carried one step ago plus a margin, not the
  // full capacity: both ends of a pair know that count (the sender counted it, the receiver
  // read it from the header), so both post the same size.  A message that outgrew its size is
  // detected on both ends too (count in the header > records carried) and the pair repeats
  // that one exchange at full capacity.
  uint32_t rec[4];  // to the left, from the left, to the right, from the right
  for (int k = 0; k < 4; ++k) {
    const uint32_t last = e->last_records[k];
    uint64_t want = last == 0xFFFFFFFFu ? e->exch_cap : (uint64_t)last + (e->halo_margin == 0 ? 0u : (last / 8u > e->halo_margin ? last / 8u : e->halo_margin));
    rec[k] = want > e->exch_cap ? e->exch_cap : (uint32_t)want;
  }
  auto exchange = [&](bool left, bool right, const uint32_t r[4]) -> int {
    if (!left && !right) return 0;
    // a fixed order (left pair, then right pair) on every rank keeps the pairs matched
    NCCL_TRY(e, a.GroupStart());
    if (left) {
      NCCL_TRY(e, a.Send(send_l, slab_message_bytes(r[0]), kNcclUint8, e->comm_rank - 1, comm, e->stream));
      NCCL_TRY(e, a.Recv(recv_l, slab_message_bytes(r[1]), kNcclUint8, e->comm_rank - 1, comm, e->stream));
      e->bytes_sent += slab_message_bytes(r[0]);
    }
    if (right) {
      NCCL_TRY(e, a.Send(send_r, slab_message_bytes(r[2]), kNcclUint8, e->comm_rank + 1, comm, e->stream));
      NCCL_TRY(e, a.Recv(recv_r, slab_message_bytes(r[3]), kNcclUint8, e->comm_rank + 1, comm, e->stream));
      e->bytes_sent += slab_message_bytes(r[2]);
    }
    NCCL_TRY(e, a.GroupEnd());
    return 0;
  };
  if (int rc = exchange(e->has_left, e->has_right, rec)) return rc;
  mark(2);
  if (int rc = slab_read_counts(e, recv_l, recv_r)) return rc;  // host sync
  {
    const uint32_t* hc = e->host_counts;
    const bool redo_l = e->has_left && (hc[8] > rec[0] || hc[0] > rec[1]);
    const bool redo_r = e->has_right && (hc[10] > rec[2] || hc[4] > rec[3]);
    if (redo_l || redo_r) {
      const uint32_t full[4] = {e->exch_cap, e->exch_cap, e->exch_cap, e->exch_cap};
      if (int rc = exchange(redo_l, redo_r, full)) return rc;
      if (int rc = slab_read_counts(e, recv_l, recv_r)) return rc;
      ++e->resends;
    }
    const uint32_t now[4] = {hc[8], hc[0], hc[10], hc[4]};
    for (int k = 0; k < 4; ++k) e->last_records[k] = now[k] > e->exch_cap ? e->exch_cap : now[k];
  }
  if (int rc = slab_apply_unpack(e, recv_l, recv_r, &n_owned, &n_ghost)) return rc;
  }
  t1 = clk::now();
  mark(3);
  if (int rc = sphb_slab_forces(e, e->gmin, e->gmax)) return rc;
  mark(4);
  float dt = dt_in, computed = dt_in;
  int attempts = 0;
  do {
    dt = computed;
    if (int rc = sphb_slab_advect(e, dt, e->red8_dev)) return rc;
    if (attempts == 0) mark(5);
    if (int rc = reduce8(e, seg && attempts == 0 ? e->seg_ev[6] : nullptr)) return rc;  // host sync
    if (attempts == 0) t2 = clk::now();
    computed = dt_from_maxima(e->params, e->red8_host[6], e->red8_host[7]);
    ++attempts;
  } while (dt - computed > 0.00001);  // sph_simulation.cpp:260
  if (int rc = sphb_slab_commit(e)) return rc;
  bbox_from_red8(e);
  if (seg) {
    cudaEventRecord(e->seg_ev[7], e->stream);  // after the last attempt's reduction
    CUDA_TRY(e, cudaEventSynchronize(e->seg_ev[7]));
    float ms = 0.f;
    if (seg_prev && cudaEventElapsedTime(&ms, e->seg_ev[8], e->seg_ev[0]) == cudaSuccess) e->seg_ms[7] += ms;
    for (int k = 0; k < 7; ++k)
      if (cudaEventElapsedTime(&ms, e->seg_ev[k], e->seg_ev[k + 1]) == cudaSuccess) e->seg_ms[k] += ms;
    cudaEventRecord(e->seg_ev[8], e->stream);
    ++e->seg_steps;
  }
  // wall-clock per phase, split at the step's host synchronisation points
  const clk::time_point t3 = clk::now();
  e->phase_ms[0] += std::chrono::duration<double, std::milli>(t1 - t0).count();
  e->phase_ms[1] += std::chrono::duration<double, std::milli>(t2 - t1).count();
  e->phase_ms[2] += std::chrono::duration<double, std::milli>(t3 - t2).count();
  if (dt_out) *dt_out = computed;
  if (attempts_out) *attempts_out = attempts;
  if (n_owned_out) *n_owned_out = n_owned;
  if (n_ghost_out) *n_ghost_out = n_ghost;
  return 0;
}

uint64_t sphb_slab_bytes_sent(sphb_engine* e) { return e ? e->bytes_sent : 0; }
int sphb_slab_p2p_active(sphb_engine* e) { return e && e->p2p_active ? 1 : 0; }
uint64_t sphb_slab_resends(sphb_engine* e) { return e ? e->resends : 0; }

int sphb_slab_gpu_segments(sphb_engine* e, double out8[8], uint64_t* steps, int reset) {
  if (!e || !out8) return SPHB_ERR_INVALID;
  for (int k = 0; k < 8; ++k) {
    out8[k] = e->seg_ms[k];
    if (reset) e->seg_ms[k] = 0.0;
  }
  if (steps) *steps = e->seg_steps;
  if (reset) e->seg_steps = 0;
  return 0;
}

int sphb_slab_phase_ms(sphb_engine* e, double out3[3], int reset) {
  if (!e || !out3) return SPHB_ERR_INVALID;
  for (int k = 0; k < 3; ++k) {
    out3[k] = e->phase_ms[k];
    if (reset) e->phase_ms[k] = 0.0;
  }
  return 0;
}

}  // extern "C"
