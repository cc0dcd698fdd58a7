/* sphb — the extern "C" boundary of the B200 SPH step engine.
 *
 * The reference (ngromko/libclsph) has no plugin / FFI interface: its hot path is the
 * private half of `class sph_simulation`, which talks to OpenCL through cl::Kernel and
 * cl::Buffer members (libclsph/sph_simulation.h:29-72).  This header is what replaces
 * that private half.  The C++11 `sph_simulation` in libclsph_b200/host/ keeps the
 * reference's public surface (libclsph/sph_simulation.h:10-27) and calls ONLY these
 * entry points; any other host language binds them the same way.
 *
 * Conventions
 *   - plain C ABI: POD arguments, caller-owned host memory, engine-owned device memory,
 *     no callbacks across the boundary, no torch / STL types in any signature;
 *   - every call returns 0 on success; > 0 is a cudaError_t (or an ncclResult_t offset by
 *     SPHB_ERR_NCCL_BASE), < 0 is one of the SPHB_ERR_* codes below; sphb_last_error()
 *     gives the text with the file:line of the failing call (the C++ wrapper prints it and
 *     exit(-1)s, the reference's EXIT_ON_CL_ERROR convention, util/cl_boilerplate.h:32-34);
 *   - one engine = one GPU = one CUDA stream; an engine is not thread-safe, but different
 *     engines may be driven from different threads of one process (sph_simulation::devices
 *     steps one slab engine per GPU that way), and sphb_download_wait may be called from a
 *     thread other than the one that steps the engine;
 *   - there is NO CPU fallback: without a CUDA device sphb_create() fails.
 *
 * Struct layouts are the reference's, byte for byte (OpenCL float3 = 16 bytes):
 * libclsph/common/structures.h:16-65.
 */
#ifndef SPHB_H_
#define SPHB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPHB_VERSION 1

/* engine error codes (< 0) */
#define SPHB_ERR_INVALID (-1)     /* bad argument / call order (e.g. step before upload) */
#define SPHB_ERR_GRID (-2)        /* a grid dimension reached 1024: the reference asserts here
                                     (libclsph/sph_simulation.cpp:722-724) */
#define SPHB_ERR_NO_DEVICE (-3)   /* no usable CUDA device */
#define SPHB_ERR_NO_SCENE (-4)    /* sphb_step before sphb_set_scene */
#define SPHB_ERR_NCCL_MISSING (-5)/* multi-GPU entry point used but libnccl could not be loaded */
#define SPHB_ERR_CAPACITY (-6)    /* slab engine: more particles / halo records than allocated */
#define SPHB_ERR_NCCL_BASE 100000 /* ncclResult_t r is returned as SPHB_ERR_NCCL_BASE + r */

typedef struct { float s[4]; } sphb_float3; /* cl_float3 */

/* libclsph/common/structures.h:40-44 — 80 bytes */
typedef struct {
  sphb_float3 position, velocity, intermediate_velocity, acceleration;
  float density, pressure;
  uint32_t grid_index;
  uint32_t pad_;
} sphb_particle;

/* libclsph/common/structures.h:16-38 — 128 bytes */
typedef struct {
  uint32_t particles_count;
  float fluid_density, total_mass, particle_mass, dynamic_viscosity;
  float simulation_time, h, simulation_scale, target_fps;
  float surface_tension_threshold, surface_tension, restitution, K;
  uint32_t pad0_[3];
  sphb_float3 constant_acceleration;
  int32_t grid_size_x, grid_size_y, grid_size_z;
  uint32_t grid_cell_count;
  sphb_float3 min_point, max_point;
} sphb_params;

/* libclsph/common/structures.h:46-52 — 20 bytes */
typedef struct {
  float poly_6, poly_6_gradient, poly_6_laplacian, spiky, viscosity;
} sphb_terms;

/* libclsph/common/structures.h:54-65 — 56 bytes */
typedef struct {
  float maxx, maxy, maxz, minx, miny, minz;
  size_t size_x, size_y, size_z, offset;
} sphb_bb;

typedef struct sphb_engine sphb_engine;

/* Replaces init_cl_single_device + make_program + the make_kernel block + the buffer
 * allocations of sph_simulation::simulate() (libclsph/sph_simulation.cpp:272-326).
 * `device` is a CUDA ordinal (-1 = current device).  params->particles_count fixes the
 * capacity; the physics constants are copied. */
int sphb_create(const sphb_params* params, const sphb_terms* terms, int device,
                sphb_engine** out);
int sphb_destroy(sphb_engine* e);

/* Replaces the bb_buffer_ upload (libclsph/sph_simulation.cpp:321-326) and
 * sph_simulation::computeDistanceField() (:607-632 -> kernels/collisionsv2.cl:57-138):
 * uploads the scene::load() products and builds the distance field on the device.
 * transforms: 12 floats per face, rvertices: 4 floats per face. */
int sphb_set_scene(sphb_engine* e, const sphb_bb* bbs, uint32_t face_count,
                   const float* transforms, const float* rvertices,
                   uint32_t total_grid_points);
/* Copies the device distance field (total_grid_points floats) to `out` — for parity tests. */
int sphb_get_distance_field(sphb_engine* e, float* out);
/* Device milliseconds the last sphb_set_scene spent in the distance-field kernel (0: no faces). */
float sphb_scene_build_ms(sphb_engine* e);

/* Replaces the initial enqueueWriteBuffer (libclsph/sph_simulation.cpp:334-336) and the
 * write-back after a pre/post-frame hook returned true (:743-745, :763-765).
 * n must equal particles_count. */
int sphb_upload(sphb_engine* e, const sphb_particle* aos, uint32_t n);
/* The same in two halves, like a non-blocking enqueueWriteBuffer: _begin starts the host ->
 * device copy (from page-locked memory, on a stream of its own, into a stage of its own) and
 * returns; the engine goes on stepping and reading back its current state; _commit orders the
 * engine's stream behind the copy and makes the uploaded particles the live state.  `aos` must
 * stay untouched until _commit's work has run (any later synchronising call).  With
 * sphb_download_begin this lets uploads, steps and read-backs of successive batches overlap
 * (PCIe is full duplex). */
int sphb_upload_begin(sphb_engine* e, const sphb_particle* aos, uint32_t n);
int sphb_upload_commit(sphb_engine* e);

/* Replaces sph_simulation::simulate_single_frame (libclsph/sph_simulation.cpp:200-263):
 * bbox -> locate_in_grid -> sort_particles (+cell table) -> density_pressure -> forces ->
 * { advection_collision(dt) ; computeTimeStep } until dt - computed <= 1e-5.
 * *dt_out = the value the reference returns (time advanced and next step's dt). */
int sphb_step(sphb_engine* e, float dt_in, float* dt_out);

/* Replaces the sub-step loop of one frame in sph_simulation::simulate()
 * (libclsph/sph_simulation.cpp:384-409, without write_all_frames):
 *     timeleft = frame_time;
 *     while (timeleft > 0) { dt = simulate_single_frame(dt); timeleft -= dt;
 *                            if (timeleft < dt) dt = timeleft; }
 * with the same results as calling sphb_step in that loop, but without a host round trip per
 * step: the next step's grid (findMinMaxPosition's host half, :668-727), the new dt
 * (computeTimeStep's host half, :598-602) and the advection retry (:260) are evaluated by
 * one-thread kernels, and a frame's steps are enqueued back to back (option "graph" = 1: one
 * CUDA graph launch per step).  The host steps in only when a cell table must grow, when dt
 * needs more than two advection attempts, and for the first step after an upload.
 * max_steps > 0 stops after that many steps.  *dt_out = the loop's dt afterwards; *steps_out =
 * steps run; timeleft_log[k] (k < log_cap) = timeleft after step k — what the reference prints
 * per step.  Several engines of one process may be inside this call on the same device at once:
 * the device-driven steps share per-process constants, so one engine at a time per device runs
 * them and any other engine steps from the host for that batch (same results, one host round
 * trip per step). */
int sphb_run_frame(sphb_engine* e, float dt_in, float frame_time, uint64_t max_steps, float* dt_out,
                   uint32_t* steps_out, float* timeleft_log, uint32_t log_cap);
/* Steps that sphb_run_frame ran without a host round trip since the engine was created. */
uint64_t sphb_queued_steps(sphb_engine* e);

/* The engine's copy of `parameters` as the reference mutates it every step
 * (min_point, max_point, grid_size_*, grid_cell_count; sph_simulation.cpp:670-727). */
int sphb_get_parameters(sphb_engine* e, sphb_params* out);

/* Replaces the enqueueReadBuffer of the whole state (libclsph/sph_simulation.cpp:415-417,
 * :736-738): AoS, in the device's current (cell-sorted) order, exactly the order the
 * reference would hold.  Padding lanes are written as 0. */
int sphb_download(sphb_engine* e, sphb_particle* aos);

/* The same read-back, split so that the device -> host copy overlaps later steps (SURVEY 8(f)-1;
 * the reference blocks in enqueueReadBuffer, sph_simulation.cpp:396-404): _begin snapshots the
 * state on the device and starts the copy on the engine's copy stream; `aos` (page-locked for
 * a truly asynchronous copy) is complete when _wait returns.  sphb_step may be called in
 * between; the next _begin / sphb_upload waits for a copy still in flight.  _wait may be
 * called from another host thread. */
int sphb_download_begin(sphb_engine* e, sphb_particle* aos);
int sphb_download_wait(sphb_engine* e);

/* Parity hooks for the last step (each pointer nullable):
 *   keys[N]   sorted grid_index of every particle,
 *   perm[N]   perm[k] = position BEFORE this step's sort of the particle now at k,
 *   cell_table[min(cap, grid_cell_count)]  the reference's cell table (kernels/sort.cl:61-78),
 *   attempts  how many times advection_collision was launched (>= 1). */
int sphb_download_debug(sphb_engine* e, uint32_t* keys, uint32_t* perm, uint32_t* cell_table,
                        uint32_t cell_table_cap, int* attempts);

/* Parity hook: runs ONLY advection_collision (kernels/sphb.cl:177-223) with time step dt over
 * the uploaded state (uploaded accelerations included), in place of a full step. */
int sphb_debug_advect(sphb_engine* e, float dt);

/* ---- slab engines: the step partitioned over the GPUs of one box ---------------------------
 * The reference is single-device (libclsph/sph_simulation.cpp:273-275); this section is new.
 * One process per GPU creates ONE slab engine that owns the particles with x in [xlo, xhi).
 * Per step the caller moves two fixed-size messages to / from its slab neighbours (NCCL send /
 * recv over NVLink) and all-reduces 8 floats per advection attempt; the engine does the rest:
 *
 *   sphb_slab_pack(e, send_left, send_right)      owned particles -> migrants + 2h ghost layer
 *   [send/recv sphb_slab_message_bytes(e) bytes with each neighbour]
 *   sphb_slab_unpack(e, recv_left, recv_right, &n_owned, &n_ghost)
 *   sphb_slab_forces(e, global_min, global_max)   grid of the GLOBAL bbox, density over owned +
 *                                                 ghosts, forces over owned
 *   do { sphb_slab_advect(e, dt, red8); [all-reduce MAX red8]; dt' = sphb_compute_dt(...); }
 *   while (dt - dt' > 1e-5)                       (sph_simulation.cpp:246-260)
 *   sphb_slab_commit(e)                           drop the ghosts; red8 = the next global bbox
 *
 * red8 (device floats): -min x, -min y, -min z, max x, max y, max z, max |v|^2, max |a|^2 over
 * the owned particles, so that ONE all-reduce(MAX) serves computeTimeStep and the next step's
 * findMinMaxPosition.  All message / red8 pointers are DEVICE pointers owned by the caller;
 * every call is enqueued on sphb_stream(e).  Particles carry a caller-chosen 31-bit id; there
 * is no global sorted order across slabs: downloads return (particle, id) pairs, with
 * grid_index the reference's Morton key under the global grid. */
int sphb_slab_create(const sphb_params* params /* particles_count = GLOBAL count */,
                     const sphb_terms* terms, int device, uint32_t capacity /* local particle slots */,
                     uint32_t exchange_capacity /* records per message */, sphb_engine** out);
size_t sphb_slab_message_bytes(sphb_engine* e); /* 16 + 48 * exchange_capacity */
/* A slab with a neighbour on both sides must be at least 3 h wide (the 2 h ghost layer + one
 * step's motion): ranks exchange with their two neighbours only.  A thinner one is refused
 * (SPHB_ERR_INVALID) — use fewer GPUs, or cut elsewhere. */
int sphb_slab_configure(sphb_engine* e, float xlo, float xhi, int has_left, int has_right);
int sphb_slab_upload(sphb_engine* e, const sphb_particle* aos, const uint32_t* ids, uint32_t n_owned);
int sphb_slab_bbox(sphb_engine* e, float* red8_dev); /* bbox of the owned positions only */
int sphb_slab_pack(sphb_engine* e, void* send_left_dev, void* send_right_dev);
int sphb_slab_unpack(sphb_engine* e, const void* recv_left_dev, const void* recv_right_dev,
                     uint32_t* n_owned, uint32_t* n_ghost); /* synchronises: the host learns the counts */
int sphb_slab_forces(sphb_engine* e, const float global_min[3], const float global_max[3]);
int sphb_slab_advect(sphb_engine* e, float dt, float* red8_dev);
int sphb_slab_commit(sphb_engine* e);
/* owned particles in local order + their ids (aos / ids nullable: just the count) */
int sphb_slab_download(sphb_engine* e, sphb_particle* aos, uint32_t* ids, uint32_t capacity,
                       uint32_t* n_owned);
/* computeTimeStep's host arithmetic (libclsph/sph_simulation.cpp:598-602) */
float sphb_compute_dt(const sphb_params* params, float max_v2, float max_a2);

/* ---- the same step driven natively: one call per step and rank (csrc/slab_nccl.cu) -------
 * The exchange (ncclSend / ncclRecv with the two x-neighbours) and the 8-float all-reduce are
 * issued on the engine's stream by the library itself; libnccl.so.2 is resolved at run time
 * (SPHB_ERR_NCCL_MISSING if it cannot be).  Set-up, collective over all ranks:
 *   rank 0: sphb_nccl_unique_id(id); [caller broadcasts the 128 bytes]
 *   sphb_slab_comm_init(e, id, rank, world)      rank r owns slab r; neighbours are r-1, r+1
 *   sphb_slab_configure + sphb_slab_upload, then sphb_slab_start(e)   (global bbox)
 * and then, per step,
 *   sphb_slab_step(e, dt_in, &dt_out, &attempts, &n_owned, &n_ghost)
 * which is pack / exchange / unpack / forces / { advect, all-reduce } / commit as listed above.
 * All ranks must pass the same dt_in; they all receive the same dt_out. */
int sphb_nccl_unique_id(unsigned char id_out[128]);
int sphb_slab_comm_init(sphb_engine* e, const unsigned char id[128], int rank, int world);
int sphb_slab_start(sphb_engine* e);
int sphb_slab_step(sphb_engine* e, float dt_in, float* dt_out, int* attempts_out, uint32_t* n_owned_out,
                   uint32_t* n_ghost_out);
uint64_t sphb_slab_bytes_sent(sphb_engine* e); /* halo bytes sent since sphb_slab_start */
/* 1 when sphb_slab_comm_init could map both neighbours' inboxes into this GPU's address space
 * (engines of one process: peer access; one process per GPU: CUDA IPC) on EVERY rank.  The
 * exchange then needs no ncclSend / ncclRecv: the pack kernel writes its records straight into
 * the neighbours' inboxes over NVLink, publishes the step's sequence number behind them, and the
 * receiving stream waits for that number (option "p2p_halo" = 0 before sphb_slab_comm_init keeps
 * the exchange on NCCL). */
int sphb_slab_p2p_active(sphb_engine* e);
/* A halo message carries last step's record count + max(1/8 of it, option "halo_margin" = 4096)
 * records (option value 0: no spare room at all, which exercises the resend path);
 * one that outgrew that is exchanged again at full capacity by the pair concerned: how often */
uint64_t sphb_slab_resends(sphb_engine* e);
/* host wall-clock milliseconds accumulated by sphb_slab_step, split at its synchronisation points:
 * pack + exchange + unpack | sort .. forces + first advection attempt + all-reduce | further
 * attempts + commit */
int sphb_slab_phase_ms(sphb_engine* e, double out3[3], int reset);
/* GPU-side time (CUDA events on the engine's stream, ms, accumulated over `*steps` steps taken with
 * sphb_profile_enable on) of the segments of sphb_slab_step: [0] pack, [1] waiting for the neighbours'
 * halos (peer memory) / send + receive (NCCL), [2] the count read-back — a host round trip during
 * which the GPU idles — and the unpack, [3] sort ... forces, [4] the first advection, [5] its
 * reduction + the all-reduce, [6] the read-back of the all-reduce and any further attempts, [7] idle
 * between the end of a step and the start of the next (host).  Measurement only. */
int sphb_slab_gpu_segments(sphb_engine* e, double out8[8], uint64_t* steps, int reset);

/* ---- measurement ------------------------------------------------------------------ */
enum {
  SPHB_PASS_BBOX = 0,    /* min/max of positions (only after an upload; otherwise fused in advect) */
  SPHB_PASS_LOCATE,      /* cell key per particle */
  SPHB_PASS_SORT,        /* key/index radix sort, all passes */
  SPHB_PASS_REORDER,     /* gather of pos / vel / v_half into sorted SoA */
  SPHB_PASS_CELLTABLE,   /* cell start offsets of the search grid */
  SPHB_PASS_DENSITY,     /* density + pressure */
  SPHB_PASS_FORCES,      /* forces -> acceleration */
  SPHB_PASS_ADVECT,      /* advection + collision + max|v|^2, max|a|^2, bbox reductions */
  SPHB_PASS_REFRANK,     /* the reference's cell table + every particle's place in its sorted array */
  SPHB_PASS_COUNT
};
/* When enabled, every pass of sphb_step is bracketed by CUDA events on the engine's
 * stream; sphb_get_pass_times returns the accumulated milliseconds and launch counts
 * since the last reset (this synchronises the stream). */
int sphb_profile_enable(sphb_engine* e, int on);
int sphb_profile_reset(sphb_engine* e);
int sphb_get_pass_times(sphb_engine* e, float ms[SPHB_PASS_COUNT],
                        uint32_t launches[SPHB_PASS_COUNT], uint32_t* steps);
/* Caller-side timing on the engine's stream: record one of 4 event slots, then read the
 * milliseconds between two recorded slots (synchronises on the later one). */
int sphb_event_record(sphb_engine* e, int slot);
int sphb_event_elapsed_ms(sphb_engine* e, int from_slot, int to_slot, float* ms);
/* number of kernel launches issued by this engine since creation */
uint64_t sphb_launch_count(sphb_engine* e);
/* What the last step's density pass handed to its forces pass (measurement only; new — the
 * reference searches twice, kernels/sphb.cl:3-35 and :37-126): out[0] = pairs of particles whose
 * neighbour list overflowed (their forces come from a second search), out[1] = pairs walked in
 * two parts, out[2] = list entries over all pairs, out[3] = entries of the longest list. */
int sphb_neighbor_list_stats(sphb_engine* e, uint64_t out[4]);
/* the engine's cudaStream_t (as void*), so a caller can record its own events on it */
void* sphb_stream(sphb_engine* e);
/* named integer options; unknown names return SPHB_ERR_INVALID.
 *   "neighbor_kernel"  0 (default) = the density pass records ONE neighbour list per pair of
 *                      consecutive particles that the forces pass consumes.  The others are kept
 *                      for measurement (same results within the tolerance): 1 = round-1 density +
 *                      forces by their own search; 2 = the round-1 kernels (one list per
 *                      particle); 4 / 5 / 6 = sub-warp-cooperative density with 4 / 8 / 16 lanes
 *                      per particle + forces by their own search; 7 = density with four
 *                      particles per thread (the default's lists and forces)
 *   "df_face_lists"    1 (default) = the distance-field build tests, per grid point, only the
 *                      faces whose box meets the box of the face that owns the point; 0 = all
 *                      faces, as kernels/collisionsv2.cl:70-76 does (set before sphb_set_scene).
 *                      Same field either way.
 *   "sort"             0 (default) = counting sort by search cell (the scan of the counts is the
 *                      cell table), made stable by a per-cell canonical order; 1 = LSD radix
 *                      sort of (key, index) pairs.  Both give the same order.
 *   "scan"             1 (default) = the prefix scans of the cell tables are one pass over the
 *                      table (decoupled look-back); 0 = three kernels (tile scans, scan of the
 *                      tile totals, add).  Same tables.  PROCESS-wide: applies to every engine.
 *   "triangle_bin"     1 (default) = scenes of 17+ faces get a uniform-grid triangle bin for the
 *                      collision pass; 0 = every particle tests every face (set before
 *                      sphb_set_scene)
 *   "side_stream"      1 (default) = the reference-order bookkeeping runs on a side stream
 *   "device_loop"      sphb_run_frame: 1 = steps are enqueued without host round trips; 0 = one
 *                      sphb_step per step; 2 (default) = 1 up to 6 M particles, where a step is
 *                      short enough for the round trip to show, else 0
 *   "graph"            1 (default) = a device-driven step is one CUDA graph launch; 0 = its
 *                      kernels are launched one by one
 *   "queued_attempts"  advection attempts enqueued per device-driven step (default 2, 1..4);
 *                      a step that needs more is finished from the host
 *   "table_headroom"   1 (default) = cell tables are allocated with 1.5x spare entries; 0 = exact
 *                      (tests: every grid growth then goes through the host)
 *   "halo_margin"      sphb_slab_step: spare records per halo message (NCCL exchange only)
 *   "p2p_halo"         1 (default) = sphb_slab_comm_init tries to set up the exchange over peer
 *                      memory (sphb_slab_p2p_active); 0 = NCCL send / receive */
int sphb_set_option(sphb_engine* e, const char* name, int value);

/* Page-locked host memory for the particle array (the reference asks OpenCL for
 * CL_MEM_ALLOC_HOST_PTR buffers, libclsph/sph_simulation.cpp:311-316): uploads and downloads
 * from such memory run at full PCIe rate.  Returns a cudaError_t. */
int sphb_host_alloc(void** out, size_t bytes);
int sphb_host_free(void* p);

/* ---- entry points that allocate nothing ---------------------------------------------------
 * No counterpart in the reference: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  sphb_cell_table_noalloc computes the reference's cell table — cell_table[c] = number
 * of particles whose Morton key (grid_index) is < c, c in [0, params->grid_cell_count): what
 * sphb_download_debug returns as `cell_table`, without the c >= N quirk of kernels/sort.cl:61-78 —
 * from DEVICE-resident positions (n x {x, y, z, -} floats, 16-byte aligned) under the grid in
 * `params` (h, min_point, grid_size_*, grid_cell_count, e.g. from sphb_get_parameters), into the
 * DEVICE array cell_table_dev (cell_table_entries >= grid_cell_count), enqueued on `stream` (a
 * cudaStream_t, NULL = the default stream) of the CURRENT device; it does not synchronise.
 * It calls no cudaMalloc / cudaFree / allocating routine: the scan's scratch (one partial sum per
 * 2048 cells) is a static device array of SPHB_NOALLOC_MAX_CELLS / 2048 words = 8 KB that every
 * GPU which loads the library holds until exit.  grid_cell_count > SPHB_NOALLOC_MAX_CELLS returns
 * SPHB_ERR_CAPACITY (text from sphb_last_error(NULL)).  Every call shares that array: calls that
 * overlap on one GPU (two streams, two threads) race; calls on one stream are ordered. */
#define SPHB_NOALLOC_MAX_CELLS 4194304u
int sphb_cell_table_noalloc(const float* positions_dev, uint32_t n, const sphb_params* params,
                            uint32_t* cell_table_dev, uint32_t cell_table_entries, void* stream);

const char* sphb_last_error(sphb_engine* e);
/* text for a return code when no engine exists (sphb_create failed) */
const char* sphb_error_string(int code);

#ifdef __cplusplus
}
#endif
#endif /* SPHB_H_ */
