/* C API shared by the two CPU checkers.  TEST INFRASTRUCTURE ONLY.
 *
 *   oracle/_ref/libsphref.so   "reference": the reference's own kernel source
 *                              (libclsph/kernels/sph.cl) compiled as C++ behind
 *                              oracle/ref_shim/clc_shim.h, driven by ref_driver.cpp
 *   oracle/liboracle.so        "port": oracle/sph_oracle.c, a plain-C restatement
 *
 * Both export exactly these symbols with exactly these semantics, so tests can
 * run either.  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
 * legs may load them; the product (libclsph_b200/) never does.
 *
 * All particle arrays are the reference's 80-byte AoS `particle`
 * (libclsph/common/structures.h:40-44 with OpenCL's 16-byte float3).
 */
#ifndef ORACLE_API_H_
#define ORACLE_API_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* layouts (bytes): particle 80, simulation_parameters 128,
 * precomputed_kernel_values 20, BB 56 */
typedef struct { float s[4]; } orc_float3;

typedef struct {
  orc_float3 position, velocity, intermediate_velocity, acceleration;
  float density, pressure;
  uint32_t grid_index;
  uint32_t pad_;
} orc_particle;

typedef struct {
  uint32_t particles_count;
  float fluid_density, total_mass, particle_mass, dynamic_viscosity;
  float simulation_time, h, simulation_scale, target_fps;
  float surface_tension_threshold, surface_tension, restitution, K;
  uint32_t pad0_[3];
  orc_float3 constant_acceleration;
  int32_t grid_size_x, grid_size_y, grid_size_z;
  uint32_t grid_cell_count;
  orc_float3 min_point, max_point;
} orc_params;

typedef struct {
  float poly_6, poly_6_gradient, poly_6_laplacian, spiky, viscosity;
} orc_terms;

typedef struct {
  float maxx, maxy, maxz, minx, miny, minz;
  size_t size_x, size_y, size_z, offset;
} orc_bb;

/* "reference" or "port" */
const char* orc_kind(void);
/* out[0..3] = sizeof particle, params, terms, BB as the checker compiled them */
void orc_sizes(uint32_t out[4]);
/* number of OpenMP threads the data-parallel loops use (0 = library default) */
void orc_set_threads(int n);
int orc_get_threads(void);

/* sph_simulation.cpp:634-728 — bbox of positions -> min/max_point, grid_size_*, grid_cell_count.
 * returns 0, or -1 if a grid dimension is >= 1024 (the reference asserts). */
int orc_find_min_max(const orc_particle* in, orc_params* params);
/* kernels/grid.cl:40-64 */
void orc_locate(const orc_particle* in, orc_particle* out, const orc_params* params);
/* sph_simulation.cpp:110-198 + kernels/sort.cl — 4-pass stable LSD radix sort of whole
 * particles + cell table.  Sorted array ends in `first`.  perm_out (nullable):
 * perm_out[k] = index in the INPUT `first` array of the particle now at k.
 * cell_table must hold params->grid_cell_count entries. thread_count = the T of the reference. */
void orc_sort(orc_particle* first, orc_particle* second, uint32_t* cell_table,
              const orc_params* params, uint32_t* perm_out, int thread_count);
/* kernels/sphb.cl:3-35 */
void orc_density(const orc_particle* in, orc_particle* out, const orc_params* params,
                 const orc_terms* terms, const uint32_t* cell_table);
/* kernels/sphb.cl:37-126 */
void orc_forces(const orc_particle* in, orc_particle* out, const orc_params* params,
                const orc_terms* terms, const uint32_t* cell_table);
/* kernels/sphb.cl:177-223 */
void orc_advect(const orc_particle* in, orc_particle* out, uint32_t n, float restitution,
                float dt, const float* df, const orc_bb* bbs, uint32_t face_count);
/* sph_simulation.cpp:536-605 */
float orc_compute_dt(const orc_particle* in, const orc_params* params);
/* kernels/collisionsv2.cl:57-138 */
void orc_compute_df(float* df, const orc_bb* bbs, const float* transforms,
                    const float* rvertices, uint32_t face_count, uint32_t gridcount);

/* sph_simulation.cpp:200-263 — one simulate_single_frame.  State in `front`, result in
 * `front`; `back` is scratch.  params is updated (min/max_point, grid sizes).
 * Outputs (each nullable): perm_out[N] as in orc_sort relative to the input order,
 * cell_table_out[cell_table_cap] (first grid_cell_count entries written),
 * attempts_out = number of advection launches.  Returns the computed dt, or NaN with
 * *attempts_out = -1 if the grid assert would fire / cell_table_cap is too small. */
float orc_step(orc_particle* front, orc_particle* back, orc_params* params,
               const orc_terms* terms, const float* df, const orc_bb* bbs,
               uint32_t face_count, float dt, uint32_t* perm_out,
               uint32_t* cell_table_out, uint32_t cell_table_cap, int* attempts_out);

/* sph_simulation.cpp:519-531 — total_mass, h and the 5 kernel constants from
 * particles_count, particle_mass, fluid_density (already set in params) */
void orc_derive(orc_params* params, orc_terms* terms, int particles_inside_influence_radius,
                float* initial_volume_out);
/* sph_simulation.cpp:52-98 — cube lattice initial state (all other fields zero) */
void orc_init_particles(orc_particle* buffer, const orc_params* params, float initial_volume);

/* per-pass wall-clock seconds of the last orc_step:
 * [0] minmax [1] locate [2] sort+cell table [3] density [4] forces [5] advect+dt */
void orc_last_step_times(double out[6]);

#ifdef __cplusplus
}
#endif
#endif
