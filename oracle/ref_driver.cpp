/* "reference" checker: drives the reference's own OpenCL kernel source on the CPU.
 * TEST INFRASTRUCTURE ONLY — never linked into the product.
 *
 * REF_KERNEL_BLOB is a build-time path to a sed-patched temporary copy of
 * /root/reference/libclsph/kernels/sph.cl (one expression patched, sort.cl:19's
 * OpenCL vector literal; see oracle/Makefile).  The host-side orchestration below
 * restates libclsph/sph_simulation.cpp (cited per function); the kernels with
 * barriers (prefix_sum_1/2/3 and the four reductions) are replaced by exact host
 * loops (an exclusive scan over the [bucket][thread] counters, min/max loops) —
 * both are order-independent, so results are unaffected.
 */
#include "ref_shim/clc_shim.h"

#include <chrono>
#include <vector>
#include <omp.h>

namespace clref {
thread_local work_item_ctx g_wi = {0, 0, 0, 1, 1};
#include REF_KERNEL_BLOB
}  // namespace clref

/* qualifier macros must not leak into the rest of this file */
#undef kernel
#undef __kernel
#undef global
#undef __global
#undef local
#undef __local
#undef constant
#undef __const

#include "oracle_api.h"

static_assert(sizeof(clref::particle) == sizeof(orc_particle), "particle layout");
static_assert(sizeof(clref::simulation_parameters) == sizeof(orc_params), "params layout");
static_assert(sizeof(clref::precomputed_kernel_values) == sizeof(orc_terms), "terms layout");
static_assert(sizeof(clref::BB) == sizeof(orc_bb), "BB layout");
static_assert(sizeof(orc_particle) == 80 && sizeof(orc_params) == 128 && sizeof(orc_bb) == 56,
              "OpenCL layouts");

namespace {

int g_threads = 0;
double g_times[6];

typedef clref::particle P;
typedef clref::simulation_parameters SP;
typedef clref::precomputed_kernel_values KV;

inline const P* cp(const orc_particle* p) { return reinterpret_cast<const P*>(p); }
inline P* mp(orc_particle* p) { return reinterpret_cast<P*>(p); }
inline const SP& sp(const orc_params* p) { return *reinterpret_cast<const SP*>(p); }
inline const KV& kv(const orc_terms* p) { return *reinterpret_cast<const KV*>(p); }

inline int nthreads() { return g_threads > 0 ? g_threads : omp_get_max_threads(); }

/* run `body()` once per work-item of an NDRange(n) with a 1-wide work-group */
template <typename F>
void for_each_work_item(size_t n, F body) {
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads())
  for (size_t gid = 0; gid < n; ++gid) {
    clref::g_wi.global_id = gid;
    clref::g_wi.local_id = 0;
    clref::g_wi.group_id = gid;
    clref::g_wi.local_size = 1;
    clref::g_wi.global_size = n;
    body();
  }
}

double now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

extern "C" {

const char* orc_kind(void) { return "reference"; }

void orc_sizes(uint32_t out[4]) {
  out[0] = sizeof(P);
  out[1] = sizeof(SP);
  out[2] = sizeof(KV);
  out[3] = sizeof(clref::BB);
}

void orc_set_threads(int n) { g_threads = n; }
int orc_get_threads(void) { return nthreads(); }

/* libclsph/sph_simulation.cpp:634-728 (host part; minimum_pos/maximum_pos replaced by loops) */
int orc_find_min_max(const orc_particle* in_, orc_params* params_) {
  const P* in = cp(in_);
  SP& parameters = *reinterpret_cast<SP*>(params_);
  const unsigned n = parameters.particles_count;
  float grid_cell_side_length = (parameters.h * 2);

  float mn[3] = {INFINITY, INFINITY, INFINITY};
  float mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (unsigned i = 0; i < n; ++i) {
    const float e[3] = {in[i].position.x, in[i].position.y, in[i].position.z};
    for (int a = 0; a < 3; ++a) {
      mn[a] = clref::min(mn[a], e[a]);
      mx[a] = clref::max(mx[a], e[a]);
    }
  }
  parameters.min_point.x = mn[0] - 2 * grid_cell_side_length;
  parameters.min_point.y = mn[1] - 2 * grid_cell_side_length;
  parameters.min_point.z = mn[2] - 2 * grid_cell_side_length;
  parameters.max_point.x = mx[0] + 2 * grid_cell_side_length;
  parameters.max_point.y = mx[1] + 2 * grid_cell_side_length;
  parameters.max_point.z = mx[2] + 2 * grid_cell_side_length;

  parameters.grid_size_x = static_cast<unsigned>(
      (parameters.max_point.x - parameters.min_point.x) / (grid_cell_side_length));
  parameters.grid_size_y = static_cast<unsigned>(
      (parameters.max_point.y - parameters.min_point.y) / (grid_cell_side_length));
  parameters.grid_size_z = static_cast<unsigned>(
      (parameters.max_point.z - parameters.min_point.z) / (grid_cell_side_length));

  if (!(parameters.grid_size_x < 1024) || !(parameters.grid_size_y < 1024) ||
      !(parameters.grid_size_z < 1024))
    return -1;

  parameters.grid_cell_count = clref::get_grid_index_z_curve(
      parameters.grid_size_x, parameters.grid_size_y, parameters.grid_size_z);
  return 0;
}

void orc_locate(const orc_particle* in, orc_particle* out, const orc_params* params) {
  const SP p = sp(params);
  const P* i = cp(in);
  P* o = mp(out);
  for_each_work_item(p.particles_count, [&] { clref::locate_in_grid(i, o, p); });
}

/* libclsph/sph_simulation.cpp:110-198 */
void orc_sort(orc_particle* first_, orc_particle* second_, uint32_t* cell_table,
              const orc_params* params, uint32_t* perm_out, int thread_count) {
  const SP parameters = sp(params);
  const unsigned n = parameters.particles_count;
  const int kBucketCount = 256, kRadixWidth = 8;
  /* the reference needs T <= N (get_start_and_end underflows otherwise, sort.cl:10-16) */
  int T = thread_count > 0 ? thread_count : 1024;
  if ((unsigned)T > n) T = n > 0 ? (int)n : 1;
  P* cur_in = mp(first_);
  P* cur_out = mp(second_);

  /* tag each particle with its input index in the padding lane of `position`; the
   * whole-struct copies of sort.cl:57 carry it through the four passes */
  if (perm_out) {
    for (unsigned i = 0; i < n; ++i) {
      uint32_t tag = i;
      std::memcpy(&cur_in[i].position.w, &tag, 4);
    }
  }

  std::vector<unsigned> counts((size_t)T * kBucketCount), starts((size_t)T * kBucketCount);
  for (int pass_number = 0; pass_number < 4; ++pass_number) {
    std::fill(counts.begin(), counts.end(), 0u); /* fillUintArray(counts, 0) */
#pragma omp parallel for num_threads(nthreads())
    for (int t = 0; t < T; ++t) {
      clref::g_wi.global_id = t;
      clref::g_wi.global_size = T;
      clref::sort_count(cur_in, counts.data(), parameters, T, pass_number, kRadixWidth);
    }
    /* prefix_sum_1/2/3 == exclusive scan over the T*256 counters */
    unsigned acc = 0;
    for (size_t k = 0; k < counts.size(); ++k) {
      starts[k] = acc;
      acc += counts[k];
    }
#pragma omp parallel for num_threads(nthreads())
    for (int t = 0; t < T; ++t) {
      clref::g_wi.global_id = t;
      clref::g_wi.global_size = T;
      clref::sort(cur_in, cur_out, starts.data(), parameters, T, pass_number, kRadixWidth);
    }
    P* tmp = cur_in;
    cur_in = cur_out;
    cur_out = tmp;
  }
  /* after 4 passes cur_in == first */

  if (perm_out) {
    for (unsigned k = 0; k < n; ++k) {
      uint32_t tag;
      std::memcpy(&tag, &cur_in[k].position.w, 4);
      perm_out[k] = tag;
      cur_in[k].position.w = 0.f;
    }
  }

  for (unsigned c = 0; c < parameters.grid_cell_count; ++c) { /* fillUintArray(cell_table, N) */
    clref::g_wi.global_id = c;
    clref::fillUintArray(cell_table, parameters.particles_count, parameters.grid_cell_count);
  }
  /* fill_cell_table: serial, in work-item order (writes of different items never overlap) */
  for (unsigned i = 0; i < n; ++i) {
    clref::g_wi.global_id = i;
    clref::fill_cell_table(cur_in, cell_table, parameters.particles_count,
                           parameters.grid_cell_count);
  }
}

void orc_density(const orc_particle* in, orc_particle* out, const orc_params* params,
                 const orc_terms* terms, const uint32_t* cell_table) {
  const SP p = sp(params);
  const KV k = kv(terms);
  const P* i = cp(in);
  P* o = mp(out);
  for_each_work_item(p.particles_count, [&] { clref::density_pressure(i, o, p, k, cell_table); });
}

void orc_forces(const orc_particle* in, orc_particle* out, const orc_params* params,
                const orc_terms* terms, const uint32_t* cell_table) {
  const SP p = sp(params);
  const KV k = kv(terms);
  const P* i = cp(in);
  P* o = mp(out);
  for_each_work_item(p.particles_count, [&] {
    P local_data[1]; /* local tile of one: get_local_id() == 0 */
    clref::forces(i, o, local_data, p, k, cell_table);
  });
}

void orc_advect(const orc_particle* in, orc_particle* out, uint32_t n, float restitution,
                float dt, const float* df, const orc_bb* bbs, uint32_t face_count) {
  const P* i = cp(in);
  P* o = mp(out);
  const clref::BB* b = reinterpret_cast<const clref::BB*>(bbs);
  const KV unused = KV();
  for_each_work_item(n, [&] {
    clref::advection_collision(i, o, restitution, dt, unused, (const unsigned*)0, df, b, face_count);
  });
}

/* libclsph/sph_simulation.cpp:536-605 (maximum_vit / maximum_accel replaced by loops) */
float orc_compute_dt(const orc_particle* in_, const orc_params* params) {
  using namespace std;
  const P* in = cp(in_);
  const SP parameters = sp(params);
  float v2 = -INFINITY, a2 = -INFINITY;
  for (unsigned i = 0; i < parameters.particles_count; ++i) {
    clref::float3 e = in[i].velocity;
    float vit = e.x * e.x + e.y * e.y + e.z * e.z;
    v2 = clref::max(v2, vit);
    e = in[i].acceleration;
    vit = e.x * e.x + e.y * e.y + e.z * e.z;
    a2 = clref::max(a2, vit);
  }
  float maxVel2 = v2;
  float maxVel = sqrt(v2);
  float maxAccel = sqrt(a2);
  float dt;
  dt = (sqrt(2 * maxAccel * parameters.h + maxVel2) - maxVel) / (2 * maxAccel);
  if (dt < 0.00001) dt = 0.00001;
  if (dt > 1.0f / (parameters.target_fps) * parameters.simulation_scale)
    dt = 1.0f / (parameters.target_fps) * parameters.simulation_scale;
  return dt;
}

void orc_compute_df(float* df, const orc_bb* bbs, const float* transforms,
                    const float* rvertices, uint32_t face_count, uint32_t gridcount) {
  const clref::BB* b = reinterpret_cast<const clref::BB*>(bbs);
  for_each_work_item(gridcount, [&] {
    clref::computeDistanceField(df, b, transforms, rvertices, face_count, gridcount);
  });
}

/* libclsph/sph_simulation.cpp:200-263 */
float orc_step(orc_particle* front, orc_particle* back, orc_params* params,
               const orc_terms* terms, const float* df, const orc_bb* bbs,
               uint32_t face_count, float dt, uint32_t* perm_out, uint32_t* cell_table_out,
               uint32_t cell_table_cap, int* attempts_out) {
  double t0 = now();
  if (attempts_out) *attempts_out = -1;
  if (orc_find_min_max(front, params) != 0) return NAN;
  double t1 = now();
  orc_locate(front, back, params);
  double t2 = now();

  std::vector<uint32_t> own_table;
  uint32_t* cell_table = cell_table_out;
  if (!cell_table) {
    own_table.resize(params->grid_cell_count);
    cell_table = own_table.data();
  } else if (cell_table_cap < params->grid_cell_count) {
    return NAN;
  }
  orc_sort(back, front, cell_table, params, perm_out, 0);
  double t3 = now();
  orc_density(back, front, params, terms, cell_table);
  double t4 = now();
  orc_forces(front, back, params, terms, cell_table);
  double t5 = now();

  int attempts = 0;
  float cumputedTime = dt;
  do {
    dt = cumputedTime;
    orc_advect(back, front, params->particles_count, params->restitution, dt, df, bbs, face_count);
    ++attempts;
    cumputedTime = orc_compute_dt(front, params);
  } while (dt - cumputedTime > 0.00001);
  double t6 = now();

  g_times[0] = t1 - t0;
  g_times[1] = t2 - t1;
  g_times[2] = t3 - t2;
  g_times[3] = t4 - t3;
  g_times[4] = t5 - t4;
  g_times[5] = t6 - t5;
  if (attempts_out) *attempts_out = attempts;
  return cumputedTime;
}

/* libclsph/sph_simulation.cpp:519-531 — derived constants, same expressions */
void orc_derive(orc_params* params, orc_terms* terms, int particles_inside_influence_radius,
                float* initial_volume_out) {
  using namespace std;
  SP& parameters = *reinterpret_cast<SP*>(params);
  KV& precomputed_terms = *reinterpret_cast<KV*>(terms);
  float initial_volume;
  parameters.total_mass = parameters.particles_count * parameters.particle_mass;
  initial_volume = parameters.total_mass / parameters.fluid_density;
  parameters.h = cbrtf(3.f * (particles_inside_influence_radius *
                              (initial_volume / parameters.particles_count)) /
                       (4.f * M_PI));

  precomputed_terms.poly_6 = 315.f / (64.f * M_PI * pow(parameters.h, 9.f));
  precomputed_terms.poly_6_gradient = -945.f / (32.f * M_PI * pow(parameters.h, 9.f));
  precomputed_terms.poly_6_laplacian = -945.f / (32.f * M_PI * pow(parameters.h, 9.f));
  precomputed_terms.spiky = -45.f / (M_PI * pow(parameters.h, 6.f));
  precomputed_terms.viscosity = 45.f / (M_PI * pow(parameters.h, 6.f));
  if (initial_volume_out) *initial_volume_out = initial_volume;
}

/* libclsph/sph_simulation.cpp:52-98 — the cube lattice (no last_frame.bin) */
void orc_init_particles(orc_particle* buffer_, const orc_params* params, float initial_volume) {
  using namespace std;
  P* buffer = mp(buffer_);
  const SP parameters = sp(params);
  int particles_per_cube_side = ceil(cbrtf(parameters.particles_count));
  float side_length = cbrtf(initial_volume);
  float spacing = side_length / (float)particles_per_cube_side;
  std::memset(buffer_, 0, sizeof(orc_particle) * (size_t)parameters.particles_count);
  for (unsigned int i = 0; i < parameters.particles_count; ++i) {
    buffer[i].position.x = (float)(i % particles_per_cube_side) * spacing - side_length / 2.f;
    buffer[i].position.y =
        ((float)((i / particles_per_cube_side) % particles_per_cube_side) * spacing);
    buffer[i].position.z =
        (float)(i / (particles_per_cube_side * particles_per_cube_side)) * spacing -
        side_length / 2.f;
  }
}

void orc_last_step_times(double out[6]) {
  for (int i = 0; i < 6; ++i) out[i] = g_times[i];
}

}  // extern "C"
