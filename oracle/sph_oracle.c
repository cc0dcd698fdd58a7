/* "port" checker: a plain-C restatement of libclsph's per-step particle pipeline.
 *
 * TEST INFRASTRUCTURE ONLY.  Loaded by tests/, __graft_entry__.smoke() and the
 * CPU-baseline legs of bench.py as the CHECKER for the CUDA engine; never linked
 * into, imported by, or used as a fallback of the product (libclsph_b200/).
 *
 * Parity status: the reference ships no tests, golden vectors or fixtures for this
 * path, so it is PINNED AGAINST THE REFERENCE ITSELF RUN HERE:
 * tests/test_oracle_golden.py requires this file to agree bit-for-bit with
 * oracle/_ref/libsphref.so (the reference's own kernels/sph.cl compiled as C++),
 * and tests/golden/ holds vectors generated from that library.  What remains
 * unpinned is only the arithmetic of the OpenCL built-ins (pown, length, '/',
 * sqrt, FMA contraction), which belongs to the OpenCL device compiler, not to the
 * reference repo; the choices are stated in oracle/ref_shim/clc_shim.h and
 * mirrored here (IEEE sqrt and '/', pown = repeated multiply, no contraction).
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/libclsph/).  Build: oracle/Makefile (-ffp-contract=off).
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <omp.h>

#include "oracle_api.h"

typedef struct { float x, y, z; } v3;

static int g_threads = 0;
static double g_times[6];

static int nthreads(void) { return g_threads > 0 ? g_threads : omp_get_max_threads(); }

static double now(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* ---- OpenCL built-ins as the shim defines them (ref_shim/clc_shim.h) -------------- */
static inline v3 V3(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 ld3(const orc_float3* p) { return V3(p->s[0], p->s[1], p->s[2]); }
static inline void st3(orc_float3* p, v3 v) { p->s[0] = v.x; p->s[1] = v.y; p->s[2] = v.z; p->s[3] = 0.f; }
static inline v3 add3(v3 a, v3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub3(v3 a, v3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 mul3s(v3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
static inline v3 smul3(float s, v3 a) { return V3(s * a.x, s * a.y, s * a.z); }
static inline v3 div3s(v3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
static inline v3 neg3(v3 a) { return V3(-a.x, -a.y, -a.z); }
static inline float dot3(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline float len3(v3 a) { return sqrtf((a.x * a.x + a.y * a.y) + a.z * a.z); }
static inline float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
static inline float pown(float b, int n) {
  float r = b;
  for (int i = 1; i < n; ++i) r = r * b;
  return n == 0 ? 1.f : r;
}
/* the gate every smoothing function starts with: 1 - clamp(floor(r/h), 0, 1) */
static inline float gate(float r, float h) { return 1.f - clampf(floorf(r / h), 0.f, 1.f); }

/* ---- common/util.h:41-62 and :4-38 — 10-bit-per-axis Morton code -------------------- */
static uint32_t spread10(uint32_t x) {
  x = (x | (x << 16)) & 0x030000FF;
  x = (x | (x << 8)) & 0x0300F00F;
  x = (x | (x << 4)) & 0x030C30C3;
  x = (x | (x << 2)) & 0x09249249;
  return x;
}
static uint32_t morton(uint32_t x, uint32_t y, uint32_t z) {
  return spread10(x) | (spread10(y) << 1) | (spread10(z) << 2);
}
static uint32_t compact10(uint32_t v) { /* uninterleave, util.h:4-19: keeps bits 0,3,...,27 */
  uint32_t r = 0;
  for (int b = 0; b < 10; ++b) r |= (v & (1u << (3 * b))) >> (2 * b);
  return r;
}

/* ---- kernels/smoothing.cl --------------------------------------------------------- */
static inline float poly_6(float r, float h, const orc_terms* t) { /* :1-4 */
  return (gate(r, h) * t->poly_6) * pown((pown(h, 2) - pown(r, 2)), 3);
}
static inline v3 poly_6_gradient(v3 r, float h, const orc_terms* t) { /* :6-10 */
  float l = len3(r);
  return mul3s(smul3(gate(l, h) * t->poly_6_gradient, r), pown((pown(h, 2) - pown(l, 2)), 2));
}
static inline float poly_6_laplacian(float r, float h, const orc_terms* t) { /* :12-17 */
  float l = fabsf(r); /* length() of a scalar */
  return ((gate(l, h) * t->poly_6_laplacian) * (pown(h, 2) - pown(r, 2))) *
         (3.f * pown(h, 2) - 7.f * pown(r, 2));
}
#define SPH_EPSILON 0.0000001f
static inline v3 spiky_gradient(v3 r, float h, const orc_terms* t) { /* :21-30 */
  float l = len3(r);
  if (l - SPH_EPSILON < 0.f && 0.f < l + SPH_EPSILON) {
    float s = (-45.f / (float)(M_PI * pown(h, 6))); /* scalar, broadcast to xyz */
    return V3(s, s, s);
  }
  return mul3s(smul3(gate(l, h) * t->spiky, div3s(r, l)), pown(h - l, 2));
}
static inline float viscosity_laplacian(float r, float h, const orc_terms* t) { /* :32-34 */
  return (gate(r, h) * t->viscosity) * (h - r);
}

/* ---- kernels/grid.cl:19-29 ---------------------------------------------------------- */
static inline void cell_range(uint32_t cell, const uint32_t* table, const orc_params* p,
                              uint32_t* start, uint32_t* end) {
  *start = table[cell];
  *end = (p->grid_cell_count > cell + 1) ? table[cell + 1] : p->particles_count;
}

const char* orc_kind(void) { return "port"; }

void orc_sizes(uint32_t out[4]) {
  out[0] = sizeof(orc_particle);
  out[1] = sizeof(orc_params);
  out[2] = sizeof(orc_terms);
  out[3] = sizeof(orc_bb);
}

void orc_set_threads(int n) { g_threads = n; }
int orc_get_threads(void) { return nthreads(); }

/* ---- sph_simulation.cpp:634-728 + kernels/reduction.cl:2-64 ------------------------- */
int orc_find_min_max(const orc_particle* in, orc_params* p) {
  const uint32_t n = p->particles_count;
  float side = (p->h * 2);
  float mn[3] = {INFINITY, INFINITY, INFINITY};
  float mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (uint32_t i = 0; i < n; ++i)
    for (int a = 0; a < 3; ++a) {
      mn[a] = fminf(mn[a], in[i].position.s[a]);
      mx[a] = fmaxf(mx[a], in[i].position.s[a]);
    }
  int32_t* gs[3] = {&p->grid_size_x, &p->grid_size_y, &p->grid_size_z};
  for (int a = 0; a < 3; ++a) {
    p->min_point.s[a] = mn[a] - 2 * side; /* :670-672 */
    p->max_point.s[a] = mx[a] + 2 * side; /* :700-702 */
    *gs[a] = (int32_t)(uint32_t)((p->max_point.s[a] - p->min_point.s[a]) / (side)); /* :709-717 */
  }
  if (!(p->grid_size_x < 1024) || !(p->grid_size_y < 1024) || !(p->grid_size_z < 1024))
    return -1; /* :722-724 asserts */
  p->grid_cell_count = morton((uint32_t)p->grid_size_x, (uint32_t)p->grid_size_y,
                              (uint32_t)p->grid_size_z); /* :726-727 */
  return 0;
}

/* ---- kernels/grid.cl:40-64 ---------------------------------------------------------- */
void orc_locate(const orc_particle* in, orc_particle* out, const orc_params* p) {
  const uint32_t n = p->particles_count;
#pragma omp parallel for schedule(static) num_threads(nthreads())
  for (uint32_t i = 0; i < n; ++i) {
    out[i] = in[i];
    uint32_t cx = (uint32_t)((in[i].position.s[0] - p->min_point.s[0]) / (p->h * 2));
    uint32_t cy = (uint32_t)((in[i].position.s[1] - p->min_point.s[1]) / (p->h * 2));
    uint32_t cz = (uint32_t)((in[i].position.s[2] - p->min_point.s[2]) / (p->h * 2));
    out[i].grid_index = morton(cx, cy, cz);
  }
}

/* ---- sph_simulation.cpp:110-198 + kernels/sort.cl:23-78 ----------------------------- *
 * The reference runs 4 passes of an 8-bit LSD radix sort over whole particles with T
 * work-items, each owning a contiguous chunk, counters laid out [bucket][work-item] and
 * exclusive-scanned: that is a STABLE counting sort per pass whatever T is, so the port
 * uses T = 1.  The sorted array ends in `first` after the 4 ping-pong passes. */
void orc_sort(orc_particle* first, orc_particle* second, uint32_t* cell_table,
              const orc_params* p, uint32_t* perm_out, int thread_count) {
  (void)thread_count;
  const uint32_t n = p->particles_count;
  uint32_t* perm_a = NULL;
  uint32_t* perm_b = NULL;
  if (perm_out) {
    perm_a = (uint32_t*)malloc(sizeof(uint32_t) * (n ? n : 1));
    perm_b = (uint32_t*)malloc(sizeof(uint32_t) * (n ? n : 1));
    for (uint32_t i = 0; i < n; ++i) perm_a[i] = i;
  }
  orc_particle* cur_in = first;
  orc_particle* cur_out = second;
  for (int pass = 0; pass < 4; ++pass) {
    uint32_t start[256], count[256];
    memset(count, 0, sizeof count);
    for (uint32_t i = 0; i < n; ++i) /* sort_count, :23-39 */
      ++count[(cur_in[i].grid_index >> (pass * 8)) & 0xFF];
    uint32_t acc = 0;
    for (int b = 0; b < 256; ++b) { /* prefix_sum_1/2/3, :108-200 */
      start[b] = acc;
      acc += count[b];
    }
    for (uint32_t i = 0; i < n; ++i) { /* sort, :41-59 */
      uint32_t at = start[(cur_in[i].grid_index >> (pass * 8)) & 0xFF]++;
      cur_out[at] = cur_in[i];
      if (perm_out) perm_b[at] = perm_a[i];
    }
    orc_particle* t = cur_in; cur_in = cur_out; cur_out = t;
    uint32_t* tp = perm_a; perm_a = perm_b; perm_b = tp;
  }
  if (perm_out) {
    memcpy(perm_out, perm_a, sizeof(uint32_t) * n);
    free(perm_a);
    free(perm_b);
  }
  /* fillUintArray(cell_table, N, grid_cell_count), sph_simulation.cpp:186-190 */
  for (uint32_t c = 0; c < p->grid_cell_count; ++c) cell_table[c] = n;
  /* fill_cell_table, sort.cl:61-78 — one work-item per particle; including the quirk that
   * entries 0..key[0] are zeroed only by work-items that exist (id < N) */
  for (uint32_t w = 0; w < n; ++w) {
    uint32_t current_index = cur_in[w].grid_index;
    if (w <= cur_in[0].grid_index) cell_table[w] = 0;
    if (w > 0) {
      uint32_t diff = current_index - cur_in[w - 1].grid_index;
      for (uint32_t i = 0; i < diff; i++) {
        cell_table[current_index] = w;
        current_index--;
      }
    }
  }
}

/* ---- kernels/sphb.cl:3-35 + kernels/forces.cl:14-42 --------------------------------- */
void orc_density(const orc_particle* in, orc_particle* out, const orc_params* p,
                 const orc_terms* t, const uint32_t* table) {
  const uint32_t n = p->particles_count;
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads())
  for (uint32_t cur = 0; cur < n; ++cur) {
    orc_particle me = in[cur];
    float density = 0.f;
    uint32_t key = in[cur].grid_index;
    uint32_t cx = compact10(key & 0x9249249), cy = compact10((key >> 1) & 0x9249249),
             cz = compact10((key >> 2) & 0x9249249);
    v3 xi = ld3(&in[cur].position);
    for (uint32_t z = cz - 1; z <= cz + 1; ++z)
      for (uint32_t y = cy - 1; y <= cy + 1; ++y)
        for (uint32_t x = cx - 1; x <= cx + 1; ++x) {
          uint32_t s, e;
          cell_range(morton(x, y, z), table, p, &s, &e);
          for (size_t i = s; i < e; ++i) {
            float r = len3(sub3(xi, ld3(&in[i].position))); /* distance() */
            density += p->particle_mass * poly_6(r, p->h, t);
          }
        }
    me.density = density;
    out[cur] = me;
    /* Tait equation, sphb.cl:32-34 */
    out[cur].pressure = p->K * (pown(me.density / p->fluid_density, 7) - 1.f);
  }
}

/* ---- kernels/sphb.cl:37-126 --------------------------------------------------------- */
void orc_forces(const orc_particle* in, orc_particle* out, const orc_params* p,
                const orc_terms* t, const uint32_t* table) {
  const uint32_t n = p->particles_count;
  const float h = p->h, m = p->particle_mass;
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads())
  for (uint32_t cur = 0; cur < n; ++cur) {
    const orc_particle me = in[cur];
    out[cur] = me;
    v3 pressure_term = {0.f, 0.f, 0.f};
    v3 viscosity_term = {0.f, 0.f, 0.f};
    v3 normal = {0.f, 0.f, 0.f};
    float color_field_laplacian = 0.f;
    const v3 xi = ld3(&me.position), vi = ld3(&me.velocity);

    uint32_t key = me.grid_index;
    uint32_t cx = compact10(key & 0x9249249), cy = compact10((key >> 1) & 0x9249249),
             cz = compact10((key >> 2) & 0x9249249);
    for (uint32_t z = cz - 1; z <= cz + 1; ++z)
      for (uint32_t y = cy - 1; y <= cy + 1; ++y)
        for (uint32_t x = cx - 1; x <= cx + 1; ++x) {
          uint32_t s, e;
          cell_range(morton(x, y, z), table, p, &s, &e);
          for (size_t i = s; i < e; ++i) {
            const orc_particle* other = &in[i];
            const v3 dx = sub3(xi, ld3(&other->position));
            if (i != cur) {
              /* [kelager] (4.11), sphb.cl:80-87 */
              float w = (other->pressure / pown(other->density, 2) +
                         me.pressure / pown(me.density, 2)) * m;
              pressure_term = add3(pressure_term, smul3(w, spiky_gradient(dx, h, t)));
              /* sphb.cl:89-94 */
              v3 dv = sub3(ld3(&other->velocity), vi);
              viscosity_term = add3(
                  viscosity_term,
                  mul3s(mul3s(dv, (m / other->density)), viscosity_laplacian(len3(dx), h, t)));
            }
            /* sphb.cl:97-106 */
            normal = add3(normal, smul3(m / other->density, poly_6_gradient(dx, h, t)));
            color_field_laplacian += m / other->density * poly_6_laplacian(len3(dx), h, t);
          }
        }

    v3 sum = add3(smul3(-me.density, pressure_term), mul3s(viscosity_term, p->dynamic_viscosity));
    if (len3(normal) > p->surface_tension_threshold) {
      sum = add3(sum, div3s(smul3(-p->surface_tension * color_field_laplacian, normal), len3(normal)));
    }
    v3 a = div3s(sum, me.density);
    a = add3(a, ld3(&p->constant_acceleration));
    st3(&out[cur].acceleration, a);
  }
}

/* ---- kernels/collisionsv2.cl:140-156 ------------------------------------------------ */
static inline float weighted_average(float x, float x1, float x2, float d1, float d2) {
  return ((x2 - x) / (x2 - x1)) * d1 + ((x - x1) / (x2 - x1)) * d2;
}
static inline float bilinear(float x, float y, float xmin, float ymin, float xmax, float ymax,
                             float d00, float d01, float d10, float d11) {
  float R1 = weighted_average(x, xmin, xmax, d00, d10);
  float R2 = weighted_average(x, xmin, xmax, d01, d11);
  return weighted_average(y, ymin, ymax, R1, R2);
}
/* the index arithmetic is done in fp32 in the reference (float x,y,z; size_t fields convert) */
static inline int df_index(const orc_bb* bb, float x, float y, float z, short a, short b, short c) {
  return (int)((float)bb->offset + (y + b) * (float)bb->size_x * (float)bb->size_z +
               (float)bb->size_x * (z + c) + x + a);
}

/* ---- kernels/advection.cl:6-18 + kernels/sphb.cl:177-223 + collisionsv2.cl:9-22,249-336 - */
void orc_advect(const orc_particle* in, orc_particle* out, uint32_t n, float restitution,
                float dt, const float* df, const orc_bb* bbs, uint32_t face_count) {
#pragma omp parallel for schedule(static) num_threads(nthreads())
  for (uint32_t cur = 0; cur < n; ++cur) {
    orc_particle o = in[cur];
    const v3 old_position = ld3(&in[cur].position);
    const v3 vhalf = ld3(&in[cur].intermediate_velocity);
    const v3 acc = ld3(&in[cur].acceleration);
    /* leapfrog */
    v3 next = add3(vhalf, mul3s(acc, dt));
    v3 pos = add3(old_position, mul3s(next, dt));

    /* handle_collisions: LAST face whose inflated AABB holds the new position */
    int indice = -1;
    for (int i = 0; i < (int)face_count; i++) {
      if (pos.x <= bbs[i].maxx && pos.x >= bbs[i].minx && pos.y <= bbs[i].maxy &&
          pos.y >= bbs[i].miny && pos.z <= bbs[i].maxz && pos.z >= bbs[i].minz)
        indice = i;
    }
    v3 rpos = pos, rvel = next;
    if (indice > -1) {
      const orc_bb* bb = &bbs[indice];
      float sidex = (bb->maxx - bb->minx) / (float)(bb->size_x - 1);
      float sidey = (bb->maxy - bb->miny) / (float)(bb->size_y - 1);
      float sidez = (bb->maxz - bb->minz) / (float)(bb->size_z - 1);
      int x = (int)((pos.x - bb->minx) / (sidex));
      int y = (int)((pos.y - bb->miny) / (sidey));
      int z = (int)((pos.z - bb->minz) / (sidez));
      float bx = x * sidex + bb->minx;
      float by = y * sidey + bb->miny;
      float bz = z * sidez + bb->minz;
#define DF(a, b, c) df[df_index(bb, (float)x, (float)y, (float)z, a, b, c)]
      float facedown = bilinear(pos.x, pos.z, bx, bz, bx + sidex, bz + sidez, DF(0, 0, 0),
                                DF(0, 0, 1), DF(1, 0, 0), DF(1, 0, 1));
      float faceup = bilinear(pos.x, pos.z, bx, bz, bx + sidex, bz + sidez, DF(0, 1, 0),
                              DF(0, 1, 1), DF(1, 1, 0), DF(1, 1, 1));
      float d = weighted_average(pos.y, by, by + sidey, facedown, faceup);
      if (d < 0.02) { /* double compare, as written */
        float faceright = bilinear(pos.y, pos.z, by, bz, by + sidey, bz + sidez, DF(1, 0, 0),
                                   DF(1, 0, 1), DF(1, 1, 0), DF(1, 1, 1));
        float faceleft = bilinear(pos.y, pos.z, by, bz, by + sidey, bz + sidez, DF(0, 0, 0),
                                  DF(0, 0, 1), DF(0, 1, 0), DF(0, 1, 1));
        float faceback = bilinear(pos.x, pos.y, bx, by, bx + sidex, by + sidey, DF(0, 0, 0),
                                  DF(0, 1, 0), DF(1, 0, 0), DF(1, 1, 0));
        float facefront = bilinear(pos.x, pos.y, bx, by, bx + sidex, by + sidey, DF(0, 0, 1),
                                   DF(0, 1, 1), DF(1, 0, 1), DF(1, 1, 1));
#undef DF
        v3 nrm = {(faceright - faceleft), (faceup - facedown), (facefront - faceback)};
        float lenn = len3(nrm);
        nrm = div3s(nrm, lenn);
        /* respond(), :9-22 */
        float ad = fabsf(d);
        rpos = add3(pos, smul3(ad, nrm));
        float f = (1.f + restitution * ad / (dt * len3(next))) * dot3(next, nrm);
        rvel = sub3(next, smul3(f, nrm));
      }
    }
    st3(&o.velocity, div3s(add3(vhalf, rvel), 2.f));
    st3(&o.intermediate_velocity, rvel);
    st3(&o.position, rpos);
    out[cur] = o;
  }
}

/* ---- sph_simulation.cpp:536-605 + kernels/reduction.cl:66-124 ----------------------- */
float orc_compute_dt(const orc_particle* in, const orc_params* p) {
  float v2 = -INFINITY, a2 = -INFINITY;
  for (uint32_t i = 0; i < p->particles_count; ++i) {
    const float* e = in[i].velocity.s;
    float vit = e[0] * e[0] + e[1] * e[1] + e[2] * e[2];
    v2 = fmaxf(v2, vit);
    e = in[i].acceleration.s;
    vit = e[0] * e[0] + e[1] * e[1] + e[2] * e[2];
    a2 = fmaxf(a2, vit);
  }
  float maxVel2 = v2, maxVel = sqrtf(v2), maxAccel = sqrtf(a2);
  float dt = (sqrtf(2 * maxAccel * p->h + maxVel2) - maxVel) / (2 * maxAccel);
  if (dt < 0.00001) dt = 0.00001; /* compared and assigned in double, :600 */
  if (dt > 1.0f / (p->target_fps) * p->simulation_scale)
    dt = 1.0f / (p->target_fps) * p->simulation_scale;
  return dt;
}

/* ---- kernels/collisionsv2.cl:24-55 -------------------------------------------------- */
static inline float det2(float x1, float y1, float x2, float y2) { return x1 * y2 - y1 * x2; }

static float dist_point_segment(float x, float y, float z, float x1, float y1, float x2, float y2) {
  float A = y - x1;
  float B = z - y1;
  float C = x2 - x1;
  float D = y2 - y1;
  float dot = A * C + B * D;
  float len_sq = C * C + D * D;
  float param = -1;
  if (len_sq != 0) param = dot / len_sq;
  float xx, yy;
  if (param < 0) {
    xx = x1;
    yy = y1;
  } else if (param > 1) {
    xx = x2;
    yy = y2;
  } else {
    xx = x1 + param * C;
    yy = y1 + param * D;
  }
  float dy = y - xx;
  float dz = z - yy;
  return sqrtf(x * x + dz * dz + dy * dy);
}

/* ---- kernels/collisionsv2.cl:57-138 ------------------------------------------------- */
void orc_compute_df(float* df, const orc_bb* bbs, const float* tr, const float* rv,
                    uint32_t face_count, uint32_t gridcount) {
#pragma omp parallel for schedule(dynamic, 1024) num_threads(nthreads())
  for (uint32_t g = 0; g < gridcount; ++g) {
    /* which face's grid does this point belong to (:58-68) */
    int indice = (int)face_count - 1;
    int toffset = (int)bbs[indice].offset;
    float temd = 20;
    while ((size_t)toffset > (size_t)g && indice > 0) {
      indice--;
      toffset = (int)bbs[indice].offset;
    }
    const orc_bb* bb = &bbs[indice];
    size_t rel = (size_t)g - (size_t)toffset;
    int x = (int)((rel % (bb->size_x * bb->size_z)) % bb->size_x);
    int z = (int)((rel % (bb->size_x * bb->size_z)) / bb->size_x);
    int y = (int)(rel / (bb->size_x * bb->size_z));
    float px = x * (bb->maxx - bb->minx) / (float)(bb->size_x - 1) + bb->minx;
    float py = y * (bb->maxy - bb->miny) / (float)(bb->size_y - 1) + bb->miny;
    float pz = z * (bb->maxz - bb->minz) / (float)(bb->size_z - 1) + bb->minz;

    for (int i = 0; i < (int)face_count; i++) {
      if (px <= bbs[i].maxx && px >= bbs[i].minx && py <= bbs[i].maxy && py >= bbs[i].miny &&
          pz <= bbs[i].maxz && pz >= bbs[i].minz) {
        const float* T = tr + i * 12;
        const float* R = rv + 4 * i;
        float tpx = px + T[3];
        float tpy = py + T[7];
        float tpz = pz + T[11];
        float rpx = T[0] * tpx + T[1] * tpy + T[2] * tpz;
        float rpy = T[4] * tpx + T[5] * tpy + T[6] * tpz;
        float rpz = T[8] * tpx + T[9] * tpy + T[10] * tpz;
        float v1y = R[1], v2x = R[2], v2y = R[3];
        float a = det2(rpy, rpz, 0, v1y) / det2(v2x, v2y, 0, v1y);
        float b = -det2(rpy, rpz, v2x, v2y) / det2(v2x, v2y, 0, v1y);
        float d, td;
        if (a > 0 && b > 0 && a + b < 1)
          d = fabsf(rpx);
        else {
          d = dist_point_segment(rpx, rpy, rpz, 0, 0, R[0], R[1]);
          td = dist_point_segment(rpx, rpy, rpz, R[0], R[1], R[2], R[3]);
          if (td < d) d = td;
          td = dist_point_segment(rpx, rpy, rpz, 0, 0, R[2], R[3]);
          if (td < d) d = td;
        }
        if (d < fabsf(temd)) temd = copysignf(d, rpx);
      }
    }
    df[g] = temd;
  }
}

/* ---- sph_simulation.cpp:200-263 ----------------------------------------------------- */
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
  uint32_t* own = NULL;
  uint32_t* table = cell_table_out;
  if (!table) {
    own = (uint32_t*)malloc(sizeof(uint32_t) * (params->grid_cell_count ? params->grid_cell_count : 1));
    table = own;
  } else if (cell_table_cap < params->grid_cell_count) {
    return NAN;
  }
  orc_sort(back, front, table, params, perm_out, 0);
  double t3 = now();
  orc_density(back, front, params, terms, table);
  double t4 = now();
  orc_forces(front, back, params, terms, table);
  double t5 = now();
  int attempts = 0;
  float computed = dt;
  do { /* :251-260 */
    dt = computed;
    orc_advect(back, front, params->particles_count, params->restitution, dt, df, bbs, face_count);
    ++attempts;
    computed = orc_compute_dt(front, params);
  } while (dt - computed > 0.00001);
  double t6 = now();
  free(own);
  g_times[0] = t1 - t0; g_times[1] = t2 - t1; g_times[2] = t3 - t2;
  g_times[3] = t4 - t3; g_times[4] = t5 - t4; g_times[5] = t6 - t5;
  if (attempts_out) *attempts_out = attempts;
  return computed;
}

/* ---- sph_simulation.cpp:519-531 ------------------------------------------------------ *
 * float/double mix as in the C++ source: pow(float,float) is the float overload there
 * (using namespace std + <cmath>), M_PI is double, results narrow to float on assignment. */
void orc_derive(orc_params* p, orc_terms* t, int n_infl, float* initial_volume_out) {
  p->total_mass = p->particles_count * p->particle_mass;
  float initial_volume = p->total_mass / p->fluid_density;
  p->h = cbrtf((float)(3.f * (n_infl * (initial_volume / p->particles_count)) / (4.f * M_PI)));
  t->poly_6 = (float)(315.f / (64.f * M_PI * powf(p->h, 9.f)));
  t->poly_6_gradient = (float)(-945.f / (32.f * M_PI * powf(p->h, 9.f)));
  t->poly_6_laplacian = (float)(-945.f / (32.f * M_PI * powf(p->h, 9.f)));
  t->spiky = (float)(-45.f / (M_PI * powf(p->h, 6.f)));
  t->viscosity = (float)(45.f / (M_PI * powf(p->h, 6.f)));
  if (initial_volume_out) *initial_volume_out = initial_volume;
}

/* ---- sph_simulation.cpp:52-98 (no last_frame.bin) ----------------------------------- */
void orc_init_particles(orc_particle* buffer, const orc_params* p, float initial_volume) {
  int side_count = (int)ceil(cbrtf((float)p->particles_count));
  float side_length = cbrtf(initial_volume);
  float spacing = side_length / (float)side_count;
  memset(buffer, 0, sizeof(orc_particle) * (size_t)p->particles_count);
  for (unsigned int i = 0; i < p->particles_count; ++i) {
    buffer[i].position.s[0] = (float)(i % side_count) * spacing - side_length / 2.f;
    buffer[i].position.s[1] = ((float)((i / side_count) % side_count) * spacing);
    buffer[i].position.s[2] = (float)(i / (side_count * side_count)) * spacing - side_length / 2.f;
  }
}

void orc_last_step_times(double out[6]) {
  for (int i = 0; i < 6; ++i) out[i] = g_times[i];
}
