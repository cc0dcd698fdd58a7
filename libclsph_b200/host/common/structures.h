// The four plain structs of libclsph's public surface (libclsph/common/structures.h:16-65),
// with the byte layout an OpenCL device gives them: simulation_parameters 128 B, particle 80 B,
// precomputed_kernel_values 20 B, BB 56 B.  include/sphb.h declares the same layouts under
// sphb_* names for the C boundary; the static_asserts below keep the two in step.
#ifndef _STRUCTURES_H_
#define _STRUCTURES_H_

#include <stddef.h>

#include "util/cl_types.h"

#define COLLISION_VOLUMES_COUNT 3

typedef struct {
  cl_uint particles_count;
  cl_float fluid_density, total_mass, particle_mass, dynamic_viscosity;
  cl_float simulation_time;
  cl_float h;  // kernel support radius
  cl_float simulation_scale, target_fps;
  cl_float surface_tension_threshold, surface_tension;
  cl_float restitution;
  cl_float K;  // Tait stiffness

  cl_float3 constant_acceleration;

  // rewritten by every step (sph_simulation.cpp:670-727 in the reference)
  cl_int grid_size_x, grid_size_y, grid_size_z;
  cl_uint grid_cell_count;
  cl_float3 min_point, max_point;
} simulation_parameters;

typedef struct {
  cl_float3 position, velocity, intermediate_velocity, acceleration;
  cl_float density, pressure;
  cl_uint grid_index;
} particle;

typedef struct {
  float poly_6, poly_6_gradient, poly_6_laplacian, spiky, viscosity;
} precomputed_kernel_values;

// Inflated bounding box of one triangle + the shape of its distance-field grid.
typedef struct {
  float maxx, maxy, maxz;
  float minx, miny, minz;
  size_t size_x, size_y, size_z;
  size_t offset;  // first grid point of this face in the flat distance field
} BB;

#ifdef __cplusplus
static_assert(sizeof(simulation_parameters) == 128, "simulation_parameters layout");
static_assert(offsetof(simulation_parameters, constant_acceleration) == 64, "layout");
static_assert(offsetof(simulation_parameters, grid_size_x) == 80, "layout");
static_assert(offsetof(simulation_parameters, min_point) == 96, "layout");
static_assert(sizeof(particle) == 80, "particle layout");
static_assert(offsetof(particle, density) == 64 && offsetof(particle, grid_index) == 72, "layout");
static_assert(sizeof(precomputed_kernel_values) == 20, "precomputed_kernel_values layout");
static_assert(sizeof(BB) == 56, "BB layout");
#endif

#endif
