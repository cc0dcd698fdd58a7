/* clsph_host — C entry points over the C++11 host classes that keep libclsph's API
 * (libclsph_b200/host/: sph_simulation, scene, houdini_file_saver).  A C++ caller uses those
 * classes directly, exactly as example/particles.cpp uses the reference's; these wrappers
 * exist for callers that cannot (ctypes in tests/ and bench.py).  Built into
 * libclsph_b200/libclsph_host.so, which links libsphb.so (include/sphb.h).
 */
#ifndef CLSPH_HOST_H_
#define CLSPH_HOST_H_

#include <stddef.h>
#include <stdint.h>

#include "sphb.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct clsph_scene clsph_scene;

/* scene::load (libclsph/scene.cpp:9-153) on an explicit OBJ path; NULL on failure. */
clsph_scene* clsph_scene_load(const char* obj_path, float dist_field_threshold);
uint32_t clsph_scene_face_count(const clsph_scene* s);
uint32_t clsph_scene_total_grid_points(const clsph_scene* s);
/* bbs[face_count], transforms[12*face_count], rvertices[4*face_count],
 * face_normals[3*face_count]; each nullable */
void clsph_scene_copy(const clsph_scene* s, sphb_bb* bbs, float* transforms, float* rvertices,
                      float* face_normals);
void clsph_scene_free(clsph_scene* s);

/* sph_simulation::load_settings (libclsph/sph_simulation.cpp:434-532). 0 = ok, -1 = error
 * (message in err). */
int clsph_load_settings(const char* fluid_json, const char* sim_json, sphb_params* params,
                        sphb_terms* terms, float* initial_volume, int* write_all_frames,
                        int* serialize, char* err, size_t err_cap);

/* The cube-lattice initial state of sph_simulation::init_particles (:75-96); every other
 * field zero. */
void clsph_lattice(sphb_particle* out, uint32_t count, float initial_volume);

/* The PGEOMETRY V5 text houdini_file_saver::writeFrameToFile writes
 * (util/houdini_geo/HoudiniFileDumpHelper.cpp:19-90).  Returns the length; copies it to `out`
 * when cap is large enough. */
size_t clsph_geo_text(const sphb_particle* particles, uint32_t count, float particle_mass,
                      uint32_t threads, char* out, size_t cap);

/* The .geo writer's float formatter on its own (tests): each value as the reference's
 * `stream << float` prints it (printf "%g"; util/houdini_geo/HoudiniFileDumpHelper.cpp:56-75),
 * followed by '\n'.  Same return convention as clsph_geo_text. */
size_t clsph_format_floats(const float* values, size_t count, char* out, size_t cap);

/* The Bgeo-V5 bytes of the reference's USE_PARTIO build for the same frame
 * (util/partio/PartioFunctions.h:5-65 + Partio::write, libclsph/file_save_delegates/
 * houdini_file_saver.cpp:78-89; partio itself is not vendored in the reference).  Same return
 * convention as clsph_geo_text. */
size_t clsph_bgeo_bytes(const sphb_particle* particles, const sphb_params* params, uint32_t threads,
                        char* out, size_t cap);

/* One houdini_file_saver::writeFrameToFile call on a fresh saver: writes
 * <frames_prefix>frames/frame0000001.geo, or .bgeo when use_partio is non-zero. */
int clsph_save_frame(const char* frames_prefix, int use_partio, sphb_particle* particles,
                     const sphb_params* params);

/* Headless sph_simulation::simulate() (:265-432) with a save_frame hook installed (frames are
 * written as .geo under <frames_prefix>frames/ when frames_prefix is non-NULL).  Stops after
 * max_steps sub-steps when max_steps > 0.  final_state / final_params receive what the last
 * save_frame call saw. */
int clsph_simulate(const char* fluid_json, const char* sim_json, const char* obj_path, int device,
                   uint64_t max_steps, const char* frames_prefix, sphb_particle* final_state,
                   sphb_params* final_params, uint64_t* steps_done, uint32_t* frames_saved,
                   char* err, size_t err_cap);

/* The slab faces sph_simulation::devices deals the particles out by: world + 1 values, -inf first,
 * +inf last, equal counts between them, a face midway between two distinct x values.  Returns
 * world + 1, or -1.  (No GPU involved.) */
int clsph_equal_count_cuts(const sphb_particle* particles, uint32_t n, int world, float* cuts_out);

/* The frame gather of sph_simulation::devices without its engines: stage[start[r] .. start[r+1])
 * holds slab r's particles (ids beside them); out = all of them stably ordered by grid_index, slabs
 * left to right, keys >= cells in one last bucket; done by `world` host threads.  (No GPU involved.) */
int clsph_gather_sort(const sphb_particle* stage, const uint32_t* stage_ids, const uint32_t* start, int world,
                      uint32_t cells, sphb_particle* particles_out, uint32_t* ids_out);

/* The upload side of the same: particles (ids beside them) dealt out to the slabs
 * [cuts[r], cuts[r+1]) — stage_out[start_out[r] .. start_out[r+1]) = slab r's particles in the order
 * they have in `particles` — by `threads` host threads, each owning one piece of the input.
 * world <= 255.  (No GPU involved.) */
int clsph_deal_out(const sphb_particle* particles, const uint32_t* ids, uint32_t n, const float* cuts, int world,
                   int threads, sphb_particle* stage_out, uint32_t* stage_ids_out, uint32_t* start_out);

/* The same over `devices` GPUs of one box (device, device + 1, ...): sph_simulation::devices —
 * the particles are partitioned into slabs along x, one slab engine and one host thread per GPU,
 * halo exchange over NCCL inside every step (libclsph_b200/host/sph_simulation_slabs.cpp).
 * rebalance: re-cut the slabs when one holds more than this multiple of its fair share (0 =
 * never); *recuts: how often that happened.  final_state is ordered by grid_index.
 * devices < 0: |devices| GPUs with sph_simulation::force_slabs (-1 = one slab engine on one GPU
 * through the same path). */
int clsph_simulate_devices(const char* fluid_json, const char* sim_json, const char* obj_path, int device,
                           int devices, float rebalance, uint64_t max_steps, const char* frames_prefix,
                           sphb_particle* final_state, sphb_params* final_params, uint64_t* steps_done,
                           uint32_t* frames_saved, uint64_t* recuts, char* err, size_t err_cap);

#ifdef __cplusplus
}
#endif
#endif /* CLSPH_HOST_H_ */
