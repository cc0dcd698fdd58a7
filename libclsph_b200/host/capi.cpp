// C entry points over the C++11 host classes, for callers without a C++ toolchain (the Python
// tests and bench.py bind these with ctypes).  Declared in include/clsph_host.h.
#include "clsph_host.h"

#include <cstring>
#include <exception>
#include <iostream>
#include <new>
#include <sstream>
#include <string>
#include <vector>

#include "file_save_delegates/houdini_file_saver.h"
#include "scene.h"
#include "sph_simulation.h"

namespace sphb_host {
void lattice_particles(particle* buffer, unsigned int count, float initial_volume);
}

namespace {
void set_err(char* err, size_t cap, const std::string& text) {
  if (!err || cap == 0) return;
  std::strncpy(err, text.c_str(), cap - 1);
  err[cap - 1] = '\0';
}

// silences std::cout for the duration of a call (the host classes print progress lines)
struct mute_cout {
  std::streambuf* saved;
  std::ostringstream sink;
  explicit mute_cout(bool on) : saved(0) {
    if (on) saved = std::cout.rdbuf(sink.rdbuf());
  }
  ~mute_cout() {
    if (saved) std::cout.rdbuf(saved);
  }
};
}  // namespace

namespace sphb_host {
std::vector<float> equal_count_cuts(const particle* p, size_t n, int world);  // sph_simulation_slabs.cpp
int gather_sort_selftest(const particle* stage, const uint32_t* stage_ids, const uint32_t* start, int world,
                         uint32_t cells, particle* particles, uint32_t* ids_out);
int deal_out_selftest(const particle* particles, const uint32_t* ids, uint32_t n, const float* cuts, int world,
                      int threads, particle* stage, uint32_t* stage_ids, uint32_t* start_out);
}

extern "C" {

clsph_scene* clsph_scene_load(const char* obj_path, float dist_field_threshold) {
  if (!obj_path) return 0;
  mute_cout quiet(true);
  scene* s = new (std::nothrow) scene();
  if (!s) return 0;
  if (!s->load_path(obj_path, dist_field_threshold)) {
    delete s;
    return 0;
  }
  return reinterpret_cast<clsph_scene*>(s);
}

uint32_t clsph_scene_face_count(const clsph_scene* s) {
  return s ? reinterpret_cast<const scene*>(s)->face_count : 0;
}

uint32_t clsph_scene_total_grid_points(const clsph_scene* s) {
  return s ? reinterpret_cast<const scene*>(s)->totalGridpoints : 0;
}

void clsph_scene_copy(const clsph_scene* h, sphb_bb* bbs, float* transforms, float* rvertices,
                      float* face_normals) {
  if (!h) return;
  const scene* s = reinterpret_cast<const scene*>(h);
  if (bbs) std::memcpy(bbs, s->bbs.data(), sizeof(BB) * s->bbs.size());
  if (transforms) std::memcpy(transforms, s->transforms.data(), sizeof(float) * s->transforms.size());
  if (rvertices) std::memcpy(rvertices, s->rvertices.data(), sizeof(float) * s->rvertices.size());
  if (face_normals) std::memcpy(face_normals, s->face_normals.data(), sizeof(float) * s->face_normals.size());
}

void clsph_scene_free(clsph_scene* s) { delete reinterpret_cast<scene*>(s); }

int clsph_load_settings(const char* fluid_json, const char* sim_json, sphb_params* params,
                        sphb_terms* terms, float* initial_volume, int* write_all_frames,
                        int* serialize, char* err, size_t err_cap) {
  if (!fluid_json || !sim_json || !params || !terms) return -1;
  try {
    mute_cout quiet(true);
    sph_simulation sim;
    sim.load_settings(fluid_json, sim_json);
    std::memcpy(params, &sim.parameters, sizeof(sphb_params));
    std::memcpy(terms, &sim.precomputed_terms, sizeof(sphb_terms));
    if (initial_volume) *initial_volume = sim.initial_volume;
    if (write_all_frames) *write_all_frames = sim.write_intermediate_frames ? 1 : 0;
    if (serialize) *serialize = sim.serialize ? 1 : 0;
  } catch (const std::exception& ex) {
    set_err(err, err_cap, ex.what());
    return -1;
  }
  return 0;
}

void clsph_lattice(sphb_particle* out, uint32_t count, float initial_volume) {
  if (!out) return;
  std::memset(out, 0, sizeof(sphb_particle) * (size_t)count);
  sphb_host::lattice_particles(reinterpret_cast<particle*>(out), count, initial_volume);
}

size_t clsph_geo_text(const sphb_particle* particles, uint32_t count, float particle_mass,
                      uint32_t threads, char* out, size_t cap) {
  const std::string text =
      sphb_host::houdini_geo_text(reinterpret_cast<const particle*>(particles), count, particle_mass, threads);
  if (out && cap >= text.size()) std::memcpy(out, text.data(), text.size());
  return text.size();
}

size_t clsph_format_floats(const float* values, size_t count, char* out, size_t cap) {
  if (!values) return 0;
  const std::string text = sphb_host::format_floats_like_ostream(values, count);
  if (out && cap >= text.size()) std::memcpy(out, text.data(), text.size());
  return text.size();
}

size_t clsph_bgeo_bytes(const sphb_particle* particles, const sphb_params* params, uint32_t threads,
                        char* out, size_t cap) {
  if (!particles || !params) return 0;
  simulation_parameters p;
  std::memcpy(&p, params, sizeof p);
  const std::string bytes =
      sphb_host::houdini_bgeo_bytes(reinterpret_cast<const particle*>(particles), p, threads);
  if (out && cap >= bytes.size()) std::memcpy(out, bytes.data(), bytes.size());
  return bytes.size();
}

int clsph_save_frame(const char* frames_prefix, int use_partio, sphb_particle* particles,
                     const sphb_params* params) {
  if (!frames_prefix || !particles || !params) return -1;
  simulation_parameters p;
  std::memcpy(&p, params, sizeof p);
  houdini_file_saver saver(frames_prefix);
  saver.use_partio = use_partio != 0;
  return saver.writeFrameToFile(reinterpret_cast<particle*>(particles), p);
}

int clsph_equal_count_cuts(const sphb_particle* particles, uint32_t n, int world, float* cuts_out) {
  if (!particles || world < 1 || !cuts_out) return -1;
  const std::vector<float> cuts = sphb_host::equal_count_cuts(reinterpret_cast<const particle*>(particles), n, world);
  for (size_t k = 0; k < cuts.size(); ++k) cuts_out[k] = cuts[k];
  return (int)cuts.size();
}

int clsph_gather_sort(const sphb_particle* stage, const uint32_t* stage_ids, const uint32_t* start, int world,
                      uint32_t cells, sphb_particle* particles_out, uint32_t* ids_out) {
  if (!stage || !stage_ids || !start || world < 1 || !particles_out || !ids_out) return -1;
  return sphb_host::gather_sort_selftest(reinterpret_cast<const particle*>(stage), stage_ids, start, world, cells,
                                         reinterpret_cast<particle*>(particles_out), ids_out);
}

int clsph_deal_out(const sphb_particle* particles, const uint32_t* ids, uint32_t n, const float* cuts, int world,
                   int threads, sphb_particle* stage_out, uint32_t* stage_ids_out, uint32_t* start_out) {
  if (!particles || !ids || !cuts || !stage_out || !stage_ids_out || !start_out) return -1;
  return sphb_host::deal_out_selftest(reinterpret_cast<const particle*>(particles), ids, n, cuts, world, threads,
                                      reinterpret_cast<particle*>(stage_out), stage_ids_out, start_out);
}

int clsph_simulate(const char* fluid_json, const char* sim_json, const char* obj_path, int device,
                   uint64_t max_steps, const char* frames_prefix, sphb_particle* final_state,
                   sphb_params* final_params, uint64_t* steps_done, uint32_t* frames_saved,
                   char* err, size_t err_cap) {
  return clsph_simulate_devices(fluid_json, sim_json, obj_path, device, 1, 0.f, max_steps, frames_prefix,
                                final_state, final_params, steps_done, frames_saved, 0, err, err_cap);
}

int clsph_simulate_devices(const char* fluid_json, const char* sim_json, const char* obj_path, int device,
                           int devices, float rebalance, uint64_t max_steps, const char* frames_prefix,
                           sphb_particle* final_state, sphb_params* final_params, uint64_t* steps_done,
                           uint32_t* frames_saved, uint64_t* recuts, char* err, size_t err_cap) {
  if (!fluid_json || !sim_json || !obj_path || devices == 0) return -1;
  try {
    mute_cout quiet(true);
    sph_simulation sim;
    sim.load_settings(fluid_json, sim_json);
    if (!sim.current_scene.load_path(obj_path, sim.parameters.h * 2)) {
      set_err(err, err_cap, std::string("Unable to load scene: ") + obj_path);
      return -1;
    }
    sim.device = device;
    sim.devices = devices < 0 ? -devices : devices;
    sim.force_slabs = devices < 0;
    if (devices != 1) sim.rebalance = rebalance;
    sim.quiet = true;
    sim.max_steps = (unsigned long)max_steps;
    houdini_file_saver saver(frames_prefix ? frames_prefix : "");
    uint32_t saved = 0;
    const uint32_t n = sim.parameters.particles_count;
    sim.save_frame = [&](particle* ps, const simulation_parameters& p) {
      if (frames_prefix) saver.writeFrameToFile(ps, p);
      if (final_state) std::memcpy(final_state, ps, sizeof(particle) * (size_t)n);
      if (final_params) std::memcpy(final_params, &p, sizeof(sphb_params));
      ++saved;
    };
    sim.simulate();
    if (steps_done) *steps_done = sim.steps_done;
    if (frames_saved) *frames_saved = saved;
    if (recuts) *recuts = sim.slab_recuts;
  } catch (const std::exception& ex) {
    set_err(err, err_cap, ex.what());
    return -1;
  }
  return 0;
}

}  // extern "C"
