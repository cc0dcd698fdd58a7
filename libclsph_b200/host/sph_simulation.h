// class sph_simulation — the public surface of libclsph/sph_simulation.h:8-27, unchanged, over
// the sphb CUDA step engine (include/sphb.h).  Everything the reference keeps in its private
// half (cl::Context, cl::CommandQueue, 16 cl::Kernel handles, 2 cl::Buffer members,
// sph_simulation.h:29-72) is replaced by one opaque engine handle.
#ifndef _SPH_SIMULATION_H_
#define _SPH_SIMULATION_H_

#include <functional>
#include <string>

#include "common/structures.h"
#include "scene.h"

struct sphb_engine;
namespace sphb_host { struct slab_team; }

class sph_simulation {
 public:
  sph_simulation()
      : parameters(), precomputed_terms(), write_intermediate_frames(false), serialize(false),
        initial_volume(0.f), device(-1), quiet(false), max_steps(0), steps_done(0), device_loop(true),
        devices(1), rebalance(1.25f), slab_recuts(0), force_slabs(false), engine_(0), team_(0) {}

  // Runs the whole simulation: frames of 1/target_fps seconds, each made of adaptive
  // sub-steps, calling the hooks below exactly where the reference does
  // (libclsph/sph_simulation.cpp:265-432).
  void simulate();

  simulation_parameters parameters;
  precomputed_kernel_values precomputed_terms;

  // (host particles in the device's current cell-sorted order, parameters of that step,
  //  isFullFrame); returning true uploads the host array back to the device.
  std::function<bool(particle*, const simulation_parameters&, bool)> pre_frame;
  std::function<void(particle*, const simulation_parameters&)> save_frame;
  std::function<bool(particle*, const simulation_parameters&, bool)> post_frame;

  // Reads the two property files and derives total_mass, initial_volume, h and the five
  // kernel constants (libclsph/sph_simulation.cpp:434-532).
  void load_settings(std::string fluid_file_name, std::string parameters_file_name);

  bool write_intermediate_frames;
  bool serialize;
  float initial_volume;
  scene current_scene;

  // ---- additions (not in the reference; defaults keep its behaviour) -------------------
  int device;               // CUDA ordinal, -1 = current device
  bool quiet;               // suppress the per-sub-step progress lines
  unsigned long max_steps;  // stop after this many sub-steps (0 = run simulation_time)
  unsigned long steps_done; // sub-steps executed by the last simulate()
  bool device_loop;         // a frame's sub-steps run as one engine call (sphb_run_frame)
  // Several GPUs of one box (sph_simulation_slabs.cpp): the particles are partitioned into
  // `devices` slabs along x, GPU device + r stepping slab r, with a halo exchange over NCCL in
  // every step.  Hooks and save_frame see all particles, ordered by grid_index.
  int devices;              // 1 = the reference's single device
  float rebalance;          // re-cut the slabs when one holds more than this multiple of its fair share (0 = never)
  unsigned long slab_recuts;  // re-cuts done by the last simulate()
  bool force_slabs;         // route even devices == 1 through the slab path (one slab: tests)

 private:
  void init_particles(particle* buffer, const simulation_parameters&);
  float simulate_single_frame(float dt);
  void run_frame_on_device(float& dt, float& timeleft);
  bool run_hook(std::function<bool(particle*, const simulation_parameters&, bool)>& hook,
                particle* particles, bool readParticle, bool isFullFrame);
  void read_particles(particle* particles);
  void upload_particles(const particle* particles);
  void check(int rc, const char* file, int line);
  void slabs_begin();
  void slabs_end();
  void slabs_upload(const particle* particles, bool recut);
  void slabs_read(particle* particles);
  float slabs_step(float dt);
  void slabs_fail(int rc, int rank, const char* what);

  sphb_engine* engine_;
  sphb_host::slab_team* team_;
};

#endif
