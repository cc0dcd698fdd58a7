// Host frame loop of the simulation.  Restates the observable behaviour of
// libclsph/sph_simulation.cpp: load_settings (:434-532), init_particles (:52-98), the frame /
// sub-step loop with its hooks and save thread (:265-432, :730-768).  The step itself
// (simulate_single_frame, :200-263, and everything under it) is one call into the CUDA engine.
#define _USE_MATH_DEFINES
#include "sph_simulation.h"

#include <cmath>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <thread>
#include <vector>

#include "sphb.h"
#include "util/json_min.h"

static_assert(sizeof(sphb_particle) == sizeof(particle), "particle layout");
static_assert(sizeof(sphb_params) == sizeof(simulation_parameters), "simulation_parameters layout");
static_assert(sizeof(sphb_terms) == sizeof(precomputed_kernel_values), "terms layout");
static_assert(sizeof(sphb_bb) == sizeof(BB), "BB layout");

namespace {
const unsigned int kPreferredGroupMultiple = 64;

inline float number_of(const sphb_host::json& obj, const char* key) {
  return (float)obj[key].number();
}
}  // namespace

// Any engine error is fatal, like the reference's EXIT_ON_CL_ERROR (util/cl_boilerplate.h:32-34).
void sph_simulation::check(int rc, const char* file, int line) {
  if (rc == 0) return;
  std::cerr << "A CUDA engine error occured (" << file << ":" << line << ")-> "
            << (engine_ ? sphb_last_error(engine_) : sphb_error_string(rc)) << std::endl;
  std::exit(-1);
}
#define SPHB_CHECK(call) check((call), __FILE__, __LINE__)

void sph_simulation::load_settings(std::string fluid_file_name, std::string parameters_file_name) {
  int particles_inside_influence_radius = 0;
  {
    std::ifstream in(fluid_file_name.c_str());
    if (!in) throw std::runtime_error("Cannot open " + fluid_file_name);
    const sphb_host::json fluid = sphb_host::json::parse(in);
    parameters.fluid_density = number_of(fluid, "fluid_density");
    parameters.dynamic_viscosity = number_of(fluid, "dynamic_viscosity");
    parameters.restitution = number_of(fluid, "restitution");
    if (parameters.restitution < 0 || parameters.restitution > 1)
      throw std::runtime_error("Restitution has an invalid value!");
    parameters.K = number_of(fluid, "k");
    parameters.surface_tension_threshold = number_of(fluid, "surface_tension_threshold");
    parameters.surface_tension = number_of(fluid, "surface_tension");
    particles_inside_influence_radius = (int)fluid["particles_inside_influence_radius"].number();
  }
  {
    std::ifstream in(parameters_file_name.c_str());
    if (!in) throw std::runtime_error("Cannot open " + parameters_file_name);
    const sphb_host::json sim = sphb_host::json::parse(in);
    parameters.particles_count = (unsigned int)sim["particles_count"].number();
    if (parameters.particles_count % kPreferredGroupMultiple != 0) {
      std::cout << std::endl
                << "\033[1;31m You should choose a number of particles that is "
                   "divisble by the preferred work group size.\033[0m";
      std::cout << std::endl << "\033[1;31m Performances will be sub-optimal.\033[0m" << std::endl;
    }
    parameters.particle_mass = number_of(sim, "particle_mass");
    parameters.simulation_time = number_of(sim, "simulation_time");
    parameters.target_fps = number_of(sim, "target_fps");
    parameters.simulation_scale = number_of(sim, "simulation_scale");
    const sphb_host::json& g = sim["constant_acceleration"];
    parameters.constant_acceleration.s[0] = number_of(g, "x");
    parameters.constant_acceleration.s[1] = number_of(g, "y");
    parameters.constant_acceleration.s[2] = number_of(g, "z");
    write_intermediate_frames = sim["write_all_frames"].boolean();
    serialize = sim["serialize"].boolean();
  }

  // Derived quantities: the float/double mix of each expression is the reference's
  // (sph_simulation.cpp:519-531: pow(float, float) is the FLOAT overload there — the file says
  // `using namespace std` — and only M_PI widens the products to double).
  parameters.total_mass = parameters.particles_count * parameters.particle_mass;
  initial_volume = parameters.total_mass / parameters.fluid_density;
  parameters.h = cbrtf(3.f * (particles_inside_influence_radius * (initial_volume / parameters.particles_count)) /
                       (4.f * M_PI));
  const float h9 = std::pow(parameters.h, 9.f), h6 = std::pow(parameters.h, 6.f);
  precomputed_terms.poly_6 = 315.f / (64.f * M_PI * h9);
  precomputed_terms.poly_6_gradient = -945.f / (32.f * M_PI * h9);
  precomputed_terms.poly_6_laplacian = -945.f / (32.f * M_PI * h9);
  precomputed_terms.spiky = -45.f / (M_PI * h6);
  precomputed_terms.viscosity = 45.f / (M_PI * h6);
}

namespace sphb_host {
// The default initial state: a cube lattice of ceil(cbrt(N)) particles per side, centred in
// x and z, resting on y = 0, at rest (sph_simulation.cpp:75-96).
void lattice_particles(particle* buffer, unsigned int count, float initial_volume) {
  const int per_side = ceil(cbrtf(count));
  const float side_length = cbrtf(initial_volume);
  const float spacing = side_length / (float)per_side;
  for (unsigned int i = 0; i < count; ++i) {
    particle& q = buffer[i];
    q.position.s[0] = (float)(i % per_side) * spacing - side_length / 2.f;
    q.position.s[1] = ((float)((i / per_side) % per_side) * spacing);
    q.position.s[2] = (float)(i / (per_side * per_side)) * spacing - side_length / 2.f;
    for (int a = 0; a < 3; ++a) q.velocity.s[a] = q.intermediate_velocity.s[a] = 0.f;
    q.density = 0.f;
    q.pressure = 0.f;
  }
}
}  // namespace sphb_host

// A resumed state (raw last_frame.bin) if one can be opened, else the cube lattice
// (sph_simulation.cpp:52-98).
void sph_simulation::init_particles(particle* buffer, const simulation_parameters& p) {
  const int per_side = ceil(cbrtf(p.particles_count));
  const float side_length = cbrtf(initial_volume);
  const float spacing = side_length / (float)per_side;
  std::cout << "volume: " << initial_volume << " side_length: " << side_length
            << " spacing: " << spacing << std::endl;

  std::ifstream resume("last_frame.bin", std::ios::in | std::ios::binary);
  if (resume.is_open()) {
    resume.read(reinterpret_cast<char*>(buffer), sizeof(particle) * (size_t)p.particles_count);
    return;
  }
  sphb_host::lattice_particles(buffer, p.particles_count, initial_volume);
}

float sph_simulation::simulate_single_frame(float dt) {
  if (team_) return slabs_step(dt);
  float computed = dt;
  SPHB_CHECK(sphb_step(engine_, dt, &computed));
  // the reference mutates `parameters` in findMinMaxPosition (:670-727); hooks see that
  sphb_params now;
  SPHB_CHECK(sphb_get_parameters(engine_, &now));
  parameters.grid_size_x = now.grid_size_x;
  parameters.grid_size_y = now.grid_size_y;
  parameters.grid_size_z = now.grid_size_z;
  parameters.grid_cell_count = now.grid_cell_count;
  for (int a = 0; a < 4; ++a) {
    parameters.min_point.s[a] = now.min_point.s[a];
    parameters.max_point.s[a] = now.max_point.s[a];
  }
  ++steps_done;
  return computed;
}

// The sub-step loop of one frame (libclsph/sph_simulation.cpp:384-409) as one engine call.
void sph_simulation::run_frame_on_device(float& dt, float& timeleft) {
  const unsigned int log_cap = 4096;
  std::vector<float> log(log_cap);
  unsigned int steps = 0;
  const unsigned long left = max_steps ? (max_steps > steps_done ? max_steps - steps_done : 1) : 0;
  SPHB_CHECK(sphb_run_frame(engine_, dt, timeleft, left, &dt, &steps, log.data(), log_cap));
  sphb_params now;
  SPHB_CHECK(sphb_get_parameters(engine_, &now));
  parameters.grid_size_x = now.grid_size_x;
  parameters.grid_size_y = now.grid_size_y;
  parameters.grid_size_z = now.grid_size_z;
  parameters.grid_cell_count = now.grid_cell_count;
  for (int a = 0; a < 4; ++a) {
    parameters.min_point.s[a] = now.min_point.s[a];
    parameters.max_point.s[a] = now.max_point.s[a];
  }
  steps_done += steps;
  if (!quiet)
    for (unsigned int k = 0; k < steps && k < log_cap; ++k)
      std::cout << "temps restant pour la frame :" << log[k] << std::endl;
  if (steps) timeleft = log[(steps < log_cap ? steps : log_cap) - 1];
}

void sph_simulation::read_particles(particle* particles) {
  if (team_) return slabs_read(particles);
  SPHB_CHECK(sphb_download(engine_, reinterpret_cast<sphb_particle*>(particles)));
}

void sph_simulation::upload_particles(const particle* particles) {
  if (team_) return slabs_upload(particles, false);
  SPHB_CHECK(sphb_upload(engine_, reinterpret_cast<const sphb_particle*>(particles),
                         parameters.particles_count));
}

// executePreFrameOpperation / executePostFrameOpperation (:730-768)
bool sph_simulation::run_hook(std::function<bool(particle*, const simulation_parameters&, bool)>& hook,
                              particle* particles, bool readParticle, bool isFullFrame) {
  if (readParticle) {
    read_particles(particles);
    readParticle = false;
  }
  if (hook(particles, parameters, isFullFrame)) upload_particles(particles);
  return readParticle;
}

void sph_simulation::simulate() {
  std::thread savet;
  bool readParticle = true;
  steps_done = 0;

  const bool slabs = devices > 1 || force_slabs;
  if (devices < 1) devices = 1;
  slab_recuts = 0;
  if (!slabs)
    SPHB_CHECK(sphb_create(reinterpret_cast<const sphb_params*>(&parameters),
                           reinterpret_cast<const sphb_terms*>(&precomputed_terms), device, &engine_));

  particle* particles = 0;
  // pinned, so the per-frame read-back runs at full PCIe speed
  SPHB_CHECK(sphb_host_alloc(reinterpret_cast<void**>(&particles),
                             sizeof(particle) * (size_t)parameters.particles_count));
  for (size_t i = 0; i < parameters.particles_count; ++i) particles[i] = particle();
  init_particles(particles, parameters);
  if (!slabs) upload_particles(particles);

  float time = 0.0;
  const float timeperframe = 1.0f / parameters.target_fps;
  int currentFrame = 2;
  float dt = timeperframe * parameters.simulation_scale;

  // computeDistanceField (:607-632): built on the device from the scene::load products
  std::cout << "totalgpoints " << current_scene.totalGridpoints << std::endl;
  if (slabs) {
    // one slab engine per GPU, each with its own copy of the scene; the particles are dealt out
    // by equal-count faces along x (sph_simulation_slabs.cpp)
    slabs_begin();
    slabs_upload(particles, true);
  } else {
    SPHB_CHECK(sphb_set_scene(engine_, reinterpret_cast<const sphb_bb*>(current_scene.bbs.data()),
                              current_scene.face_count, current_scene.transforms.data(),
                              current_scene.rvertices.data(), current_scene.totalGridpoints));
  }

  // one save in flight at most; the host array is borrowed until the next join
  auto spawn_save = [&]() {
    const simulation_parameters snapshot = parameters;
    savet = std::thread([this, particles, snapshot] { save_frame(particles, snapshot); });
  };
  if (save_frame) spawn_save();

  // Without pre/post hooks nobody but the save thread reads the host array, so the per-frame
  // read-back need not block the step loop: the state is snapshot on the device,
  // the copy to the host runs beside the next frame's steps, and the save thread waits for it.
  const bool async_frames = save_frame && !pre_frame && !post_frame && !slabs;
  auto read_and_save = [&]() {
    savet.join();  // the previous save still reads the host array
    if (!async_frames) {
      read_particles(particles);
      spawn_save();
      return;
    }
    SPHB_CHECK(sphb_download_begin(engine_, reinterpret_cast<sphb_particle*>(particles)));
    const simulation_parameters snapshot = parameters;
    sphb_engine* const eng = engine_;
    savet = std::thread([this, eng, particles, snapshot] {
      const int rc = sphb_download_wait(eng);
      if (rc != 0) {
        std::cerr << "sphb_download_wait: " << sphb_last_error(eng) << std::endl;
        exit(-1);
      }
      save_frame(particles, snapshot);
    });
  };

  bool stop = false;
  while (time < parameters.simulation_time && !stop) {
    std::cout << "Simulating frame " << currentFrame << " (" << time << "s)" << std::endl;

    if (!write_intermediate_frames && pre_frame)
      readParticle = run_hook(pre_frame, particles, readParticle, true);

    float timeleft = timeperframe;
    if (!write_intermediate_frames && device_loop && !slabs) {
      // Nothing looks at the state between the sub-steps of this frame: the whole loop below runs
      // inside the engine, without a host round trip per step, and the lines the
      // reference prints per step are printed afterwards from the engine's log.
      run_frame_on_device(dt, timeleft);
      readParticle = true;
      if (max_steps && steps_done >= max_steps) stop = true;
    }
    while (timeleft > 0.0 && !stop) {
      if (write_intermediate_frames && pre_frame)
        readParticle = run_hook(pre_frame, particles, readParticle, false);
      readParticle = true;
      dt = simulate_single_frame(dt);
      timeleft -= dt;
      if (timeleft < dt) dt = timeleft;
      if (!quiet) std::cout << "temps restant pour la frame :" << timeleft << std::endl;
      if (save_frame && write_intermediate_frames) {
        read_and_save();
        readParticle = false;
      }
      if (write_intermediate_frames && post_frame)
        readParticle = run_hook(post_frame, particles, readParticle, false);
      if (max_steps && steps_done >= max_steps) {
        stop = true;
        break;
      }
    }
    time += timeperframe;
    ++currentFrame;

    if (!write_intermediate_frames && save_frame) {
      read_and_save();
      readParticle = false;
    }
    if (!write_intermediate_frames && post_frame)
      readParticle = run_hook(post_frame, particles, readParticle, true);
  }
  if (save_frame) savet.join();
  sphb_host_free(particles);
  if (slabs) slabs_end();
  else sphb_destroy(engine_);
  engine_ = 0;
}
