// Command-line driver with the interface of the reference's example/particles.cpp:
//   ./sph <fluid_name> <simulation_properties_name> <scene_name> <frames_folder_prefix> [-y] [-g N]
// reads fluid_properties/<fluid>.json, simulation_properties/<sim>.json, scenes/<scene>,
// writes <prefix>frames/frameNNNNNNN.geo and (with "serialize") last_frame.bin, asks for a key
// before starting (skipped with -y) and prints the CPU-clock duration of simulate().
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <fstream>
#include <iostream>
#include <string>

#include "file_save_delegates/houdini_file_saver.h"
#include "sph_simulation.h"

static bool resume_file_matches(size_t expected_bytes, bool& present) {
  std::ifstream in("last_frame.bin", std::ios::in | std::ios::binary);
  present = in.is_open();
  if (!present) return true;
  in.seekg(0, std::ios_base::end);
  return (size_t)in.tellg() == expected_bytes;
}

int main(int argc, char** argv) {
  if (argc < 5) {
    std::cout << "Too few arguments" << std::endl
              << "Usage: ./sph <fluid_name> <simulation_properties_name> <scene_name> "
                 "<frames_folder_prefix>" << std::endl;
    return -1;
  }
  // additions after the reference's four arguments: -y (do not wait for a key), -g N (N GPUs)
  bool assume_yes = false;
  int gpus = 1;
  for (int a = 5; a < argc; ++a) {
    if (std::strcmp(argv[a], "-y") == 0) assume_yes = true;
    else if (std::strcmp(argv[a], "-g") == 0 && a + 1 < argc) gpus = std::atoi(argv[++a]);
  }

  sph_simulation simulation;
  simulation.devices = gpus < 1 ? 1 : gpus;
  houdini_file_saver saver = houdini_file_saver(std::string(argv[4]));
  try {
    simulation.load_settings(std::string("fluid_properties/") + argv[1] + ".json",
                             std::string("simulation_properties/") + argv[2] + ".json");
  } catch (const std::exception& ex) {
    std::cerr << ex.what() << std::endl;
    std::exit(-1);
  }

  simulation.save_frame = [&](particle* particles, const simulation_parameters& params) {
    saver.writeFrameToFile(particles, params);
    if (simulation.serialize) {  // raw dump, no header: what init_particles reads back
      std::ofstream out("last_frame.bin", std::ios::out | std::ios::trunc | std::ios::binary);
      out.write(reinterpret_cast<const char*>(particles),
                (std::streamsize)(sizeof(particle) * (size_t)params.particles_count));
    }
  };

  const simulation_parameters& p = simulation.parameters;
  std::cout << std::endl
            << "Loaded parameters          " << std::endl
            << "-----------------          " << std::endl
            << "Simulation time:           " << p.simulation_time << std::endl
            << "Target FPS:                " << p.target_fps << std::endl
            << "Simulation scale:          " << p.simulation_scale << std::endl
            << "Write intermediate frames: " << (simulation.write_intermediate_frames ? "true" : "false") << std::endl
            << "Serialize frames:          " << (simulation.serialize ? "true" : "false") << std::endl
            << std::endl
            << "Particle count:            " << p.particles_count << std::endl
            << "Particle mass:             " << p.particle_mass << std::endl
            << "Total mass:                " << p.total_mass << std::endl
            << "Initial volume:            " << simulation.initial_volume << std::endl
            << std::endl
            << "Fluid density:             " << p.fluid_density << std::endl
            << "Dynamic viscosity:         " << p.dynamic_viscosity << std::endl
            << "Surface tension threshold: " << p.surface_tension_threshold << std::endl
            << "Surface tension:           " << p.surface_tension << std::endl
            << "Stiffness (k):             " << p.K << std::endl
            << "Restitution:               " << p.restitution << std::endl
            << std::endl
            << "Kernel support radius (h): " << p.h << std::endl
            << std::endl
            << "Saving to folder:          " << saver.frames_folder_prefix + "frames/" << std::endl;

  if (!simulation.current_scene.load(argv[3], p.h * 2)) {
    std::cerr << "Unable to load scene: " << argv[3] << std::endl;
    return -1;
  }

  bool present = false;
  if (!resume_file_matches(sizeof(particle) * (size_t)p.particles_count, present)) {
    std::cout << std::endl
              << "\033[1;31m Serialized frame of incorrect size found. Revert to last know settings "
                 "or delete it, then try again. \033[0m" << std::endl;
    return 0;
  }
  if (present) {
    std::cout << std::endl
              << "\033[1;32m Serialized frame found.  Simulation will pick up where last run left off.\033[0m";
    std::cout << std::endl
              << "\033[1;32m To start a new simulation, delete last_frame.bin. \033[0m" << std::endl;
  }

  char response = 'y';
  if (!assume_yes) {
    std::cout << std::endl
              << "Revise simulation parameters.  Press q to quit, any other key to proceed with simulation"
              << std::endl;
    std::cin >> response;
  }
  double duration = 0;
  const std::clock_t start = std::clock();
  if (response != 'q') {
    simulation.simulate();
    duration = (std::clock() - start) / (double)CLOCKS_PER_SEC;
  }
  std::cout << "Duration : " << duration << '\n';
  return 0;
}
