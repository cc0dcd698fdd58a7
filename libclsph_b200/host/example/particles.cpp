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
  // the parameter summary, line for line what the reference's CLI prints: labels padded to 27
  // columns, a blank line between the groups
  struct summary {
    static void text(const char* label, const std::string& value) {
      std::string cell(label);
      cell.resize(27, ' ');
      std::cout << cell << value << std::endl;
    }
    static void number(const char* label, double value) {
      std::string cell(label);
      cell.resize(27, ' ');
      std::cout << cell << (float)value << std::endl;
    }
    static void count(const char* label, unsigned long value) {
      std::string cell(label);
      cell.resize(27, ' ');
      std::cout << cell << value << std::endl;
    }
    static void gap() { std::cout << std::endl; }
  };
  summary::gap();
  summary::text("Loaded parameters", "");
  summary::text("-----------------", "");
  summary::number("Simulation time:", p.simulation_time);
  summary::number("Target FPS:", p.target_fps);
  summary::number("Simulation scale:", p.simulation_scale);
  summary::text("Write intermediate frames:", simulation.write_intermediate_frames ? "true" : "false");
  summary::text("Serialize frames:", simulation.serialize ? "true" : "false");
  summary::gap();
  summary::count("Particle count:", p.particles_count);
  summary::number("Particle mass:", p.particle_mass);
  summary::number("Total mass:", p.total_mass);
  summary::number("Initial volume:", simulation.initial_volume);
  summary::gap();
  summary::number("Fluid density:", p.fluid_density);
  summary::number("Dynamic viscosity:", p.dynamic_viscosity);
  summary::number("Surface tension threshold:", p.surface_tension_threshold);
  summary::number("Surface tension:", p.surface_tension);
  summary::number("Stiffness (k):", p.K);
  summary::number("Restitution:", p.restitution);
  summary::gap();
  summary::number("Kernel support radius (h):", p.h);
  summary::gap();
  summary::text("Saving to folder:", saver.frames_folder_prefix + "frames/");

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
