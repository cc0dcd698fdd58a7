// houdini_file_saver — same surface as libclsph/file_save_delegates/houdini_file_saver.h:9-20.
// Writes one ASCII Houdini "PGEOMETRY V5" .geo per call to
// <prefix>frames/frame<7 digits>.geo, byte-identical to what the reference's
// HoudiniFileDumpParticles::dump produces (util/houdini_geo/HoudiniFileDumpHelper.cpp:19-90),
// but formatted in parallel into memory and written with one write per chunk.
//
// The reference's other compile-time branch (USE_PARTIO, houdini_file_saver.cpp:78-89) hands the
// frame to the partio library, which writes a binary Houdini "Bgeo V5" .bgeo.  partio is not
// vendored in the reference (only four API headers under util/partio/), so that branch is served
// here by the repo's own Bgeo-V5 writer (sphb_host::houdini_bgeo_bytes): defining USE_PARTIO when
// this class is compiled, or setting use_partio at run time, selects it.
#ifndef _HOUDINI_FILE_SAVER_H_
#define _HOUDINI_FILE_SAVER_H_

#include <string>

#include "common/structures.h"

class houdini_file_saver {
 public:
  houdini_file_saver(std::string frames_folder_prefix)
      : frames_folder_prefix(frames_folder_prefix),
#ifdef USE_PARTIO
        use_partio(true),
#else
        use_partio(false),
#endif
        frame_count(0) {
  }

  int writeFrameToFile(particle* particles, const simulation_parameters& parameters);

  std::string frames_folder_prefix;
  // addition: true writes frames/frame<7 digits>.bgeo (the reference's USE_PARTIO build)
  bool use_partio;

 private:
  int frame_count;
};

namespace sphb_host {
// The .geo text of `count` particles (what writeFrameToFile writes), for tests.
// The writer's float formatter on its own (tests): every value as operator<<(ostream&, float)
// prints it, each followed by '\n'.
std::string format_floats_like_ostream(const float* values, size_t count);

std::string houdini_geo_text(const particle* particles, unsigned int count, float particle_mass,
                             unsigned int threads = 0);
// The Bgeo-V5 bytes the reference's USE_PARTIO build writes for the same frame: makeData
// (util/partio/PartioFunctions.h:5-65) followed by Partio::write to a ".bgeo" name.
std::string houdini_bgeo_bytes(const particle* particles, const simulation_parameters& parameters,
                               unsigned int threads = 0);
}

#endif
