// houdini_file_saver — same surface as libclsph/file_save_delegates/houdini_file_saver.h:9-20.
// Writes one ASCII Houdini "PGEOMETRY V5" .geo per call to
// <prefix>frames/frame<7 digits>.geo, byte-identical to what the reference's
// HoudiniFileDumpParticles::dump produces (util/houdini_geo/HoudiniFileDumpHelper.cpp:19-90),
// but formatted in parallel into memory and written with one write per chunk.
#ifndef _HOUDINI_FILE_SAVER_H_
#define _HOUDINI_FILE_SAVER_H_

#include <string>

#include "common/structures.h"

class houdini_file_saver {
 public:
  houdini_file_saver(std::string frames_folder_prefix)
      : frames_folder_prefix(frames_folder_prefix), frame_count(0) {}

  int writeFrameToFile(particle* particles, const simulation_parameters& parameters);

  std::string frames_folder_prefix;

 private:
  int frame_count;
};

namespace sphb_host {
// The .geo text of `count` particles (what writeFrameToFile writes), for tests.
std::string houdini_geo_text(const particle* particles, unsigned int count, float particle_mass,
                             unsigned int threads = 0);
}

#endif
