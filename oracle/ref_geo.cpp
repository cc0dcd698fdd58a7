/* "reference" frame writer: wraps the reference's own houdini_file_saver + Houdini dump helper,
 * compiled where they lie under /root/reference (see oracle/Makefile), so tests can check the
 * product's .geo writer byte for byte.  TEST INFRASTRUCTURE ONLY. */
#include <string>

#include "file_save_delegates/houdini_file_saver.h" /* -I/root/reference/libclsph */

extern "C" {
/* writes <prefix>frames/frame0000001.geo ... one file per call, `calls` times */
int orcgeo_write(const char* prefix, void* particles, const void* params, int calls) {
  houdini_file_saver saver((std::string(prefix)));
  for (int k = 0; k < calls; ++k)
    saver.writeFrameToFile(static_cast<particle*>(particles),
                           *static_cast<const simulation_parameters*>(params));
  return 0;
}
}
