// Host-side stand-ins for the OpenCL scalar / vector typedefs that libclsph's public structs
// are written in (the reference gets them from <CL/cl.hpp> through util/cl_boilerplate.h:10-20).
// No OpenCL exists behind this engine; only the memory layout is kept, because user code reads
// `particles[i].position.s[0]` (libclsph/file_save_delegates/houdini_file_saver.cpp:40-45) and
// last_frame.bin is a raw dump of 80-byte particles (example/particles.cpp:85).
#ifndef SPHB_HOST_CL_TYPES_H_
#define SPHB_HOST_CL_TYPES_H_

#include <stdint.h>

// The reference's common/structures.h reaches <iostream>, <string>, <vector>, <fstream> and
// <sstream> through util/cl_boilerplate.h:4-8, and its callers rely on that (example/particles.cpp
// uses std::ofstream / std::filebuf without including <fstream>): the same headers come along here.
#ifdef __cplusplus
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>
#endif

typedef int32_t cl_int;
typedef uint32_t cl_uint;
typedef float cl_float;

// OpenCL's float3 / uint3 occupy (and align to) four lanes.
typedef union alignas(16) {
  cl_float s[4];
  struct { cl_float x, y, z, w; };
} cl_float3;

typedef union alignas(16) {
  cl_uint s[4];
  struct { cl_uint x, y, z, w; };
} cl_uint3;

#endif
