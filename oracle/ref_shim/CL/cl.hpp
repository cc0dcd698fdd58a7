/* Stand-in for <CL/cl.hpp> so that the reference's HOST files that only need the
 * cl_* scalar/vector typedefs (libclsph/scene.cpp via common/structures.h ->
 * util/cl_boilerplate.h) compile without an OpenCL SDK.  TEST INFRASTRUCTURE ONLY. */
#ifndef ORACLE_CL_HPP_STUB_
#define ORACLE_CL_HPP_STUB_
#include <cstdint>
#include <limits>
typedef int32_t cl_int;
typedef uint32_t cl_uint;
typedef float cl_float;
typedef union { float s[4]; struct { float x, y, z, w; }; } __attribute__((aligned(16))) cl_float3;
typedef union { uint32_t s[4]; struct { uint32_t x, y, z, w; }; } __attribute__((aligned(16))) cl_uint3;
#define CL_SUCCESS 0
namespace cl {
class Context {};
class Device {};
class Program {};
}
#endif
