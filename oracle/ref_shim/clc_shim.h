/* OpenCL-C -> C++11 shim.  TEST INFRASTRUCTURE ONLY (see oracle/README.md).
 *
 * Lets the reference's own kernel source (libclsph/kernels/sph.cl, read from
 * /root/reference at build time, never copied into this repo) compile with g++
 * so that it can be executed one work-item at a time on the CPU.  Everything a
 * real OpenCL device compiler would supply (vector types, built-ins, address
 * space qualifiers, work-item functions) is defined here.  The arithmetic of
 * the built-ins is the part of the reference that is NOT pinned by the
 * reference repo (it belongs to whichever OpenCL compiler runs it); the choices
 * made here are: correctly rounded sqrt and '/', pown as repeated left-to-right
 * multiplication, length = sqrt((x*x + y*y) + z*z), no FMA contraction
 * (the library is built with -ffp-contract=off).
 */
#ifndef ORACLE_CLC_SHIM_H_
#define ORACLE_CLC_SHIM_H_

/* every std header BEFORE the qualifier macros (#define global breaks <locale>) */
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <limits>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

namespace clref {

typedef unsigned int uint;

struct alignas(16) float3 {
  float x, y, z, w;
  float3() : x(0.f), y(0.f), z(0.f), w(0.f) {}
  float3(float s) : x(s), y(s), z(s), w(0.f) {} /* scalar broadcast (smoothing.cl:24) */
  float3(float a, float b, float c) : x(a), y(b), z(c), w(0.f) {}
};

struct alignas(16) uint3 {
  uint x, y, z, w;
  uint3() : x(0), y(0), z(0), w(0) {}
  uint3(uint a, uint b, uint c) : x(a), y(b), z(c), w(0) {}
};

struct alignas(8) uint2 {
  uint x, y;
  uint2() : x(0), y(0) {}
  uint2(size_t a, size_t b) : x((uint)a), y((uint)b) {}
};

inline float3 operator+(float3 a, float3 b) { return float3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline float3 operator-(float3 a, float3 b) { return float3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline float3 operator*(float3 a, float3 b) { return float3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline float3 operator/(float3 a, float3 b) { return float3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline float3 operator*(float3 a, float s) { return float3(a.x * s, a.y * s, a.z * s); }
inline float3 operator*(float s, float3 a) { return float3(s * a.x, s * a.y, s * a.z); }
inline float3 operator/(float3 a, float s) { return float3(a.x / s, a.y / s, a.z / s); }
inline float3 operator-(float3 a) { return float3(-a.x, -a.y, -a.z); }
inline float3& operator+=(float3& a, float3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
inline float3& operator-=(float3& a, float3 b) { a.x -= b.x; a.y -= b.y; a.z -= b.z; return a; }
inline float3& operator/=(float3& a, float s) { a.x /= s; a.y /= s; a.z /= s; return a; }

/* scalar built-ins, float overloads only, declared here so that unqualified
 * lookup from the kernel source finds these before any ::/std:: overload. */
inline float sqrt(float v) { return std::sqrt(v); }
inline float floor(float v) { return std::floor(v); }
inline float fabs(float v) { return std::fabs(v); }
inline float copysign(float a, float b) { return std::copysign(a, b); }
inline float min(float a, float b) { return std::fmin(a, b); }
inline float max(float a, float b) { return std::fmax(a, b); }
inline float clamp(float v, float lo, float hi) { return std::fmin(std::fmax(v, lo), hi); }
inline float convert_float(float v) { return v; }
inline float convert_float(double v) { return (float)v; }
inline float pown(float b, int n) {
  float r = b;
  for (int i = 1; i < n; ++i) r = r * b;
  return n == 0 ? 1.f : r;
}
inline float dot(float3 a, float3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float length(float3 a) { return std::sqrt((a.x * a.x + a.y * a.y) + a.z * a.z); }
inline float length(float a) { return std::fabs(a); } /* smoothing.cl:14 calls length() on a scalar */
inline float distance(float3 a, float3 b) { return length(a - b); }

/* work-item functions: one work-item at a time, ids set by the driver */
struct work_item_ctx {
  size_t global_id, local_id, group_id, local_size, global_size;
};
extern thread_local work_item_ctx g_wi;
inline size_t get_global_id(int) { return g_wi.global_id; }
inline size_t get_local_id(int) { return g_wi.local_id; }
inline size_t get_group_id(int) { return g_wi.group_id; }
inline size_t get_local_size(int) { return g_wi.local_size; }
inline size_t get_global_size(int) { return g_wi.global_size; }
inline void barrier(int) {}

}  // namespace clref

#define CLK_LOCAL_MEM_FENCE 1

/* address-space / function qualifiers */
#define kernel
#define __kernel
#define global
#define __global
#define local
#define __local
#define constant
#define __const const

#endif
