// Host scene preparation; behaviour of libclsph/scene.cpp:9-153 (all arithmetic in fp32, same
// expression order, because the distance field and the collision pass are compared with the
// reference's bit for bit).
#include "scene.h"

#include <cmath>
#include <iostream>
#include <limits>

#include "util/obj_reader.h"

namespace {
inline float len3(float x, float y, float z) { return std::sqrt(x * x + y * y + z * z); }
inline size_t grid_points_along(float lo, float hi, float inflate) {
  // spacing ~ inflate/2 (= h): ceil(extent / inflate * 2) + 1     (scene.cpp:109-114)
  return (size_t)(std::ceil((hi - lo) / inflate * 2) + 1);
}
}  // namespace

void scene::add_face(const float* a, const float* b, const float* c, float inflate) {
  const float ux = b[0] - a[0], uy = b[1] - a[1], uz = b[2] - a[2];
  const float vx = c[0] - a[0], vy = c[1] - a[1], vz = c[2] - a[2];
  float nx = uy * vz - uz * vy;
  float ny = uz * vx - ux * vz;
  float nz = ux * vy - uy * vx;
  const float nlen = len3(nx, ny, nz);
  const float ulen = len3(ux, uy, uz);
  nx = nx / nlen; ny = ny / nlen; nz = nz / nlen;
  face_normals.push_back(nx);
  face_normals.push_back(ny);
  face_normals.push_back(nz);

  // AABB of the three vertices; the running min/max start from INT limits converted to
  // float, exactly as scene.cpp:47-52 does
  BB box;
  box.maxx = box.maxy = box.maxz = (float)std::numeric_limits<cl_int>::min();
  box.minx = box.miny = box.minz = (float)std::numeric_limits<cl_int>::max();
  const float* tri[3] = {a, b, c};
  for (int k = 0; k < 3; ++k) {
    if (tri[k][0] > box.maxx) box.maxx = tri[k][0];
    if (tri[k][1] > box.maxy) box.maxy = tri[k][1];
    if (tri[k][2] > box.maxz) box.maxz = tri[k][2];
    if (tri[k][0] < box.minx) box.minx = tri[k][0];
    if (tri[k][1] < box.miny) box.miny = tri[k][1];
    if (tri[k][2] < box.minz) box.minz = tri[k][2];
  }
  box.maxx += inflate; box.maxy += inflate; box.maxz += inflate;
  box.minx -= inflate; box.miny -= inflate; box.minz -= inflate;
  box.size_x = grid_points_along(box.minx, box.maxx, inflate);
  box.size_y = grid_points_along(box.miny, box.maxy, inflate);
  box.size_z = grid_points_along(box.minz, box.maxz, inflate);
  box.offset = totalGridpoints;
  totalGridpoints += box.size_x * box.size_y * box.size_z;
  bbs.push_back(box);

  // orthonormal face frame: n, w = u^ x n, u^   (scene.cpp:122-143)
  const float hx = ux / ulen, hy = uy / ulen, hz = uz / ulen;
  const float wx = hy * nz - hz * ny;
  const float wy = hz * nx - hx * nz;
  const float wz = hx * ny - hy * nx;
  const float rows[12] = {nx, ny, nz, -a[0], wx, wy, wz, -a[1], hx, hy, hz, -a[2]};
  transforms.insert(transforms.end(), rows, rows + 12);

  // the two other vertices in the (w, u^) plane   (scene.cpp:145-149)
  rvertices.push_back(ux * wx + uy * wy + uz * wz);
  rvertices.push_back(ux * hx + uy * hy + uz * hz);
  rvertices.push_back(wx * vx + wy * vy + wz * vz);
  rvertices.push_back(hx * vx + hy * vy + hz * vz);
}

bool scene::load_path(const std::string& path, float distFieldThreshold) {
  totalGridpoints = 0;
  face_count = 0;
  std::vector<sphb_host::obj_shape> shapes;
  const std::string err = sphb_host::read_obj(shapes, path);
  if (!err.empty()) {
    std::cerr << err << std::endl;
    return false;
  }
  std::cout << "Scene Loading - number of shapes in file [" << path << "]: " << shapes.size()
            << std::endl;
  for (size_t s = 0; s < shapes.size(); ++s) {
    indices = shapes[s].indices;
    if (indices.size() % 3 != 0) {
      std::cerr << "Meshes must be made of triangles only" << std::endl;
      return false;
    }
    vertices = shapes[s].positions;
    const size_t nfaces = indices.size() / 3;
    face_count += (unsigned int)nfaces;
    for (size_t f = 0; f < nfaces; ++f)
      add_face(&vertices[3 * indices[3 * f + 0]], &vertices[3 * indices[3 * f + 1]],
               &vertices[3 * indices[3 * f + 2]], distFieldThreshold);
  }
  return true;
}

bool scene::load(std::string filename, float distFieldThreshold) {
  return load_path(std::string("scenes/") + filename, distFieldThreshold);
}
