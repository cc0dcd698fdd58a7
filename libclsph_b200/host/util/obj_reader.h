// Minimal Wavefront OBJ reader with the observable behaviour of the loader the reference uses
// (tinyobjloader 0.9.6, util/tinyobj/tiny_obj_loader.cc:480-664): a shape is flushed at every
// `g` / `o` line and at end of file; polygons are fan-triangulated in file order
// (:217-240); each shape carries its own flattened position array, de-duplicated on the
// (v, vt, vn) index triple in first-use order (:156-193); indices may be negative (relative)
// (:63-75).  Materials are not read: nothing on the simulation path uses them.
#ifndef SPHB_HOST_OBJ_READER_H_
#define SPHB_HOST_OBJ_READER_H_

#include <string>
#include <vector>

namespace sphb_host {

struct obj_shape {
  std::string name;
  std::vector<float> positions;       // xyz per flattened vertex
  std::vector<unsigned int> indices;  // 3 per triangle, into positions
};

// Returns an empty string on success, else an error message.
std::string read_obj(std::vector<obj_shape>& shapes, const std::string& path);

}  // namespace sphb_host
#endif
