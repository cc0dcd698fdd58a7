// class scene — same public surface as libclsph/scene.h:8-20.
#ifndef _SCENE_H_
#define _SCENE_H_
#include "common/structures.h"

#include <string>
#include <vector>

class scene {
 public:
  scene() : face_count(0), totalGridpoints(0) {}

  // Loads "scenes/<filename>" (triangles only) and derives, per face, what the collision pass
  // needs: normal, AABB inflated by distFieldThreshold with its distance-field grid shape,
  // the face-local frame and the face's other two vertices in that frame.
  bool load(std::string filename, float distFieldThreshold);
  // Same, from an explicit path (no "scenes/" prefix) — used by tools that generate scenes.
  bool load_path(const std::string& path, float distFieldThreshold);

  unsigned int face_count;
  std::vector<float> face_normals;    // 3 per face
  std::vector<float> vertices;        // positions of the LAST shape read (as in the reference)
  std::vector<unsigned int> indices;  // indices of the LAST shape read
  std::vector<BB> bbs;                // 1 per face, all shapes
  std::vector<float> transforms;      // 12 per face: rows [n | -v0.x], [u x n | -v0.y], [u | -v0.z]
  std::vector<float> rvertices;       // 4 per face: v1, v2 in the (u x n, u) plane frame
  cl_uint totalGridpoints;

 private:
  void add_face(const float* a, const float* b, const float* c, float inflate);
};

#endif
