/* "reference" scene loader: wraps the reference's own libclsph/scene.cpp (+ its vendored
 * tinyobj), compiled where it lies under /root/reference (see oracle/Makefile), behind a C
 * API so tests can check the product's scene loader against it.  TEST INFRASTRUCTURE ONLY. */
#include <unistd.h>
#include <climits>
#include <cstring>
#include <string>

#include "scene.h" /* -I/root/reference/libclsph */

extern "C" {

/* scene::load reads "scenes/<filename>" relative to the CWD (libclsph/scene.cpp:16-17) */
void* orcscene_load(const char* root_dir, const char* filename, float threshold) {
  char cwd[PATH_MAX];
  if (!getcwd(cwd, sizeof cwd)) return 0;
  if (chdir(root_dir) != 0) return 0;
  scene* s = new scene();
  bool ok = s->load(filename, threshold);
  if (chdir(cwd) != 0) { /* nothing sensible to do */ }
  if (!ok) {
    delete s;
    return 0;
  }
  return s;
}

unsigned orcscene_face_count(void* h) { return static_cast<scene*>(h)->face_count; }
unsigned orcscene_total_gridpoints(void* h) { return static_cast<scene*>(h)->totalGridpoints; }
unsigned orcscene_bb_size(void) { return sizeof(BB); }

/* bbs: face_count*56 B, transforms: 12 f32/face, rvertices: 4 f32/face, normals: 3 f32/face */
void orcscene_copy(void* h, void* bbs, float* transforms, float* rvertices, float* normals) {
  scene* s = static_cast<scene*>(h);
  if (bbs) std::memcpy(bbs, s->bbs.data(), s->bbs.size() * sizeof(BB));
  if (transforms) std::memcpy(transforms, s->transforms.data(), s->transforms.size() * 4);
  if (rvertices) std::memcpy(rvertices, s->rvertices.data(), s->rvertices.size() * 4);
  if (normals) std::memcpy(normals, s->face_normals.data(), s->face_normals.size() * 4);
}

void orcscene_free(void* h) { delete static_cast<scene*>(h); }
}
