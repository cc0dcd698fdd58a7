// ThreadSanitizer driver for the threaded host-side code (no GPU): frame gather, deal-out, .geo/.bgeo formatting
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "clsph_host.h"
int main() {
  const uint32_t n = 200000;
  const int world = 8;
  std::vector<sphb_particle> p(n), stage(n), out(n);
  std::vector<uint32_t> ids(n), sids(n), oids(n), start(world + 1);
  std::memset(p.data(), 0, sizeof(sphb_particle) * n);
  srand(1);
  for (uint32_t i = 0; i < n; ++i) {
    p[i].position.s[0] = (float)rand() / RAND_MAX;
    p[i].grid_index = rand() % 5000;
    ids[i] = i;
  }
  std::vector<float> cuts(world + 1);
  if (clsph_equal_count_cuts(p.data(), n, world, cuts.data()) != world + 1) return 1;
  for (int rep = 0; rep < 3; ++rep) {
    if (clsph_deal_out(p.data(), ids.data(), n, cuts.data(), world, world, stage.data(), sids.data(), start.data())) return 2;
    if (clsph_gather_sort(stage.data(), sids.data(), start.data(), world, 4096, out.data(), oids.data())) return 3;
  }
  sphb_params prm;
  std::memset(&prm, 0, sizeof prm);
  prm.particles_count = n;
  const size_t len = clsph_geo_text(p.data(), n, 0.02f, 8, 0, 0);
  std::vector<char> text(len + 1);
  if (clsph_geo_text(p.data(), n, 0.02f, 8, text.data(), text.size()) != len) return 4;
  const size_t blen = clsph_bgeo_bytes(p.data(), &prm, 8, 0, 0);
  std::vector<char> bytes(blen + 1);
  if (clsph_bgeo_bytes(p.data(), &prm, 8, bytes.data(), bytes.size()) != blen) return 5;
  std::printf("ok %u %zu\n", start[world], len);
  return 0;
}
