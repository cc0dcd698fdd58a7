#include "file_save_delegates/houdini_file_saver.h"

#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <thread>
#include <vector>

namespace sphb_host {
namespace {

// operator<<(ostream&, float) with default flags is printf("%g") (precision 6).
inline char* put_float(char* p, float f) { return p + std::sprintf(p, "%g", (double)f); }
inline char* put_char(char* p, char c) { *p = c; return p + 1; }

// density -> colour ramp of houdini_file_saver.cpp:47-61
inline void density_colour(float d, float& r, float& g, float& b) {
  r = (d > 1000.f && d <= 2000.f) ? (d - 1000.f) / 1000.f : 0.f;
  g = (d >= 0.f && d < 1000.f) ? 1.f - (d) / 1000.f : 0.f;
  b = (d >= 500.f && d <= 1000.f) ? (d - 500.f) / 500.f
      : (d >= 1000.f && d <= 1500.f) ? 1.f - (d - 1000.f) / 500.f : 0.f;
}

// "px py pz 0 (vx vy vz\tr g b\tmass)\n" for particles [lo, hi)
void format_points(const particle* ps, unsigned int lo, unsigned int hi, float mass, std::string* out) {
  std::vector<char> line(16 * 11);
  out->reserve((size_t)(hi - lo) * 96);
  for (unsigned int i = lo; i < hi; ++i) {
    const particle& q = ps[i];
    float r, g, b;
    density_colour(q.density, r, g, b);
    char* p = line.data();
    p = put_float(p, q.position.s[0]); p = put_char(p, ' ');
    p = put_float(p, q.position.s[1]); p = put_char(p, ' ');
    p = put_float(p, q.position.s[2]); p = put_char(p, ' ');
    p = put_float(p, 0.f); p = put_char(p, ' ');
    p = put_char(p, '(');
    p = put_float(p, q.velocity.s[0]); p = put_char(p, ' ');
    p = put_float(p, q.velocity.s[1]); p = put_char(p, ' ');
    p = put_float(p, q.velocity.s[2]); p = put_char(p, '\t');
    p = put_float(p, r); p = put_char(p, ' ');
    p = put_float(p, g); p = put_char(p, ' ');
    p = put_float(p, b); p = put_char(p, '\t');
    p = put_float(p, mass);
    p = put_char(p, ')'); p = put_char(p, '\n');
    out->append(line.data(), (size_t)(p - line.data()));
  }
}

void format_ids(unsigned int lo, unsigned int hi, std::string* out) {
  char buf[16];
  out->reserve((size_t)(hi - lo) * 9);
  for (unsigned int i = lo; i < hi; ++i) {
    const int n = std::sprintf(buf, " %d", (int)i);
    out->append(buf, (size_t)n);
  }
}

void geo_chunks(const particle* ps, unsigned int count, float mass, unsigned int threads,
                std::vector<std::string>& chunks) {
  if (threads == 0) {
    threads = std::thread::hardware_concurrency();
    if (threads == 0) threads = 1;
  }
  if (threads > 64) threads = 64;
  if (count < 4096) threads = 1;
  char head[256];
  std::snprintf(head, sizeof head,
                "PGEOMETRY V5\nNPoints %d NPrims 1\nNPointGroups 0 NPrimGroups 1\n"
                "NPointAttrib 3 NVertexAttrib 0 NPrimAttrib 2 NAttrib 0\nPointAttrib\n"
                "v 3 float 1 1 1\ncolor 3 float 1 1 1\nmass 1 float 1\n",
                (int)count);
  std::vector<std::string> pts(threads), ids(threads);
  std::vector<std::thread> pool;
  for (unsigned int t = 0; t < threads; ++t) {
    const unsigned int lo = (unsigned int)((unsigned long long)count * t / threads);
    const unsigned int hi = (unsigned int)((unsigned long long)count * (t + 1) / threads);
    pool.push_back(std::thread([=, &pts, &ids] {
      format_points(ps, lo, hi, mass, &pts[t]);
      format_ids(lo, hi, &ids[t]);
    }));
  }
  for (size_t t = 0; t < pool.size(); ++t) pool[t].join();
  chunks.clear();
  chunks.push_back(head);
  for (unsigned int t = 0; t < threads; ++t) chunks.push_back(std::string()), chunks.back().swap(pts[t]);
  char mid[160];
  std::snprintf(mid, sizeof mid,
                "PrimitiveAttrib\ngenerator 1 index 1 location1\n"
                "dopobject 1 index 1 /obj/AutoDopNetwork:1\nPart %d", (int)count);
  chunks.push_back(mid);
  for (unsigned int t = 0; t < threads; ++t) chunks.push_back(std::string()), chunks.back().swap(ids[t]);
  chunks.push_back(" [0\t0]\nbox_object1 unordered\n1 1\nbeginExtra\nendExtra\n");
}

}  // namespace

std::string houdini_geo_text(const particle* particles, unsigned int count, float particle_mass,
                             unsigned int threads) {
  std::vector<std::string> chunks;
  geo_chunks(particles, count, particle_mass, threads, chunks);
  std::string all;
  for (size_t c = 0; c < chunks.size(); ++c) all += chunks[c];
  return all;
}

}  // namespace sphb_host

int houdini_file_saver::writeFrameToFile(particle* particles, const simulation_parameters& parameters) {
  char name[32];
  ++frame_count;
  // setw(9)/setfill('0') then keep the last 7 characters (houdini_file_saver.cpp:15-23)
  std::snprintf(name, sizeof name, "%07d", frame_count % 10000000);
  const std::string fileName = frames_folder_prefix + "frames/frame" + name + ".geo";

  std::vector<std::string> chunks;
  sphb_host::geo_chunks(particles, parameters.particles_count, parameters.particle_mass, 0, chunks);
  std::ofstream out(fileName.c_str(), std::ios::out | std::ios::trunc | std::ios::binary);
  if (!out.is_open()) {
    std::cerr << "Error while writing to " << fileName << std::endl;
    return 0;
  }
  for (size_t c = 0; c < chunks.size(); ++c) out.write(chunks[c].data(), (std::streamsize)chunks[c].size());
  return 0;
}
