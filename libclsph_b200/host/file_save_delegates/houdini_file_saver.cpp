#include "file_save_delegates/houdini_file_saver.h"

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <thread>
#include <vector>

namespace sphb_host {
namespace {

// operator<<(ostream&, float) with default flags is printf("%g") (precision 6).  put_float writes
// exactly those bytes without going through printf for almost every value: the 6 significant
// digits come from one multiplication by a power of ten and one rounding, and whenever that
// rounding could be a close call (the scaled value within 1e-6 of a tie — the multiplication is
// good to ~1e-10 there), or the value is not an ordinary finite number, printf itself is asked.
inline char* put_float_slow(char* p, float f) { return p + std::sprintf(p, "%g", (double)f); }

struct Pow10Table {
  double v[121];  // 10^-60 ... 10^60
  Pow10Table() {
    for (int k = 0; k <= 120; ++k) {
      char lit[16];
      std::snprintf(lit, sizeof lit, "1e%d", k - 60);
      v[k] = std::strtod(lit, nullptr);  // correctly rounded
    }
  }
  double operator()(int e) const { return v[e + 60]; }
};
const Pow10Table kPow10;

inline char* put_float(char* p, float f) {
  uint32_t bits;
  std::memcpy(&bits, &f, 4);
  const uint32_t mag = bits & 0x7fffffffu;
  if (mag == 0u) {  // "0" / "-0"
    if (bits) *p++ = '-';
    *p++ = '0';
    return p;
  }
  if (mag < 0x00800000u || mag >= 0x7f800000u) return put_float_slow(p, f);  // denormal, inf, nan
  const double a = std::fabs((double)f);
  // decimal exponent: from the binary one (log10(2) = 0.30103), corrected by the scaled value itself
  int e10 = (int)(((int)(mag >> 23) - 127) * 0.30102999566398120);
  if ((int)(mag >> 23) - 127 < 0) --e10;
  double scaled = a * kPow10(5 - e10);
  if (scaled < 1e5) { --e10; scaled = a * kPow10(5 - e10); }
  else if (scaled >= 1e6) { ++e10; scaled = a * kPow10(5 - e10); }
  if (!(scaled >= 1e5 && scaled < 1e6)) return put_float_slow(p, f);
  const double fl = std::floor(scaled), frac = scaled - fl;
  if (std::fabs(frac - 0.5) < 1e-6) return put_float_slow(p, f);  // a tie, or too close to call here
  uint32_t digits = (uint32_t)fl + (frac > 0.5 ? 1u : 0u);
  if (digits == 1000000u) { digits = 100000u; ++e10; }
  if (bits >> 31) *p++ = '-';
  char d[6];
  for (int k = 5; k >= 0; --k) { d[k] = (char)('0' + digits % 10u); digits /= 10u; }
  int nd = 6;
  while (nd > 1 && d[nd - 1] == '0') --nd;  // %g drops trailing zeros
  if (e10 < -4 || e10 >= 6) {  // d.ddddde+XX
    *p++ = d[0];
    if (nd > 1) { *p++ = '.'; for (int k = 1; k < nd; ++k) *p++ = d[k]; }
    *p++ = 'e';
    int e = e10;
    if (e < 0) { *p++ = '-'; e = -e; } else *p++ = '+';
    if (e >= 100) { *p++ = (char)('0' + e / 100); e %= 100; }
    *p++ = (char)('0' + e / 10);
    *p++ = (char)('0' + e % 10);
  } else if (e10 >= 0) {  // ddd.ddd
    for (int k = 0; k <= e10; ++k) *p++ = k < nd ? d[k] : '0';
    if (nd > e10 + 1) { *p++ = '.'; for (int k = e10 + 1; k < nd; ++k) *p++ = d[k]; }
  } else {  // 0.000ddd
    *p++ = '0'; *p++ = '.';
    for (int k = -1; k > e10; --k) *p++ = '0';
    for (int k = 0; k < nd; ++k) *p++ = d[k];
  }
  return p;
}
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
    char* q = buf + sizeof buf;  // " %d", digits written backwards
    unsigned int v = i;
    do { *--q = (char)('0' + v % 10u); v /= 10u; } while (v);
    *--q = ' ';
    out->append(q, (size_t)(buf + sizeof buf - q));
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

// ---- Bgeo V5 (binary Houdini geometry), every multi-byte value big-endian ----
inline char* put_be32(char* p, uint32_t v) {
  p[0] = (char)(v >> 24); p[1] = (char)(v >> 16); p[2] = (char)(v >> 8); p[3] = (char)v;
  return p + 4;
}
inline char* put_be16(char* p, unsigned int v) {
  p[0] = (char)(v >> 8); p[1] = (char)v;
  return p + 2;
}
inline char* put_bef(char* p, float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  return put_be32(p, u);
}
inline void append_be32(std::string& s, uint32_t v) { char b[4]; put_be32(b, v); s.append(b, 4); }
inline void append_be16(std::string& s, unsigned int v) { char b[2]; put_be16(b, v); s.append(b, 2); }
inline void append_hstr(std::string& s, const char* text) {  // 16-bit length, then the characters
  const size_t n = std::strlen(text);
  append_be16(s, (unsigned int)n);
  s.append(text, n);
}
// a point attribute's dictionary entry: name, size, Houdini type (0 float, 1 int, 5 vector),
// `size` zero defaults
inline void append_attr(std::string& s, const char* name, unsigned int size, uint32_t type) {
  append_hstr(s, name);
  append_be16(s, size);
  append_be32(s, type);
  for (unsigned int k = 0; k < size; ++k) append_be32(s, 0);
}

constexpr size_t kBgeoWordsPerPoint = 4 + 3 + 3 + 1 + 1 + 1;

// points [lo, hi): position + homogeneous 1, then the attributes in the order makeData adds
// them (velocity, color, id, mass, pscale)
void bgeo_points(const particle* ps, unsigned int lo, unsigned int hi, float mass, float h, char* out) {
  for (unsigned int i = lo; i < hi; ++i) {
    const particle& q = ps[i];
    float r, g, b;
    density_colour(q.density, r, g, b);
    out = put_bef(out, q.position.s[0]);
    out = put_bef(out, q.position.s[1]);
    out = put_bef(out, q.position.s[2]);
    out = put_bef(out, 1.f);
    // PartioFunctions.h:36-38 assigns vel[0] three times: y and z are never written and keep
    // what partio's freshly grown attribute array held — zero here
    out = put_bef(out, q.velocity.s[0]);
    out = put_bef(out, 0.f);
    out = put_bef(out, 0.f);
    out = put_bef(out, r);
    out = put_bef(out, g);
    out = put_bef(out, b);
    out = put_be32(out, (uint32_t)i);
    out = put_bef(out, mass);
    out = put_bef(out, h);
  }
}

void bgeo_indices(unsigned int lo, unsigned int hi, bool wide, char* out) {
  if (wide)
    for (unsigned int i = lo; i < hi; ++i) out = put_be32(out, i);
  else
    for (unsigned int i = lo; i < hi; ++i) out = put_be16(out, i);
}

}  // namespace

std::string format_floats_like_ostream(const float* values, size_t count) {
  std::string out;
  out.reserve(count * 12);
  char buf[32];
  for (size_t k = 0; k < count; ++k) {
    char* p = put_float(buf, values[k]);
    *p++ = '\n';
    out.append(buf, (size_t)(p - buf));
  }
  return out;
}

std::string houdini_bgeo_bytes(const particle* particles, const simulation_parameters& parameters,
                               unsigned int threads) {
  const unsigned int count = parameters.particles_count;
  if (threads == 0) {
    threads = std::thread::hardware_concurrency();
    if (threads == 0) threads = 1;
  }
  if (threads > 64) threads = 64;
  if (count < 4096) threads = 1;
  std::string head;
  append_be32(head, 0x4267656fu);  // "Bgeo"
  head.push_back('V');
  append_be32(head, 5);
  append_be32(head, count);  // NPoints
  append_be32(head, 1);      // NPrims: one particle-system primitive
  append_be32(head, 0);      // NPointGroups
  append_be32(head, 0);      // NPrimGroups
  append_be32(head, 5);      // NPointAttrib: everything but the position
  append_be32(head, 0);      // NVertexAttrib
  append_be32(head, 1);      // NPrimAttrib
  append_be32(head, 0);      // NAttrib
  append_attr(head, "velocity", 3, 5);
  append_attr(head, "color", 3, 5);
  append_attr(head, "id", 1, 1);
  append_attr(head, "mass", 1, 0);
  append_attr(head, "pscale", 1, 0);
  std::string mid;  // primitive attribute dictionary, then the primitive's header
  append_hstr(mid, "generator");
  append_be16(mid, 1);
  append_be32(mid, 4);  // index-string attribute ...
  append_be32(mid, 1);  // ... with one string
  append_hstr(mid, "papi");
  append_be32(mid, 0x8000);  // particle system
  append_be32(mid, count);
  const bool wide = (int)count > (1 << 16);  // vertex numbers: 16-bit up to 65536 points
  std::string tail;
  append_be32(tail, 0);  // the primitive's "generator" index
  tail.push_back((char)0x00);
  tail.push_back((char)0xff);

  const size_t points_bytes = (size_t)count * kBgeoWordsPerPoint * 4, index_bytes = (size_t)count * (wide ? 4 : 2);
  std::string all;
  all.resize(head.size() + points_bytes + mid.size() + index_bytes + tail.size());
  char* base = &all[0];
  std::memcpy(base, head.data(), head.size());
  char* points = base + head.size();
  std::memcpy(points + points_bytes, mid.data(), mid.size());
  char* indices = points + points_bytes + mid.size();
  std::memcpy(indices + index_bytes, tail.data(), tail.size());
  std::vector<std::thread> pool;
  for (unsigned int t = 0; t < threads; ++t) {
    const unsigned int lo = (unsigned int)((unsigned long long)count * t / threads);
    const unsigned int hi = (unsigned int)((unsigned long long)count * (t + 1) / threads);
    const float mass = parameters.particle_mass, h = parameters.h;
    pool.push_back(std::thread([=] {
      bgeo_points(particles, lo, hi, mass, h, points + (size_t)lo * kBgeoWordsPerPoint * 4);
      bgeo_indices(lo, hi, wide, indices + (size_t)lo * (wide ? 4 : 2));
    }));
  }
  for (size_t t = 0; t < pool.size(); ++t) pool[t].join();
  return all;
}

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
  if (use_partio) {  // houdini_file_saver.cpp:78-89
    const std::string fileName = frames_folder_prefix + "frames/frame" + name + ".bgeo";
    const std::string bytes = sphb_host::houdini_bgeo_bytes(particles, parameters, 0);
    std::ofstream out(fileName.c_str(), std::ios::out | std::ios::binary);
    // Partio::write reports failure on stderr and returns; the delegate still returns 0
    if (!out.is_open()) {
      std::cerr << "Partio: Unable to open file " << fileName << std::endl;
      return 0;
    }
    out.write(bytes.data(), (std::streamsize)bytes.size());
    return 0;
  }
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
