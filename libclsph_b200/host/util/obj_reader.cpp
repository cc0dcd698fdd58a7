#include "util/obj_reader.h"

#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>

namespace sphb_host {
namespace {

struct corner {
  int v, vt, vn;
  bool operator<(const corner& o) const {
    if (v != o.v) return v < o.v;
    if (vn != o.vn) return vn < o.vn;
    return vt < o.vt;
  }
};

inline bool blank(char c) { return c == ' ' || c == '\t'; }
inline bool eol(char c) { return c == '\r' || c == '\n' || c == '\0'; }

// 1-based -> 0-based; 0 stays 0; negative counts back from the `count` elements read so far
inline int resolve(int idx, int count) { return idx > 0 ? idx - 1 : (idx == 0 ? 0 : count + idx); }

float next_float(const char*& p) {
  p += strspn(p, " \t");
  const float f = (float)atof(p);
  p += strcspn(p, " \t\r");
  return f;
}

// i, i/j, i//k, i/j/k
corner next_corner(const char*& p, int nv, int nvn, int nvt) {
  corner c = {-1, -1, -1};
  c.v = resolve(atoi(p), nv);
  p += strcspn(p, "/ \t\r");
  if (*p != '/') return c;
  ++p;
  if (*p == '/') {
    ++p;
    c.vn = resolve(atoi(p), nvn);
    p += strcspn(p, "/ \t\r");
    return c;
  }
  c.vt = resolve(atoi(p), nvt);
  p += strcspn(p, "/ \t\r");
  if (*p != '/') return c;
  ++p;
  c.vn = resolve(atoi(p), nvn);
  p += strcspn(p, "/ \t\r");
  return c;
}

struct builder {
  std::vector<float> v;
  int nvn, nvt;
  std::vector<std::vector<corner> > faces;
  std::string name;
  builder() : nvn(0), nvt(0) {}

  bool flush(std::vector<obj_shape>& shapes, std::string& err) {
    if (faces.empty()) return true;
    obj_shape s;
    s.name = name;
    std::map<corner, unsigned int> seen;
    for (size_t f = 0; f < faces.size(); ++f) {
      const std::vector<corner>& poly = faces[f];
      for (size_t k = 2; k < poly.size(); ++k) {
        const corner tri[3] = {poly[0], poly[k - 1], poly[k]};
        for (int t = 0; t < 3; ++t) {
          std::map<corner, unsigned int>::iterator it = seen.find(tri[t]);
          unsigned int id;
          if (it != seen.end()) {
            id = it->second;
          } else {
            if (tri[t].v < 0 || (size_t)(3 * tri[t].v + 2) >= v.size()) {
              std::ostringstream m;
              m << "OBJ face refers to vertex " << tri[t].v + 1 << " of " << v.size() / 3;
              err = m.str();
              return false;
            }
            s.positions.push_back(v[3 * tri[t].v + 0]);
            s.positions.push_back(v[3 * tri[t].v + 1]);
            s.positions.push_back(v[3 * tri[t].v + 2]);
            id = (unsigned int)(s.positions.size() / 3 - 1);
            seen[tri[t]] = id;
          }
          s.indices.push_back(id);
        }
      }
    }
    shapes.push_back(s);
    faces.clear();
    return true;
  }
};

}  // namespace

std::string read_obj(std::vector<obj_shape>& shapes, const std::string& path) {
  shapes.clear();
  std::ifstream in(path.c_str());
  if (!in) return "Cannot open file [" + path + "]";
  builder b;
  std::string line, err;
  while (std::getline(in, line)) {
    const char* p = line.c_str();
    p += strspn(p, " \t");
    if (*p == '\0' || *p == '#') continue;
    if (p[0] == 'v' && blank(p[1])) {
      p += 2;
      const float x = next_float(p), y = next_float(p), z = next_float(p);
      b.v.push_back(x); b.v.push_back(y); b.v.push_back(z);
    } else if (p[0] == 'v' && p[1] == 'n' && blank(p[2])) {
      ++b.nvn;
    } else if (p[0] == 'v' && p[1] == 't' && blank(p[2])) {
      ++b.nvt;
    } else if (p[0] == 'f' && blank(p[1])) {
      p += 2;
      p += strspn(p, " \t");
      std::vector<corner> poly;
      while (!eol(*p)) {
        poly.push_back(next_corner(p, (int)(b.v.size() / 3), b.nvn, b.nvt));
        p += strspn(p, " \t\r");
      }
      b.faces.push_back(poly);
    } else if ((p[0] == 'g' || p[0] == 'o') && blank(p[1])) {
      if (!b.flush(shapes, err)) return err;
      std::istringstream words(p + 2);
      b.name.clear();
      words >> b.name;
    }
    // anything else (mtllib, usemtl, s, ...) has no effect on the geometry
  }
  if (!b.flush(shapes, err)) return err;
  return std::string();
}

}  // namespace sphb_host
