// A small recursive-descent JSON reader for the two property files (fluid_properties/*.json,
// simulation_properties/*.json).  Like the reader the reference uses (picojson's operator>>,
// libclsph/sph_simulation.cpp:441-443), it stops after the first complete value, so the
// trailing ';' those files end with is tolerated.
#ifndef SPHB_HOST_JSON_MIN_H_
#define SPHB_HOST_JSON_MIN_H_

#include <cstdlib>
#include <istream>
#include <iterator>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace sphb_host {

class json {
 public:
  enum kind { null_k, bool_k, number_k, string_k, array_k, object_k };
  json() : k_(null_k), b_(false), n_(0) {}

  static json parse(std::istream& in) {
    const std::string text((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
    size_t pos = 0;
    return value(text, pos);
  }

  kind type() const { return k_; }
  bool has(const std::string& key) const { return k_ == object_k && o_.count(key) != 0; }
  const json& operator[](const std::string& key) const {
    if (k_ != object_k) throw std::runtime_error("JSON: not an object (looking up '" + key + "')");
    std::map<std::string, json>::const_iterator it = o_.find(key);
    if (it == o_.end()) throw std::runtime_error("JSON: missing key '" + key + "'");
    return it->second;
  }
  double number() const {
    if (k_ != number_k) throw std::runtime_error("JSON: expected a number");
    return n_;
  }
  bool boolean() const {
    if (k_ != bool_k) throw std::runtime_error("JSON: expected true/false");
    return b_;
  }
  const std::string& string() const {
    if (k_ != string_k) throw std::runtime_error("JSON: expected a string");
    return s_;
  }
  const std::vector<json>& array() const { return a_; }

 private:
  kind k_;
  bool b_;
  double n_;
  std::string s_;
  std::vector<json> a_;
  std::map<std::string, json> o_;

  static void skip(const std::string& t, size_t& p) {
    while (p < t.size() && (t[p] == ' ' || t[p] == '\t' || t[p] == '\n' || t[p] == '\r')) ++p;
  }
  static void expect(const std::string& t, size_t& p, char c) {
    skip(t, p);
    if (p >= t.size() || t[p] != c)
      throw std::runtime_error(std::string("JSON: expected '") + c + "' at offset " + std::to_string(p));
    ++p;
  }
  static std::string quoted(const std::string& t, size_t& p) {
    expect(t, p, '"');
    std::string out;
    while (p < t.size() && t[p] != '"') {
      if (t[p] == '\\' && p + 1 < t.size()) {
        const char e = t[++p];
        out += e == 'n' ? '\n' : e == 't' ? '\t' : e == 'r' ? '\r' : e;
      } else {
        out += t[p];
      }
      ++p;
    }
    if (p >= t.size()) throw std::runtime_error("JSON: unterminated string");
    ++p;
    return out;
  }
  static json value(const std::string& t, size_t& p) {
    skip(t, p);
    if (p >= t.size()) throw std::runtime_error("JSON: empty input");
    json j;
    const char c = t[p];
    if (c == '{') {
      j.k_ = object_k;
      ++p;
      skip(t, p);
      if (p < t.size() && t[p] == '}') { ++p; return j; }
      for (;;) {
        const std::string key = quoted(t, p);
        expect(t, p, ':');
        j.o_[key] = value(t, p);
        skip(t, p);
        if (p < t.size() && t[p] == ',') { ++p; skip(t, p); continue; }
        expect(t, p, '}');
        return j;
      }
    }
    if (c == '[') {
      j.k_ = array_k;
      ++p;
      skip(t, p);
      if (p < t.size() && t[p] == ']') { ++p; return j; }
      for (;;) {
        j.a_.push_back(value(t, p));
        skip(t, p);
        if (p < t.size() && t[p] == ',') { ++p; continue; }
        expect(t, p, ']');
        return j;
      }
    }
    if (c == '"') {
      j.k_ = string_k;
      j.s_ = quoted(t, p);
      return j;
    }
    if (t.compare(p, 4, "true") == 0) { j.k_ = bool_k; j.b_ = true; p += 4; return j; }
    if (t.compare(p, 5, "false") == 0) { j.k_ = bool_k; j.b_ = false; p += 5; return j; }
    if (t.compare(p, 4, "null") == 0) { p += 4; return j; }
    char* end = 0;
    const double d = strtod(t.c_str() + p, &end);
    if (end == t.c_str() + p) throw std::runtime_error("JSON: unexpected character at offset " + std::to_string(p));
    p = (size_t)(end - t.c_str());
    j.k_ = number_k;
    j.n_ = d;
    return j;
  }
};

}  // namespace sphb_host
#endif
