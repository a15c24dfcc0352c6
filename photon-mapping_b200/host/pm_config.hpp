// pm_config.hpp — reader for the reference's config.toml (common/src/configLoader.h:6-29,
// config.toml.example:1-27). Only the TOML subset that file uses: [section] headers, key = value with
// strings, integers (with '_' separators), floats, booleans and flat arrays of numbers; '#' comments.
#pragma once
#include <cctype>
#include <cstdlib>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace pmh {

struct TomlValue {
  enum Kind { Str, Num, Arr, Boolean } kind = Num;
  std::string s;
  double num = 0;
  std::vector<double> arr;
};

class Config {
 public:
  // the reference reads "../config.toml" relative to the working directory (configLoader.h:6)
  static Config load(const std::string& path) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("cannot open config file " + path);
    Config c;
    std::string line, section;
    int ln = 0;
    while (std::getline(f, line)) {
      ln++;
      line = strip(uncomment(line));
      if (line.empty()) continue;
      if (line.front() == '[') {
        if (line.back() != ']') throw std::runtime_error(path + ":" + std::to_string(ln) + ": bad section header");
        section = strip(line.substr(1, line.size() - 2));
        continue;
      }
      const size_t eq = line.find('=');
      if (eq == std::string::npos) throw std::runtime_error(path + ":" + std::to_string(ln) + ": expected key = value");
      const std::string key = strip(line.substr(0, eq));
      c.v_[section + "." + key] = parse_value(strip(line.substr(eq + 1)), path, ln);
    }
    return c;
  }

  bool has(const std::string& sec, const std::string& key) const { return v_.count(sec + "." + key) != 0; }
  const TomlValue& at(const std::string& sec, const std::string& key) const {
    auto it = v_.find(sec + "." + key);
    if (it == v_.end()) throw std::runtime_error("config: missing [" + sec + "] " + key);  // toml::find throws too
    return it->second;
  }
  std::string str(const std::string& sec, const std::string& key) const {
    const TomlValue& v = at(sec, key);
    if (v.kind != TomlValue::Str) throw std::runtime_error("config: [" + sec + "] " + key + " is not a string");
    return v.s;
  }
  double num(const std::string& sec, const std::string& key) const {
    const TomlValue& v = at(sec, key);
    if (v.kind != TomlValue::Num) throw std::runtime_error("config: [" + sec + "] " + key + " is not a number");
    return v.num;
  }
  long long integer(const std::string& sec, const std::string& key) const { return (long long)num(sec, key); }
  // toml_to_vec3f / toml_to_vec2i (configLoader.h:17-29)
  std::vector<double> array(const std::string& sec, const std::string& key, size_t n) const {
    const TomlValue& v = at(sec, key);
    if (v.kind != TomlValue::Arr || v.arr.size() != n)
      throw std::runtime_error("config: [" + sec + "] " + key + " must be an array of " + std::to_string(n) + " numbers");
    return v.arr;
  }

 private:
  std::map<std::string, TomlValue> v_;

  static std::string strip(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && std::isspace((unsigned char)s[a])) a++;
    while (b > a && std::isspace((unsigned char)s[b - 1])) b--;
    return s.substr(a, b - a);
  }
  static std::string uncomment(const std::string& s) {
    bool in_str = false;
    for (size_t i = 0; i < s.size(); i++) {
      if (s[i] == '"' && (i == 0 || s[i - 1] != '\\')) in_str = !in_str;
      if (s[i] == '#' && !in_str) return s.substr(0, i);
    }
    return s;
  }
  static double parse_number(std::string t, const std::string& path, int ln) {
    std::string u;
    for (char ch : t) if (ch != '_') u += ch;
    char* end = nullptr;
    const double d = std::strtod(u.c_str(), &end);
    if (u.empty() || end != u.c_str() + u.size()) throw std::runtime_error(path + ":" + std::to_string(ln) + ": bad number '" + t + "'");
    return d;
  }
  static TomlValue parse_value(const std::string& t, const std::string& path, int ln) {
    TomlValue v;
    if (t.empty()) throw std::runtime_error(path + ":" + std::to_string(ln) + ": empty value");
    if (t.front() == '"' || t.front() == '\'') {
      const char q = t.front();
      const size_t e = t.rfind(q);
      if (e == 0) throw std::runtime_error(path + ":" + std::to_string(ln) + ": unterminated string");
      v.kind = TomlValue::Str;
      const std::string raw = t.substr(1, e - 1);
      if (q == '\'') { v.s = raw; return v; }
      for (size_t i = 0; i < raw.size(); i++) {  // basic-string escapes
        if (raw[i] == '\\' && i + 1 < raw.size()) {
          i++;
          v.s += raw[i] == 'n' ? '\n' : raw[i] == 't' ? '\t' : raw[i];
        } else {
          v.s += raw[i];
        }
      }
      return v;
    }
    if (t.front() == '[') {
      if (t.back() != ']') throw std::runtime_error(path + ":" + std::to_string(ln) + ": arrays must close on the same line");
      v.kind = TomlValue::Arr;
      std::stringstream ss(t.substr(1, t.size() - 2));
      std::string item;
      while (std::getline(ss, item, ',')) {
        item = strip(item);
        if (!item.empty()) v.arr.push_back(parse_number(item, path, ln));
      }
      return v;
    }
    if (t == "true" || t == "false") { v.kind = TomlValue::Boolean; v.num = t == "true"; return v; }
    v.kind = TomlValue::Num;
    v.num = parse_number(t, path, ln);
    return v;
  }
};

}  // namespace pmh
