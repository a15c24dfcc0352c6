// pm_json.hpp — the small JSON reader the glb loader needs (objects, arrays, numbers, strings, bools, null).
#pragma once
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace pmh {

struct Json {
  enum Type { Null, Bool, Number, String, Array, Object } type = Null;
  bool b = false;
  double num = 0;
  std::string str;
  std::vector<Json> arr;
  std::map<std::string, Json> obj;

  bool has(const std::string& k) const { return type == Object && obj.count(k) != 0; }
  const Json& at(const std::string& k) const {
    auto it = obj.find(k);
    if (type != Object || it == obj.end()) throw std::runtime_error("json: missing key " + k);
    return it->second;
  }
  const Json& at(size_t i) const {
    if (type != Array || i >= arr.size()) throw std::runtime_error("json: index out of range");
    return arr[i];
  }
  size_t size() const { return type == Array ? arr.size() : obj.size(); }
  long long as_int() const { return (long long)num; }
  long long get_int(const std::string& k, long long dflt) const { return has(k) ? at(k).as_int() : dflt; }
};

class JsonParser {
 public:
  JsonParser(const char* p, size_t n) : p_(p), e_(p + n) {}
  Json parse() {
    Json v = value();
    ws();
    return v;
  }

 private:
  const char *p_, *e_;
  void ws() { while (p_ < e_ && (*p_ == ' ' || *p_ == '\n' || *p_ == '\t' || *p_ == '\r')) p_++; }
  [[noreturn]] void fail(const char* m) { throw std::runtime_error(std::string("json: ") + m); }
  bool lit(const char* s) {
    const char* q = p_;
    while (*s) { if (q >= e_ || *q != *s) return false; q++; s++; }
    p_ = q;
    return true;
  }
  Json value() {
    ws();
    if (p_ >= e_) fail("unexpected end");
    Json v;
    const char c = *p_;
    if (c == '{') {
      v.type = Json::Object; p_++; ws();
      if (p_ < e_ && *p_ == '}') { p_++; return v; }
      while (true) {
        ws();
        Json k = string_();
        ws();
        if (p_ >= e_ || *p_ != ':') fail("expected ':'");
        p_++;
        v.obj[k.str] = value();
        ws();
        if (p_ < e_ && *p_ == ',') { p_++; continue; }
        if (p_ < e_ && *p_ == '}') { p_++; break; }
        fail("expected ',' or '}'");
      }
    } else if (c == '[') {
      v.type = Json::Array; p_++; ws();
      if (p_ < e_ && *p_ == ']') { p_++; return v; }
      while (true) {
        v.arr.push_back(value());
        ws();
        if (p_ < e_ && *p_ == ',') { p_++; continue; }
        if (p_ < e_ && *p_ == ']') { p_++; break; }
        fail("expected ',' or ']'");
      }
    } else if (c == '"') {
      v = string_();
    } else if (lit("true")) { v.type = Json::Bool; v.b = true;
    } else if (lit("false")) { v.type = Json::Bool; v.b = false;
    } else if (lit("null")) { v.type = Json::Null;
    } else {
      char* end = nullptr;
      std::string tmp(p_, (size_t)std::min<ptrdiff_t>(e_ - p_, 64));
      v.num = std::strtod(tmp.c_str(), &end);
      if (end == tmp.c_str()) fail("bad number");
      p_ += end - tmp.c_str();
      v.type = Json::Number;
    }
    return v;
  }
  Json string_() {
    if (p_ >= e_ || *p_ != '"') fail("expected string");
    p_++;
    Json v; v.type = Json::String;
    while (p_ < e_ && *p_ != '"') {
      if (*p_ == '\\' && p_ + 1 < e_) {
        p_++;
        switch (*p_) {
          case 'n': v.str += '\n'; break;
          case 't': v.str += '\t'; break;
          case 'r': v.str += '\r'; break;
          case 'b': v.str += '\b'; break;
          case 'f': v.str += '\f'; break;
          case 'u': { /* keep ASCII, replace the rest: names are only compared with .mtl keys */
            if (p_ + 4 < e_) { unsigned cp = (unsigned)std::strtoul(std::string(p_ + 1, 4).c_str(), nullptr, 16); v.str += cp < 128 ? (char)cp : '?'; p_ += 4; }
            break;
          }
          default: v.str += *p_;
        }
        p_++;
      } else {
        v.str += *p_++;
      }
    }
    if (p_ >= e_) fail("unterminated string");
    p_++;
    return v;
  }
};

}  // namespace pmh
