// pm_assets.hpp — scene import for the C++ hosts: what assets::import_scene does
// (common/src/assetImporter.cxx:16-215) for binary glTF files, without assimp:
//   * breadth-first node walk accumulating transforms (assetImporter.cxx:33-47);
//   * one Mesh per triangle primitive, vertices transformed to world space and de-duplicated per mesh
//     by exact equality in first-seen order over the face stream (:54-78) — done with a hash map
//     instead of the reference's O(V^2) std::find, same output;
//   * mesh name = material name (:88-92), used to look the material up in <model>.mtl (:139-214);
//   * lights.txt beside the model, "posx posy posz r g b power" (:98-137); missing file throws. Four more values
//     on a line ("nx ny nz side") make the light a square emitter (not in the reference).
// Paths are used as given (the reference rewrites '/' to '\\', which only works on Windows).
#pragma once
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <deque>
#include <fstream>
#include <iostream>
#include <sstream>
#include <unordered_map>
#include <utility>

#include <pmhot.h>
#include "pm_json.hpp"

namespace pmh {

struct Mesh {
  std::string name;
  std::vector<float> vertices;   // 3 per vertex
  std::vector<int32_t> indices;  // 3 per triangle, into this mesh's vertices
  pm_material material{{1.f, 1.f, 1.f}, 1.f, 0.f, 0.f, 0.f};  // white diffuse default (assetImporter.cxx:199-207)
};

struct World {
  std::vector<Mesh> meshes;
  std::vector<pm_light> light_sources;
  std::vector<pm_light_shape> light_shapes;  // one per light; lights.txt lines with 4 extra values are square emitters
  // flattened for pm_scene_create
  std::vector<float> verts;
  std::vector<int32_t> tris, tri_mesh;
  std::vector<pm_material> mats;
  void flatten() {
    verts.clear(); tris.clear(); tri_mesh.clear(); mats.clear();
    for (size_t m = 0; m < meshes.size(); m++) {
      const int32_t base = (int32_t)(verts.size() / 3);
      verts.insert(verts.end(), meshes[m].vertices.begin(), meshes[m].vertices.end());
      for (int32_t i : meshes[m].indices) tris.push_back(base + i);
      tri_mesh.insert(tri_mesh.end(), meshes[m].indices.size() / 3, (int32_t)m);
      mats.push_back(meshes[m].material);
    }
  }
};

struct Mat4 {
  float m[4][4];
  static Mat4 identity() { Mat4 r{}; for (int i = 0; i < 4; i++) r.m[i][i] = 1.f; return r; }
};
inline Mat4 mul(const Mat4& a, const Mat4& b) {
  Mat4 r{};
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      float s = 0.f;
      for (int k = 0; k < 4; k++) s += a.m[i][k] * b.m[k][j];
      r.m[i][j] = s;
    }
  return r;
}

inline Mat4 node_matrix(const Json& node) {
  Mat4 r = Mat4::identity();
  if (node.has("matrix")) {  // column-major in glTF
    const Json& a = node.at("matrix");
    for (int c = 0; c < 4; c++) for (int rr = 0; rr < 4; rr++) r.m[rr][c] = (float)a.at((size_t)(c * 4 + rr)).num;
    return r;
  }
  float t[3] = {0, 0, 0}, q[4] = {0, 0, 0, 1}, s[3] = {1, 1, 1};
  if (node.has("translation")) for (int i = 0; i < 3; i++) t[i] = (float)node.at("translation").at((size_t)i).num;
  if (node.has("rotation")) for (int i = 0; i < 4; i++) q[i] = (float)node.at("rotation").at((size_t)i).num;
  if (node.has("scale")) for (int i = 0; i < 3; i++) s[i] = (float)node.at("scale").at((size_t)i).num;
  const float x = q[0], y = q[1], z = q[2], w = q[3];
  const float R[3][3] = {{1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)},
                         {2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)},
                         {2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)}};
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) r.m[i][j] = R[i][j] * s[j];
    r.m[i][3] = t[i];
  }
  return r;
}

class GlbFile {
 public:
  explicit GlbFile(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("Unable to open file: " + path);
    blob_.assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
    if (blob_.size() < 20 || rd32(0) != 0x46546C67u) throw std::runtime_error(path + ": not a binary glTF file");
    size_t off = 12;
    while (off + 8 <= blob_.size()) {
      const uint32_t clen = rd32(off), ctype = rd32(off + 4);
      if (off + 8 + clen > blob_.size()) throw std::runtime_error(path + ": truncated chunk");
      if (ctype == 0x4E4F534Au) json = JsonParser(blob_.data() + off + 8, clen).parse();
      else if (ctype == 0x004E4942u) { bin_ = blob_.data() + off + 8; bin_len_ = clen; }
      off += 8 + (size_t)clen;
    }
    if (json.type != Json::Object) throw std::runtime_error(path + ": no JSON chunk");
  }
  Json json;

  // accessor -> doubles, `ncomp` per element
  std::vector<double> accessor(long long idx, int* ncomp) const {
    const Json& acc = json.at("accessors").at((size_t)idx);
    const Json& bv = json.at("bufferViews").at((size_t)acc.at("bufferView").as_int());
    const long long ct = acc.at("componentType").as_int();
    const std::string& ty = acc.at("type").str;
    const int nc = ty == "SCALAR" ? 1 : ty == "VEC2" ? 2 : ty == "VEC3" ? 3 : ty == "VEC4" ? 4 : ty == "MAT4" ? 16 : 0;
    const int cs = ct == 5120 || ct == 5121 ? 1 : ct == 5122 || ct == 5123 ? 2 : 4;
    if (nc == 0) throw std::runtime_error("glb: unsupported accessor type " + ty);
    const size_t off = (size_t)bv.get_int("byteOffset", 0) + (size_t)acc.get_int("byteOffset", 0);
    size_t stride = (size_t)bv.get_int("byteStride", 0);
    if (stride == 0) stride = (size_t)cs * nc;
    const size_t cnt = (size_t)acc.at("count").as_int();
    if (cnt && off + (cnt - 1) * stride + (size_t)cs * nc > bin_len_) throw std::runtime_error("glb: accessor out of range");
    std::vector<double> out(cnt * nc);
    for (size_t i = 0; i < cnt; i++)
      for (int c = 0; c < nc; c++) {
        const char* p = bin_ + off + i * stride + (size_t)c * cs;
        double v = 0;
        switch (ct) {
          case 5120: { int8_t t; memcpy(&t, p, 1); v = t; break; }
          case 5121: { uint8_t t; memcpy(&t, p, 1); v = t; break; }
          case 5122: { int16_t t; memcpy(&t, p, 2); v = t; break; }
          case 5123: { uint16_t t; memcpy(&t, p, 2); v = t; break; }
          case 5125: { uint32_t t; memcpy(&t, p, 4); v = t; break; }
          case 5126: { float t; memcpy(&t, p, 4); v = t; break; }
          default: throw std::runtime_error("glb: unsupported component type");
        }
        out[i * nc + c] = v;
      }
    *ncomp = nc;
    return out;
  }

 private:
  std::vector<char> blob_;
  const char* bin_ = nullptr;
  size_t bin_len_ = 0;
  uint32_t rd32(size_t off) const { uint32_t v; memcpy(&v, blob_.data() + off, 4); return v; }
};

struct Vec3Hash {
  size_t operator()(const std::array<uint32_t, 3>& k) const {
    uint64_t h = 1469598103934665603ull;
    for (uint32_t w : k) { h ^= w; h *= 1099511628211ull; }
    return (size_t)h;
  }
};

inline std::vector<Mesh> extract_objects(const GlbFile& glb) {
  const Json& j = glb.json;
  std::vector<Mesh> meshes;
  if (!j.has("scenes")) return meshes;
  const Json& scene = j.at("scenes").at((size_t)j.get_int("scene", 0));
  std::deque<std::pair<long long, Mat4>> queue;
  if (scene.has("nodes")) for (const Json& n : scene.at("nodes").arr) queue.emplace_back(n.as_int(), Mat4::identity());
  while (!queue.empty()) {
    const auto [ni, parent] = queue.front();
    queue.pop_front();
    const Json& node = j.at("nodes").at((size_t)ni);
    const Mat4 m = mul(parent, node_matrix(node));
    if (node.has("children")) for (const Json& c : node.at("children").arr) queue.emplace_back(c.as_int(), m);
    if (!node.has("mesh")) continue;
    for (const Json& prim : j.at("meshes").at((size_t)node.at("mesh").as_int()).at("primitives").arr) {
      if (prim.get_int("mode", 4) != 4) continue;  // aiProcess_SortByPType: only triangle lists reach the face loop
      int nc = 0;
      const std::vector<double> pos = glb.accessor(prim.at("attributes").at("POSITION").as_int(), &nc);
      if (nc != 3) throw std::runtime_error("glb: POSITION is not VEC3");
      std::vector<long long> idx;
      if (prim.has("indices")) {
        int ic = 0;
        for (double d : glb.accessor(prim.at("indices").as_int(), &ic)) idx.push_back((long long)d);
      } else {
        for (size_t i = 0; i < pos.size() / 3; i++) idx.push_back((long long)i);
      }
      Mesh out;
      std::unordered_map<std::array<uint32_t, 3>, int32_t, Vec3Hash> seen;
      for (long long vi : idx) {
        if (vi < 0 || (size_t)vi >= pos.size() / 3) throw std::runtime_error("glb: vertex index out of range");
        const float p[3] = {(float)pos[3 * vi], (float)pos[3 * vi + 1], (float)pos[3 * vi + 2]};
        float w[3];
        for (int r = 0; r < 3; r++) w[r] = ((p[0] * m.m[r][0] + p[1] * m.m[r][1]) + p[2] * m.m[r][2]) + m.m[r][3];
        for (int r = 0; r < 3; r++) w[r] = w[r] + 0.0f;  // -0 == +0 for std::find's operator==
        std::array<uint32_t, 3> key;
        memcpy(key.data(), w, 12);
        auto it = seen.find(key);
        if (it == seen.end()) {
          it = seen.emplace(key, (int32_t)(out.vertices.size() / 3)).first;
          out.vertices.insert(out.vertices.end(), w, w + 3);
        }
        out.indices.push_back(it->second);
      }
      out.indices.resize(out.indices.size() / 3 * 3);
      if (prim.has("material") && j.has("materials")) {
        const Json& mat = j.at("materials").at((size_t)prim.at("material").as_int());
        if (mat.has("name")) out.name = mat.at("name").str;
      }
      meshes.push_back(std::move(out));
    }
  }
  return meshes;
}

inline std::vector<pm_light> extract_lights(const std::string& model_path, std::vector<pm_light_shape>* shapes = nullptr) {
  const size_t slash = model_path.find_last_of("/\\");
  const std::string file = (slash == std::string::npos ? std::string(".") : model_path.substr(0, slash)) + "/lights.txt";
  std::ifstream f(file);
  if (!f) throw std::runtime_error("Unable to open file: " + file);  // assetImporter.cxx:108-110
  std::vector<pm_light> out;
  std::string line;
  while (std::getline(f, line)) {
    if (!line.empty() && line.back() == '\r') line.pop_back();
    if (line.empty() || line[0] == '#') continue;
    std::istringstream iss(line);
    double v[7];
    for (double& d : v) if (!(iss >> d)) throw std::runtime_error("Invalid light source data format");  // :124-128
    pm_light l{{(float)v[0], (float)v[1], (float)v[2]}, {(float)v[3], (float)v[4], (float)v[5]}, v[6]};
    out.push_back(l);
    // extension (world.h:17-20 fields the reference never fills): "... power nx ny nz side" = square emitter
    pm_light_shape sh{PM_LIGHT_POINT, {0.f, 0.f, 0.f}, 0.f};
    double e[4];
    if (iss >> e[0] >> e[1] >> e[2] >> e[3]) sh = pm_light_shape{PM_LIGHT_SQUARE, {(float)e[0], (float)e[1], (float)e[2]}, (float)e[3]};
    if (shapes) shapes->push_back(sh);
  }
  return out;
}

inline void assign_materials(std::vector<Mesh>& meshes, const std::string& model_path) {
  const size_t dot = model_path.find_last_of('.');
  const std::string file = model_path.substr(0, dot) + ".mtl";
  std::unordered_map<std::string, pm_material> table;
  std::ifstream f(file);
  if (!f) {
    std::cerr << "Error: Unable to open file " << file << std::endl;  // warns only (assetImporter.cxx:149-152)
  } else {
    std::string line;
    while (std::getline(f, line)) {
      if (!line.empty() && line.back() == '\r') line.pop_back();
      if (line.empty() || line[0] == '#') continue;
      std::istringstream iss(line);
      std::string name;
      float a[7];
      bool ok = (bool)(iss >> name);
      for (float& x : a) ok = ok && (bool)(iss >> x);
      if (!ok) { std::cerr << "Warning: Invalid line format: " << line << std::endl; continue; }
      table[name] = pm_material{{a[0], a[1], a[2]}, a[3], a[4], a[5], a[6]};
    }
  }
  for (Mesh& m : meshes) {
    auto it = table.find(m.name);
    if (it != table.end()) m.material = it->second;
  }
}

// ≙ assets::import_scene (common/src/assetImporter.h:8)
inline World import_scene(const std::string& path) {
  World w;
  GlbFile glb(path);
  w.meshes = extract_objects(glb);
  w.light_sources = extract_lights(path, &w.light_shapes);
  assign_materials(w.meshes, path);
  w.flatten();
  return w;
}

}  // namespace pmh
