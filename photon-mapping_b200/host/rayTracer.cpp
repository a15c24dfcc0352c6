// rayTracer — the render pass host (drop-in for ray-tracer/src/hostCode.cu:180-246).
// Reads the same config.toml keys ([camera], [data], [ray-tracer]), the scene, and the two photon text
// files written by photonMapping; builds the two photon maps (≙ cukd::buildTree ×2, hostCode.cu:93-98),
// renders through libpmhot (≙ owlRayGenLaunch2D, :231-237) and writes the PNG (:239-240).
//
//   rayTracer [config.toml] [--rank R --world G --comm-file PATH [--device D]]
// With --world G (one process per GPU, pm_ranks.hpp) every rank reads the photon files, the global map is built with
// pm_map_build_sharded (same tree, levels below log2(G) split over the ranks), rank R renders pixels
// [R*W*H/G, (R+1)*W*H/G) and the disjoint ranges meet in one all-reduce; rank 0 writes the PNG.
#include <chrono>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <memory>

#include "pm_assets.hpp"
#include "pm_config.hpp"
#include "pm_png.hpp"
#include "pm_ranks.hpp"

#define LOG(message) std::cout << "\033[1;37m#pm.host: " << message << "\033[0m" << std::endl
#define LOG_OK(message) std::cout << "\033[1;32m#pm.host: " << message << "\033[0m" << std::endl

static void check(int rc) {
  if (rc != PM_OK) throw std::runtime_error(std::string("libpmhot: ") + pm_last_error());
}

struct V3 { float x, y, z; };
static V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
// owl/common/math/vec.h:330-356: cross, dot = (x*x + y*y) + z*z, normalize = v * (1 / sqrt(dot))
static V3 cross(V3 a, V3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
static V3 normalize(V3 v) { return (1.f / std::sqrt(v.x * v.x + v.y * v.y + v.z * v.z)) * v; }

// setupCamera (ray-tracer/src/hostCode.cu:100-108): cos(fovy) scales the image plane
static pm_camera setup_camera(V3 from, V3 at, V3 up, float fovy, int W, int H) {
  const float aspect = W / static_cast<float>(H);
  const float cosFovy = std::cos(fovy);
  V3 d00 = normalize(at - from);
  const V3 du = (cosFovy * aspect) * normalize(cross(d00, up));
  const V3 dv = cosFovy * normalize(cross(du, d00));
  d00 = d00 - 0.5f * (du + dv);
  pm_camera c{{from.x, from.y, from.z}, {d00.x, d00.y, d00.z}, {du.x, du.y, du.z}, {dv.x, dv.y, dv.z}};
  return c;
}

// readPhotonsFromFile (ray-tracer/src/hostCode.cu:26-52): an unreadable file is reported and yields no photons — the
// render goes on with direct light only. One pass over the file: a text line holds nine "%.6f" numbers, so it is at
// least 81 bytes long and size / 64 + 16 records always suffice.
static std::vector<pm_photon40> read_photons(const std::string& path) {
  const bool bin = path.size() > 4 && path.compare(path.size() - 4, 4, ".bin") == 0;  // written by photonMapping for *.bin names
  int64_t n = 0, cap = 0;
  if (bin) {
    if (pm_photons_read_bin(path.c_str(), nullptr, 0, &cap) != PM_OK) {
      std::cerr << "Error reading photons file " << path << ": " << pm_last_error() << std::endl;
      return {};
    }
  } else {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) {
      std::cerr << "Error opening photons file " << path << std::endl;
      return {};
    }
    std::fseek(f, 0, SEEK_END);
    cap = (int64_t)std::ftell(f) / 64 + 16;
    std::fclose(f);
  }
  std::vector<pm_photon40> v((size_t)cap);
  if ((bin ? pm_photons_read_bin : pm_photons_read_text)(path.c_str(), v.data(), cap, &n) != PM_OK) {
    std::cerr << "Error reading photons file " << path << ": " << pm_last_error() << std::endl;
    return {};
  }
  v.resize((size_t)std::min(n, cap));
  return v;
}

int main(int argc, char** argv) {
  try {
    std::string config_path = "../config.toml";
    pmh::Ranks ranks;
    for (int i = 1; i < argc; i++) {
      if (ranks.parse(argc, argv, i)) continue;
      config_path = argv[i];
    }
    LOG("Loading Config file...");
    const pmh::Config cfg = pmh::Config::load(config_path);
    const std::string global_file = cfg.str("data", "photons_file");
    const std::string caustic_file = cfg.str("data", "caustics_photons_file");
    const std::string model_path = cfg.str("data", "model_path");
    const auto at = cfg.array("camera", "look_at", 3), from = cfg.array("camera", "look_from", 3), up = cfg.array("camera", "look_up", 3);
    const float fovy = (float)cfg.num("camera", "fovy");
    const auto sky = cfg.array("ray-tracer", "sky_colour", 3);
    const std::string output = cfg.str("ray-tracer", "output_filename");
    const auto fb_size = cfg.array("ray-tracer", "fb_size", 2);
    const int W = (int)fb_size[0], H = (int)fb_size[1];
    const int spp = (int)cfg.integer("ray-tracer", "samples_per_pixel");
    const int depth = (int)cfg.integer("ray-tracer", "depth");

    LOG("Loading Model...");
    pmh::World world = pmh::import_scene(model_path);
    pm_ctx* ctx = nullptr;
    pm_scene* scene = nullptr;
    check(pm_ctx_create(ranks.pick_device(), &ctx));
    ranks.connect(ctx);
    check(pm_scene_create(ctx, world.verts.data(), (int)(world.verts.size() / 3), world.tris.data(), (int)(world.tris.size() / 3),
                          world.tri_mesh.data(), world.mats.data(), (int)world.mats.size(), &scene));

    // loadPhotons (hostCode.cu:54-99)
    const std::vector<pm_photon40> g40 = read_photons(global_file), c40 = read_photons(caustic_file);
    printf("Loaded %d photons and %d caustic photons\n", (int)g40.size(), (int)c40.size());
    std::vector<pm_photon44> g44(g40.size() + c40.size()), c44(c40.size());
    check(pm_photons_to_maps(ctx, g40.data(), (int64_t)g40.size(), c40.data(), (int64_t)c40.size(), g44.data(), c44.data()));
    pm_map *gmap = nullptr, *cmap = nullptr;
    float gb[6], cb[6];
    std::unique_ptr<pmh::DevBuf> g_dev, fb_dev;
    const auto t0 = std::chrono::steady_clock::now();
    if (ranks.world == 1 || g44.empty()) {
      check(pm_map_build(ctx, g44.data(), (int64_t)g44.size(), gb, &gmap));
    } else {
      g_dev.reset(new pmh::DevBuf(ctx, (int64_t)(g44.size() * sizeof(pm_photon44))));
      check(pm_dev_copy(ctx, g_dev->p, g44.data(), (int64_t)(g44.size() * sizeof(pm_photon44))));
      check(pm_map_build_sharded(ctx, ranks.comm, g_dev->as<pm_photon44>(), (int64_t)g44.size(), gb, &gmap));
    }
    check(pm_map_build(ctx, c44.data(), (int64_t)c44.size(), cb, &cmap));  // the caustic map is small: built whole
    const auto t1 = std::chrono::steady_clock::now();
    printf("Photon maps built in %.3f ms\n", std::chrono::duration<double, std::milli>(t1 - t0).count());

    const pm_camera cam = setup_camera({(float)from[0], (float)from[1], (float)from[2]}, {(float)at[0], (float)at[1], (float)at[2]},
                                       {(float)up[0], (float)up[1], (float)up[2]}, fovy, W, H);
    const float sky3[3] = {(float)sky[0], (float)sky[1], (float)sky[2]};
    std::vector<uint32_t> fb((size_t)W * H, 0u);
    LOG("Launching...");
    const auto t2 = std::chrono::steady_clock::now();
    const int64_t npx = (int64_t)W * H;
    if (ranks.world == 1) {
      check(pm_render(ctx, scene, world.light_sources.data(), (int)world.light_sources.size(), gmap, cmap, &cam, W, H, spp, depth, sky3,
                      0, npx, fb.data()));
    } else {
      fb_dev.reset(new pmh::DevBuf(ctx, npx * (int64_t)sizeof(uint32_t)));  // zero-filled
      const int64_t b = ranks.rank * npx / ranks.world, e = (ranks.rank + 1) * npx / ranks.world;
      check(pm_render(ctx, scene, world.light_sources.data(), (int)world.light_sources.size(), gmap, cmap, &cam, W, H, spp, depth, sky3,
                      b, e, fb_dev->as<uint32_t>()));
      check(pm_comm_allreduce_sum(ctx, ranks.comm, fb_dev->p, npx, PM_DTYPE_I32));  // disjoint ranges: the sum is a merge
      check(pm_dev_copy(ctx, fb.data(), fb_dev->p, npx * (int64_t)sizeof(uint32_t)));
    }
    const auto t3 = std::chrono::steady_clock::now();
    printf("Rendered in %.3f ms (gather kernels %.3f ms)\n", std::chrono::duration<double, std::milli>(t3 - t2).count(),
           pm_ctx_last_kernel_ms(ctx));
    if (ranks.rank == 0) {
      if (!pmh::write_png_rgba8(output, fb.data(), W, H)) throw std::runtime_error("cannot write " + output);
      LOG_OK("Written rendered frame buffer to file " << output);
    }
    ranks.close();
    pm_map_destroy(gmap);
    pm_map_destroy(cmap);
    fb_dev.reset();  // device buffers go before their context
    g_dev.reset();
    pm_scene_destroy(scene);
    pm_ctx_destroy(ctx);
    return 0;
  } catch (const std::exception& e) {
    std::cerr << "rayTracer: " << e.what() << std::endl;
    return 2;
  }
}
