// photonViewer — the photon splat host (drop-in for photon-viewer/src/hostCode.cu:115-177).
// Reads the same config.toml keys ([camera], [data], [photon-viewer]), the scene and the two photon files written by
// photonMapping, and writes one PNG per photon file: every photon visible from the camera colours its pixel
// (≙ loadPhotons + photonViewerRayGen, run by pm_view_photons as CUDA kernels — no graphics API).
//
//   photonViewer [config.toml]
#include <chrono>
#include <cstring>
#include <iostream>

#include "pm_assets.hpp"
#include "pm_config.hpp"
#include "pm_png.hpp"

#define LOG(message) std::cout << "\033[1;37m#pm.host: " << message << "\033[0m" << std::endl
#define LOG_OK(message) std::cout << "\033[1;32m#pm.host: " << message << "\033[0m" << std::endl

static void check(int rc) {
  if (rc != PM_OK) throw std::runtime_error(std::string("libpmhot: ") + pm_last_error());
}

static std::vector<pm_photon40> read_photons(const std::string& path) {
  int64_t n = 0;
  const bool bin = path.size() > 4 && path.compare(path.size() - 4, 4, ".bin") == 0;
  auto rd = bin ? pm_photons_read_bin : pm_photons_read_text;
  check(rd(path.c_str(), nullptr, 0, &n));
  std::vector<pm_photon40> v((size_t)n);
  if (n) check(rd(path.c_str(), v.data(), n, &n));
  return v;
}

// run (photon-viewer/src/hostCode.cu:115-155): one image per photon file
static void run(const pmh::Config& cfg, pm_ctx* ctx, pm_scene* scene, const std::string& photons_file, const std::string& output) {
  const auto at = cfg.array("camera", "look_at", 3), from = cfg.array("camera", "look_from", 3), up = cfg.array("camera", "look_up", 3);
  const float fovy = (float)cfg.num("camera", "fovy");
  const auto fb_size = cfg.array("photon-viewer", "fb_size", 2);
  const int W = (int)fb_size[0], H = (int)fb_size[1];
  const float f3[3] = {(float)from[0], (float)from[1], (float)from[2]}, a3[3] = {(float)at[0], (float)at[1], (float)at[2]},
              u3[3] = {(float)up[0], (float)up[1], (float)up[2]};
  const std::vector<pm_photon40> photons = read_photons(photons_file);
  std::vector<uint32_t> fb((size_t)W * H);
  const auto t0 = std::chrono::steady_clock::now();
  check(pm_view_photons(ctx, scene, photons.data(), (int64_t)photons.size(), f3, a3, u3, fovy, W, H, fb.data()));
  const auto t1 = std::chrono::steady_clock::now();
  printf("Splatted %d photons in %.3f ms (kernels %.3f ms)\n", (int)photons.size(),
         std::chrono::duration<double, std::milli>(t1 - t0).count(), pm_ctx_last_kernel_ms(ctx));
  if (!pmh::write_png_rgba8(output, fb.data(), W, H)) throw std::runtime_error("cannot write " + output);
}

int main(int argc, char** argv) {
  try {
    const std::string config_path = argc > 1 ? argv[1] : "../config.toml";
    LOG("Loading Config file...");
    const pmh::Config cfg = pmh::Config::load(config_path);
    const std::string photons_filename = cfg.str("data", "photons_file");
    const std::string caustics_photons_filename = cfg.str("data", "caustics_photons_file");
    const std::string output_filename = cfg.str("photon-viewer", "output_filename");
    const std::string caustics_output_filename = cfg.str("photon-viewer", "caustics_output_filename");
    pmh::World world = pmh::import_scene(cfg.str("data", "model_path"));
    LOG_OK("Loaded world.");
    pm_ctx* ctx = nullptr;
    pm_scene* scene = nullptr;
    check(pm_ctx_create(0, &ctx));
    check(pm_scene_create(ctx, world.verts.data(), (int)(world.verts.size() / 3), world.tris.data(), (int)(world.tris.size() / 3),
                          world.tri_mesh.data(), world.mats.data(), (int)world.mats.size(), &scene));
    LOG("Running photon viewer...");
    run(cfg, ctx, scene, photons_filename, output_filename);
    LOG_OK("Done with photon viewer.");
    LOG("Running caustics viewer...");
    run(cfg, ctx, scene, caustics_photons_filename, caustics_output_filename);
    LOG_OK("Done with caustics viewer.");
    pm_scene_destroy(scene);
    pm_ctx_destroy(ctx);
    return 0;
  } catch (const std::exception& e) {
    std::cerr << "photonViewer: " << e.what() << std::endl;
    return 2;
  }
}
