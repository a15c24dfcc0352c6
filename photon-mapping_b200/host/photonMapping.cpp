// photonMapping — the photon pass host (drop-in for photon-mapping/src/hostCode.cu:140-188).
// Same inputs (config.toml keys [data] / [photon-mapper], the glb scene, lights.txt, <model>.mtl) and the
// same outputs (two photon text files); the device work goes through libpmhot's C ABI instead of OWL.
//
//   photonMapping [config.toml] [--dump-scene] [--counter-rng SEED] [--rank R --world G --comm-file PATH [--device D]]
// Without arguments the config is read from ../config.toml like the reference (common/src/configLoader.h:6).
// With --world G (one process per GPU, pm_ranks.hpp) rank R traces launch indices R, R + G, ... of every light, the
// shards meet in ONE exchange (pm_exchange_photons over NCCL) and rank 0 writes the same files a single process writes.
#include <chrono>
#include <cstdio>
#include <cstring>
#include <iostream>

#include "pm_assets.hpp"
#include "pm_config.hpp"
#include "pm_ranks.hpp"

#define LOG(message) std::cout << "\033[1;37m#pm.host: " << message << "\033[0m" << std::endl
#define LOG_OK(message) std::cout << "\033[1;32m#pm.host: " << message << "\033[0m" << std::endl

static void check(int rc) {
  if (rc != PM_OK) throw std::runtime_error(std::string("libpmhot: ") + pm_last_error());
}

int main(int argc, char** argv) {
  try {
    std::string config_path = "../config.toml";
    bool dump = false;
    int rng_mode = PM_RNG_LCG_PARITY;
    uint64_t seed = 0;
    pmh::Ranks ranks;
    for (int i = 1; i < argc; i++) {
      if (ranks.parse(argc, argv, i)) continue;
      if (!strcmp(argv[i], "--dump-scene")) dump = true;
      else if (!strcmp(argv[i], "--counter-rng") && i + 1 < argc) { rng_mode = PM_RNG_COUNTER; seed = strtoull(argv[++i], nullptr, 10); }
      else config_path = argv[i];
    }
    LOG("Loading Config file...");
    const pmh::Config cfg = pmh::Config::load(config_path);
    const std::string photons_filename = cfg.str("data", "photons_file");
    const std::string caustics_photons_filename = cfg.str("data", "caustics_photons_file");
    const std::string model_path = cfg.str("data", "model_path");
    const long long casted_diffuse = cfg.integer("photon-mapper", "casted_diffuse_photons");
    const long long casted_caustics = cfg.integer("photon-mapper", "casted_caustics_photons");
    const int max_depth = (int)cfg.integer("photon-mapper", "max_depth");

    LOG("Loading Model...");
    pmh::World world = pmh::import_scene(model_path);
    LOG_OK("Loaded world: " << world.meshes.size() << " meshes, " << world.tris.size() / 3 << " triangles, "
                            << world.light_sources.size() << " lights");
    if (dump) {  // machine-readable scene summary: used by the CPU tests, needs no GPU
      printf("meshes %zu\n", world.meshes.size());
      for (const pmh::Mesh& m : world.meshes) {
        printf("mesh \"%s\" %zu %zu mat %.9g %.9g %.9g %.9g %.9g %.9g %.9g\n", m.name.c_str(), m.vertices.size() / 3,
               m.indices.size() / 3, m.material.albedo[0], m.material.albedo[1], m.material.albedo[2], m.material.diffuse,
               m.material.specular, m.material.transmission, m.material.refraction_idx);
        for (size_t v = 0; v < m.vertices.size(); v += 3) printf("v %.9g %.9g %.9g\n", m.vertices[v], m.vertices[v + 1], m.vertices[v + 2]);
        for (size_t t = 0; t < m.indices.size(); t += 3) printf("f %d %d %d\n", m.indices[t], m.indices[t + 1], m.indices[t + 2]);
      }
      for (const pm_light& l : world.light_sources)
        printf("light %.9g %.9g %.9g %.9g %.9g %.9g %.17g\n", l.pos[0], l.pos[1], l.pos[2], l.rgb[0], l.rgb[1], l.rgb[2], l.power);
      printf("photon-mapper %lld %lld %d\n", casted_diffuse, casted_caustics, max_depth);
      return 0;
    }

    pm_ctx* ctx = nullptr;
    pm_scene* scene = nullptr;
    check(pm_ctx_create(ranks.pick_device(), &ctx));
    ranks.connect(ctx);
    check(pm_scene_create(ctx, world.verts.data(), (int)(world.verts.size() / 3), world.tris.data(), (int)(world.tris.size() / 3),
                          world.tri_mesh.data(), world.mats.data(), (int)world.mats.size(), &scene));
    LOG_OK("Geometry uploaded, LBVH built");

    // runNormal then runCaustics (hostCode.cu:112-138); buffers sized casted * maxDepth (hostCode.cu:93,97)
    struct Pass { const char* what; long long casted; int caustics; const std::string* file; };
    const Pass passes[2] = {{"global", casted_diffuse, 0, &photons_filename}, {"caustic", casted_caustics, 1, &caustics_photons_filename}};
    for (const Pass& p : passes) {
      const int64_t capacity = (int64_t)std::max<long long>(p.casted, 0) * (int64_t)std::max(max_depth, 1);
      std::vector<pm_photon40> photons;
      int64_t count = 0, paths = 0, segments = 0;
      const auto t0 = std::chrono::steady_clock::now();
      if (ranks.world == 1) {
        photons.resize((size_t)capacity);
        check(pm_emit_photons_shaped(ctx, scene, world.light_sources.data(), world.light_shapes.data(), (int)world.light_sources.size(),
                                     p.casted, max_depth, p.caustics, seed, rng_mode, 0, 1, photons.data(), nullptr, capacity, &count,
                                     &paths, &segments));
      } else {
        // this rank's launch indices -> device shard + one count byte per path -> the exchange -> all photons, in the
        // single-process order, as 44-byte records on every rank; rank 0 turns them back into the file's 40-byte records
        const int64_t cap_local = capacity / ranks.world + (int64_t)max_depth * ((int64_t)world.light_sources.size() + 1);
        const int64_t paths_max = p.casted / ranks.world + (int64_t)world.light_sources.size() + 16;
        pmh::DevBuf shard(ctx, cap_local * (int64_t)sizeof(pm_photon40)), counts(ctx, paths_max), all(ctx, capacity * (int64_t)sizeof(pm_photon44));
        int64_t stored = 0;
        check(pm_emit_photons_counted(ctx, scene, world.light_sources.data(), world.light_shapes.data(), (int)world.light_sources.size(),
                                      p.casted, max_depth, p.caustics, seed, rng_mode, ranks.rank, ranks.world, shard.as<pm_photon40>(),
                                      counts.as<uint8_t>(), cap_local, &stored, &paths, &segments));
        float coll_ms = 0.f;
        check(pm_exchange_photons(ctx, ranks.comm, world.light_sources.data(), (int)world.light_sources.size(), p.casted,
                                  shard.as<pm_photon40>(), stored, counts.as<uint8_t>(), 1.0f, all.as<pm_photon44>(), capacity, &count,
                                  &coll_ms));
        if (ranks.rank == 0) {
          std::vector<pm_photon44> rec((size_t)count);
          check(pm_dev_copy(ctx, rec.data(), all.p, count * (int64_t)sizeof(pm_photon44)));
          photons.resize((size_t)count);
          for (int64_t i = 0; i < count; i++) {
            pm_photon40& q = photons[(size_t)i];
            memcpy(q.pos, rec[(size_t)i].pos, 12); memcpy(q.dir, rec[(size_t)i].dir, 12); memcpy(q.color, rec[(size_t)i].color, 12);
            q.power = 0;
          }
        }
      }
      const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
      LOG_OK(p.what << " pass" << (ranks.world > 1 ? " (rank " + std::to_string(ranks.rank) + " of " + std::to_string(ranks.world) + ")" : std::string())
                    << ": " << paths << " paths, " << segments << " segments, " << count << " photons stored in " << ms
                    << " ms (device " << pm_ctx_last_kernel_ms(ctx) << " ms)");
      if (ranks.rank == 0) {
        const bool bin = p.file->size() > 4 && p.file->compare(p.file->size() - 4, 4, ".bin") == 0;  // lossless hand-off
        check(bin ? pm_photons_write_bin(p.file->c_str(), photons.data(), count) : pm_photons_write_text(p.file->c_str(), photons.data(), count));
        LOG_OK("Wrote " << *p.file);
      }
    }
    ranks.close();
    pm_scene_destroy(scene);
    pm_ctx_destroy(ctx);
    return 0;
  } catch (const std::exception& e) {
    std::cerr << "photonMapping: " << e.what() << std::endl;
    return 2;
  }
}
