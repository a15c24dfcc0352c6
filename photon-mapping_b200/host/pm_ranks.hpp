// pm_ranks.hpp — multi-process mode of the shipped hosts: one process per GPU, no launcher library.
//   <host> config.toml --rank R --world G --comm-file PATH [--device D]
// Rank 0 draws the NCCL id (pm_comm_unique_id) and publishes its 128 bytes in PATH (written to PATH.tmp, then renamed);
// the other ranks poll for the file. Any process starter works: a shell loop, mpirun, srun. The reference has no
// counterpart (owlContextCreate(nullptr, 1), photon-mapping/src/hostCode.cu:145).
#pragma once
#include <chrono>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>

#include "pmhot.h"

namespace pmh {

struct Ranks {
  int rank = 0, world = 1, device = -1;
  std::string comm_file;
  pm_comm* comm = nullptr;

  // consumes --rank / --world / --comm-file / --device from argv[i]; returns true if it did
  bool parse(int argc, char** argv, int& i) {
    auto need = [&](const char* what) { if (i + 1 >= argc) throw std::runtime_error(std::string(what) + " needs a value"); return argv[++i]; };
    if (!strcmp(argv[i], "--rank")) { rank = atoi(need("--rank")); return true; }
    if (!strcmp(argv[i], "--world")) { world = atoi(need("--world")); return true; }
    if (!strcmp(argv[i], "--device")) { device = atoi(need("--device")); return true; }
    if (!strcmp(argv[i], "--comm-file")) { comm_file = need("--comm-file"); return true; }
    return false;
  }

  int pick_device() const {
    if (device >= 0) return device;
    const int n = pm_device_count();
    return n > 0 ? rank % n : 0;
  }

  void connect(pm_ctx* ctx) {
    if (world <= 1) return;
    if (rank < 0 || rank >= world) throw std::runtime_error("--rank must be in [0, --world)");
    if (comm_file.empty()) throw std::runtime_error("--world > 1 needs --comm-file PATH (a path every rank can read)");
    unsigned char id[PM_COMM_ID_BYTES];
    if (rank == 0) {
      if (pm_comm_unique_id(id, (int)sizeof id) != PM_OK) throw std::runtime_error(std::string("libpmhot: ") + pm_last_error());
      const std::string tmp = comm_file + ".tmp";
      FILE* f = fopen(tmp.c_str(), "wb");
      if (!f || fwrite(id, 1, sizeof id, f) != sizeof id || fclose(f) != 0) throw std::runtime_error("cannot write " + tmp);
      if (rename(tmp.c_str(), comm_file.c_str()) != 0) throw std::runtime_error("cannot publish " + comm_file);
    } else {
      bool got = false;
      for (int tries = 0; tries < 1200 && !got; tries++) {  // up to two minutes
        if (FILE* f = fopen(comm_file.c_str(), "rb")) {
          got = fread(id, 1, sizeof id, f) == sizeof id;
          fclose(f);
        }
        if (!got) std::this_thread::sleep_for(std::chrono::milliseconds(100));
      }
      if (!got) throw std::runtime_error("rank 0 never published " + comm_file);
    }
    if (pm_comm_create(ctx, id, rank, world, &comm) != PM_OK) throw std::runtime_error(std::string("libpmhot: ") + pm_last_error());
  }

  void close() {
    if (comm) pm_comm_destroy(comm);
    comm = nullptr;
    if (world > 1 && rank == 0) remove(comm_file.c_str());
  }
};

// device buffer owned by a host that links libpmhot only
struct DevBuf {
  pm_ctx* ctx = nullptr;
  void* p = nullptr;
  DevBuf(pm_ctx* c, int64_t bytes) : ctx(c) {
    if (pm_dev_alloc(c, bytes, &p) != PM_OK) throw std::runtime_error(std::string("libpmhot: ") + pm_last_error());
  }
  ~DevBuf() { pm_dev_free(ctx, p); }
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  template <class T> T* as() const { return static_cast<T*>(p); }
};

}  // namespace pmh
