/* pm_io.cu — the photon wire format between the two passes.
 * Text, one photon per line: "pos.x pos.y pos.z dir.x dir.y dir.z color.x color.y color.z",
 * std::fixed with 6 decimals (writer: photon-mapping/src/hostCode.cu:31-49; reader:
 * ray-tracer/src/hostCode.cu:26-52, which parses with operator>> until the first failure). */
#include <cstdio>
#include <cerrno>
#include "pm_internal.h"

extern "C" {

int pm_photons_write_text(const char* path, const pm_photon40* p, int64_t n) {
  PM_REQUIRE(path && (p || n == 0) && n >= 0, "bad arguments");
  FILE* f = fopen(path, "w");
  if (!f) { pm::set_error("pm_photons_write_text: cannot open %s: %s", path, strerror(errno)); return PM_ERR_IO; }
  std::vector<char> buf(1 << 20);
  setvbuf(f, buf.data(), _IOFBF, buf.size());
  for (int64_t i = 0; i < n; i++) {
    const pm_photon40& q = p[i];
    /* %f of a float promoted to double == std::fixed << setprecision(6) of the float */
    fprintf(f, "%.6f %.6f %.6f %.6f %.6f %.6f %.6f %.6f %.6f\n", q.pos[0], q.pos[1], q.pos[2], q.dir[0], q.dir[1],
            q.dir[2], q.color[0], q.color[1], q.color[2]);
  }
  const bool bad = ferror(f) != 0;
  if (fclose(f) != 0 || bad) { pm::set_error("pm_photons_write_text: write to %s failed", path); return PM_ERR_IO; }
  return PM_OK;
}

int pm_photons_read_text(const char* path, pm_photon40* out, int64_t capacity, int64_t* count) {
  PM_REQUIRE(path && count && capacity >= 0, "bad arguments");
  *count = 0;
  FILE* f = fopen(path, "r");
  if (!f) { pm::set_error("pm_photons_read_text: cannot open %s: %s", path, strerror(errno)); return PM_ERR_IO; }
  std::vector<char> buf(1 << 20);
  setvbuf(f, buf.data(), _IOFBF, buf.size());
  int64_t n = 0;
  float v[9];
  while (fscanf(f, "%f %f %f %f %f %f %f %f %f", &v[0], &v[1], &v[2], &v[3], &v[4], &v[5], &v[6], &v[7], &v[8]) == 9) {
    if (out && n < capacity) {
      pm_photon40& q = out[n];
      q.pos[0] = v[0]; q.pos[1] = v[1]; q.pos[2] = v[2];
      q.dir[0] = v[3]; q.dir[1] = v[4]; q.dir[2] = v[5];
      q.power = 0;
      q.color[0] = v[6]; q.color[1] = v[7]; q.color[2] = v[8];
    }
    n++;
  }
  fclose(f);
  *count = n;
  return PM_OK;
}

/* Binary hand-off between the passes: "PMPH", version, count, then the 40-byte records
 * as they are in memory. Lossless (the text format rounds to 1e-6) and ~60x faster to parse. */
static const uint32_t kBinMagic = 0x48504D50u; /* "PMPH" little endian */

int pm_photons_write_bin(const char* path, const pm_photon40* p, int64_t n) {
  PM_REQUIRE(path && (p || n == 0) && n >= 0, "bad arguments");
  FILE* f = fopen(path, "wb");
  if (!f) { pm::set_error("pm_photons_write_bin: cannot open %s: %s", path, strerror(errno)); return PM_ERR_IO; }
  const uint32_t head[2] = {kBinMagic, 1u};
  bool ok = fwrite(head, sizeof head, 1, f) == 1 && fwrite(&n, sizeof n, 1, f) == 1;
  if (ok && n > 0) ok = fwrite(p, sizeof(pm_photon40), (size_t)n, f) == (size_t)n;
  if (fclose(f) != 0 || !ok) { pm::set_error("pm_photons_write_bin: write to %s failed", path); return PM_ERR_IO; }
  return PM_OK;
}

int pm_photons_read_bin(const char* path, pm_photon40* out, int64_t capacity, int64_t* count) {
  PM_REQUIRE(path && count && capacity >= 0, "bad arguments");
  *count = 0;
  FILE* f = fopen(path, "rb");
  if (!f) { pm::set_error("pm_photons_read_bin: cannot open %s: %s", path, strerror(errno)); return PM_ERR_IO; }
  uint32_t head[2] = {0, 0};
  int64_t n = -1;
  if (fread(head, sizeof head, 1, f) != 1 || fread(&n, sizeof n, 1, f) != 1 || head[0] != kBinMagic || head[1] != 1u || n < 0) {
    fclose(f);
    pm::set_error("pm_photons_read_bin: %s is not a photon file of this format", path);
    return PM_ERR_IO;
  }
  *count = n;
  if (out && capacity > 0) {
    const int64_t take = n < capacity ? n : capacity;
    if (take > 0 && fread(out, sizeof(pm_photon40), (size_t)take, f) != (size_t)take) {
      fclose(f);
      pm::set_error("pm_photons_read_bin: %s is truncated", path);
      return PM_ERR_IO;
    }
  }
  fclose(f);
  return PM_OK;
}

}  // extern "C"
