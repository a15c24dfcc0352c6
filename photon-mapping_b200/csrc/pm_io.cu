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

}  // extern "C"
