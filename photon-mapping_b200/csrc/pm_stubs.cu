/* temporary: entry points not implemented yet */
#include "pm_internal.h"
#define STUB(name) { pm::set_error(#name ": not implemented"); return PM_ERR_UNSUPPORTED; }
extern "C" {
int pm_render(pm_ctx*, const pm_scene*, const pm_light*, int, const pm_map*, const pm_map*, const pm_camera*, int, int, int, int, const float*, int64_t, int64_t, uint32_t*) STUB(pm_render)
int pm_scene_create(pm_ctx*, const float*, int, const int32_t*, int, const int32_t*, const pm_material*, int, pm_scene**) STUB(pm_scene_create)
void pm_scene_destroy(pm_scene*) {}
int pm_trace_closest(pm_ctx*, const pm_scene*, const pm_ray*, int64_t, pm_hit*) STUB(pm_trace_closest)
int pm_trace_occluded(pm_ctx*, const pm_scene*, const pm_ray*, int64_t, uint8_t*) STUB(pm_trace_occluded)
int pm_emit_photons(pm_ctx*, const pm_scene*, const pm_light*, int, int64_t, int, int, uint64_t, int, int64_t, int64_t, pm_photon40*, int64_t, int64_t*, int64_t*, int64_t*) STUB(pm_emit_photons)
int pm_photons_to_maps(pm_ctx*, const pm_photon40*, int64_t, const pm_photon40*, int64_t, pm_photon44*, pm_photon44*) STUB(pm_photons_to_maps)
}
