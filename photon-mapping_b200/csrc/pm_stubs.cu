/* temporary: entry points not implemented yet */
#include "pm_internal.h"
#define STUB(name) { pm::set_error(#name ": not implemented"); return PM_ERR_UNSUPPORTED; }
extern "C" {
int pm_render(pm_ctx*, const pm_scene*, const pm_light*, int, const pm_map*, const pm_map*, const pm_camera*, int, int, int, int, const float*, int64_t, int64_t, uint32_t*) STUB(pm_render)
}
