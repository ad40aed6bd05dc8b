// yb_shade_sd_path.cu — build 0 of the shade kernel for waves of path vertices in scenes whose materials are all shinydiffuse
// (see yb_shade.inc: YB_SHADE_KIND, YB_FIX)
#include "yb_shade_common.cuh"
namespace yb {
#define YB_SHADE_NS shade_sd_path
#define YB_EXT 0
#define YB_SHADE_KIND 1
#define YB_FIX MAT_SHINYDIFFUSE
#include "yb_shade.inc"
void launch_shade_sd_path(uint32_t n, const RenderArgs &a, rt_stream s) { shade_sd_path::k_shade(n, a, s); }
void launch_shade_sd_int(uint32_t n, const RenderArgs &a, rt_stream s);
void launch_shade_sd_kind(uint32_t kind, uint32_t n, const RenderArgs &a, rt_stream s)
{
    if (kind == TASK_INTEGRATE) launch_shade_sd_int(n, a, s); else launch_shade_sd_path(n, a, s);
}
}  // namespace yb
