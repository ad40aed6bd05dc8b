// yb_shade_sd_int.cu — build 0 of the shade kernel for waves of integrate() tasks in scenes whose materials are all shinydiffuse
// (see yb_shade.inc: YB_SHADE_KIND, YB_FIX)
#include "yb_shade_common.cuh"
namespace yb {
#define YB_SHADE_NS shade_sd_int
#define YB_EXT 0
#define YB_SHADE_KIND 0
#define YB_FIX MAT_SHINYDIFFUSE
#include "yb_shade.inc"
void launch_shade_sd_int(uint32_t n, const RenderArgs &a, rt_stream s) { shade_sd_int::k_shade(n, a, s); }
}  // namespace yb
