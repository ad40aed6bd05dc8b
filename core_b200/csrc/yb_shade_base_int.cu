// yb_shade_base_int.cu — build 0 of the shade kernel for waves that hold integrate() tasks only (see yb_shade.inc, YB_SHADE_KIND)
#include "yb_shade_common.cuh"
namespace yb {
#define YB_SHADE_NS shade_base_int
#define YB_EXT 0
#define YB_SHADE_KIND 0
#include "yb_shade.inc"
void launch_shade_base_int(uint32_t n, const RenderArgs &a, rt_stream s) { shade_base_int::k_shade(n, a, s); }
}  // namespace yb
