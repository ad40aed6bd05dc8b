// yb_shade_ext.cu — build 1 of the shade kernel (see yb_shade.inc for what each build adds)
#include "yb_shade_common.cuh"
namespace yb {
#define YB_SHADE_NS shade_ext
#define YB_EXT 1
#include "yb_shade.inc"
void launch_shade_ext(uint32_t n, const RenderArgs &a, rt_stream s) { shade_ext::k_shade(n, a, s); }
}  // namespace yb
