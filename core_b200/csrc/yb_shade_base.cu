// yb_shade_base.cu — build 0 of the shade kernel (see yb_shade.inc for what each build adds)
#include "yb_shade_common.cuh"
namespace yb {
#define YB_SHADE_NS shade_base
#define YB_EXT 0
#include "yb_shade.inc"
void launch_shade_base(uint32_t n, const RenderArgs &a, rt_stream s) { shade_base::k_shade(n, a, s); }
}  // namespace yb
