// yb_shade_base_path.cu — build 0 of the shade kernel for waves that hold path vertices only (see yb_shade.inc, YB_SHADE_KIND)
#include "yb_shade_common.cuh"
namespace yb {
#define YB_SHADE_NS shade_base_path
#define YB_EXT 0
#define YB_SHADE_KIND 1
#include "yb_shade.inc"
void launch_shade_base_path(uint32_t n, const RenderArgs &a, rt_stream s) { shade_base_path::k_shade(n, a, s); }
void launch_shade_base_int(uint32_t n, const RenderArgs &a, rt_stream s);
void launch_shade_base_kind(uint32_t kind, uint32_t n, const RenderArgs &a, rt_stream s)
{
    if (kind == TASK_INTEGRATE) launch_shade_base_int(n, a, s); else launch_shade_base_path(n, a, s);
}
}  // namespace yb
