// yb_shade_nodes.cu — build 2 of the shade kernel (see yb_shade.inc for what each build adds)
#include "yb_shade_common.cuh"
namespace yb {
#define YB_SHADE_NS shade_nodes
#define YB_EXT 2
#include "yb_shade.inc"
void launch_shade_nodes(uint32_t n, const RenderArgs &a, rt_stream s) { shade_nodes::k_shade(n, a, s); }
}  // namespace yb
