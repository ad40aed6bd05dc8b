// yb_shade_staged.cu — the base build of the shade kernel (no rough_glass / blend_mat / mesh lights / shader nodes) with direct
// lighting taken out: estimate{All,One}DirectLight only queue (task, light) jobs, which the kernels of yb_direct.cu evaluate
// group by group.  Used for scenes whose tasks fan out (glossy / specular recursion, several paths per sample); see yb_direct.cu.
#include "yb_shade_common.cuh"
namespace yb {
#define YB_SHADE_NS shade_staged
#define YB_EXT 0
#define YB_STAGE_DIRECT 1
#include "yb_shade.inc"
void launch_shade_staged(uint32_t n, const RenderArgs &a, rt_stream s) { shade_staged::k_shade(n, a, s); }
}  // namespace yb
