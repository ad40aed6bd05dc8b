// yb_ctx.h — the object behind the opaque yb200_ctx of include/yb200.h (shared by the translation units that
// define extern "C" entry points).
#pragma once
#include "yb_scene.h"

namespace yb {
// device staging of the host-pointer forms of the C ABI (ray batches, the resolved frame)
struct IoBuffers { DevBuf<char> in, out; DevBuf<float> film, rgba; };
}
struct yb200_ctx { yb::Context c; yb::IoBuffers io; };
