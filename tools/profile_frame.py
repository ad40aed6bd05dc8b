"""One frame of the bench workload at reduced spp for `ncu` (BVH build + 2 batches of ~4 M camera samples):
    python tools/profile_frame.py                                   # must exit 0 on its own first
    ncu --set full --import-source on --clock-control none --profile-from-start off \
        -o gpurun_out/prof python tools/profile_frame.py
The first (untimed) frame warms caches/allocations outside the profiled range."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402  (device memory / profiler range only)

from core_b200 import scenes  # noqa: E402
from core_b200.interface import YafrayInterface  # noqa: E402

spp = int(os.environ.get("YB_PROFILE_SPP", "4"))
yi = YafrayInterface(device=0)
if os.environ.get("YB_PROFILE_SCENE", "c2") == "c4":      # BASELINE configs[3]: materials scene, pathtracing 8 bounces
    scenes.materials_scene(yi, width=1920, height=1080, spp=spp, bounces=8)
else:
    scenes.icosphere_field(yi, n_tris=1_000_000, width=1920, height=1080, spp=spp, bounces=4)
yi.render(1920, 1080)
torch.cuda.synchronize()
torch.cuda.profiler.start()
img = yi.render(1920, 1080)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
st = yi.getStats()
print("frame: %.1f ms, %d rays, %d launches, mean %.4f" % (st["render_ms"], st["rays_closest"] + st["rays_shadow"],
                                                             st["kernel_launches"], float(img[..., :3].mean())))
