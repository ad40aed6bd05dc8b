import os, sys, time
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, torch
from core_b200 import scenes
from core_b200.interface import YafrayInterface
yi = YafrayInterface(device=0)
scenes.icosphere_field(yi, n_tris=1_000_000, width=1920, height=1080, spp=64, bounces=4)
out = np.zeros((1080, 1920, 4), np.float32)
for i in range(4):
    yi.touchGeometry()
    t0 = time.time()
    ok = yi._call("render", out.ctypes.data, 1920, 1080, None)
    t1 = time.time()
    st = yi.getStats()
    print("frame %d: e2e %.1f ms, render_ms %.1f build_ms %.1f kernels %.1f" % (i, (t1 - t0) * 1e3, st["render_ms"], st["build_ms"],
          st["trace_closest_ms"] + st["trace_shadow_ms"] + st["shade_ms"] + st["film_ms"]), flush=True)
# pieces
for i in range(3):
    yi.touchGeometry()
    t0 = time.time(); yi.updateScene(); t1 = time.time()
    print("updateScene %.1f ms" % ((t1 - t0) * 1e3), flush=True)
film = torch.zeros(1080, 1920, 5, device="cuda"); torch.cuda.synchronize()
for i in range(3):
    t0 = time.time(); yi._call("film_resolve", film.data_ptr(), out.ctypes.data, 0); t1 = time.time()
    print("film_resolve to host %.1f ms" % ((t1 - t0) * 1e3), flush=True)
