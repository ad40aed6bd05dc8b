"""The five BASELINE.json configs on one B200, one line of JSON each (configs[1] is bench.py's workload and only listed for
completeness; the others are parity-test cases, rendered here at the stated size for their throughput):
    python tools/run_configs.py [c1 c3 c4 c5 ...] > gpurun_out/configs.jsonl
Rays are the device counters of yb200_get_stats; time is render only, scene already resident (like bench.py's `value`)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from core_b200 import scenes  # noqa: E402
from core_b200.interface import YafrayInterface  # noqa: E402

CONFIGS = {
    "c1": ("Cornell box, shinydiffuse + area light, directlighting, 512x512, 16 spp", scenes.cornell_box,
           dict(width=512, height=512, spp=16, light_samples=8), 3),
    "c2": ("1M-triangle icosphere field, pathtracing 4 bounces, 1920x1080, 64 spp", scenes.icosphere_field,
           dict(n_tris=1_000_000, width=1920, height=1080, spp=64, bounces=4), 2),
    "c3": ("10M-triangle soup, directlighting + ambient occlusion 32 samples (IntersectS stress), 1920x1080, 4 spp", scenes.triangle_soup,
           dict(n_tris=10_000_000, width=1920, height=1080, spp=4, ao_samples=32, ao_distance=0.1), 2),
    "c4": ("glossy / coated glossy / glass, sun + point + area, pathtracing 8 bounces, 1920x1080, 256 spp", scenes.materials_scene,
           dict(width=1920, height=1080, spp=256, bounces=8), 1),
    "c5": ("50M-triangle terrain, pathtracing 4 bounces, 3840x2160, 64 spp of the config's 1024 (one GPU; time scales linearly in spp)",
           scenes.terrain, dict(grid=5000, width=3840, height=2160, spp=64, bounces=4), 1),
}

sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402  (the textured / bumped material sets of the parity cases, at full resolution)
CONFIGS["c4tex"] = ("materials scene with procedural textures + shader nodes on every slot (tests/cases.py _TEXTURED), pathtracing 4 bounces, 1080p, 16 spp",
                    scenes.materials_scene, dict(width=1920, height=1080, spp=16, bounces=4, **cases._TEXTURED), 2)
CONFIGS["c4img"] = ("materials scene with image textures, projections, bump and normal maps (tests/cases.py _IMAGED), pathtracing 4 bounces, 1080p, 16 spp",
                    scenes.materials_scene, dict(width=1920, height=1080, spp=16, bounces=4, **cases._IMAGED), 2)

for key in (sys.argv[1:] or ["c1", "c2", "c3", "c4", "c5"]):
    key, _, over = key.partition(":")              # "c4:spp=4" overrides keyword arguments of the scene (profiling runs)
    desc, fn, kw, reps = CONFIGS[key]
    if over:
        kw = dict(kw, **{k: int(v) for k, v in (o.split("=") for o in over.split(","))})
        desc += " [" + over + "]"
    yi = YafrayInterface(device=0)
    t0 = time.time()
    fn(yi, **kw)
    t_scene = time.time() - t0
    t0 = time.time()
    yi.updateScene()
    t_build = time.time() - t0
    yi.render(kw["width"], kw["height"]) if reps > 1 else None       # warm-up where the frame is short
    best = None
    for _ in range(max(1, reps - 1)):
        img = yi.render(kw["width"], kw["height"])
        st = yi.getStats()
        if best is None or st["render_ms"] < best["render_ms"]:
            best = dict(st)
    rays = best["rays_closest"] + best["rays_shadow"]
    print(json.dumps({"config": key, "what": desc, "n_triangles": best["n_triangles"], "camera_samples": best["camera_samples"],
                      "rays": rays, "sec_per_frame": best["render_ms"] / 1e3, "Mrays_per_s": rays / best["render_ms"] / 1e3,
                      "trace_closest_ms": best["trace_closest_ms"], "trace_shadow_ms": best["trace_shadow_ms"], "shade_ms": best["shade_ms"],
                      "film_ms": best["film_ms"], "bvh_build_ms": best["build_ms"], "host_scene_gen_s": t_scene, "update_scene_s": t_build,
                      "image_mean": float(img[..., :3].mean())}), flush=True)
    del yi
