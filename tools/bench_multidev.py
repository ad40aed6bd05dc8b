"""One process, one context, N GPUs (yb200_set_devices): the reference-facing call yb200_render (yafrayInterface_t::render) on
BASELINE.json configs[1] with a host RGBA buffer, geometry re-flattened / uploaded / rebuilt every step.
    python tools/bench_multidev.py --gpus N [--steps K] [--warmup W] [--workload c2|c5]
Prints one JSON line in the shape of bench.py's (e2e only: this IS the end-to-end call)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from core_b200 import scenes  # noqa: E402
from core_b200.interface import YafrayInterface  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--gpus-list", default="", help="comma-separated device counts measured one after the other on ONE scene (e.g. 1,2,4,8)")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--workload", default="c2", choices=["c2", "c5"])
    ap.add_argument("--spp", type=int, default=0)
    args = ap.parse_args()
    counts = [int(x) for x in args.gpus_list.split(",")] if args.gpus_list else [args.gpus]
    yi = YafrayInterface()
    yi.setDevices(list(range(counts[0])))
    if args.workload == "c2":
        W, H, spp = 1920, 1080, args.spp or 64
        n_tris = scenes.icosphere_field(yi, n_tris=1_000_000, width=W, height=H, spp=spp, bounces=4)
        name = "c2_field_1M_1080p_%dspp" % spp
    else:
        W, H, spp = 3840, 2160, args.spp or 64
        n_tris = scenes.terrain(yi, grid=5000, width=W, height=H, spp=spp, bounces=4)
        name = "c5_terrain_50M_4k_%dspp" % spp
    out = np.zeros((H, W, 4), np.float32)
    for k, ngpu in enumerate(counts):
        if k:
            yi.setDevices(list(range(ngpu)))
        args.gpus = ngpu
        measure(yi, args, out, W, H, name, n_tris)


def measure(yi, args, out, W, H, name, n_tris):
    times, rays = [], 0
    for i in range(args.warmup + args.steps):
        yi.touchGeometry()
        t0 = time.time()
        if not yi._call("render", out.ctypes.data, W, H, None):
            raise RuntimeError(yi.lastError())
        dt = time.time() - t0
        st = yi.getStats()
        if i >= args.warmup:
            times.append(dt)
            rays = st["rays_closest"] + st["rays_shadow"]
    sec = float(np.mean(times))
    st = yi.getStats()
    print(json.dumps({"metric": "Mrays/s (pathtracer, rays = scene_t::intersect + isShadowed calls)", "value": rays / sec / 1e6, "unit": "Mrays/s",
                      "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
                      "scaling": "strong", "dtype": "f32", "data": "synthetic",
                      "config": {"workload": name, "n_triangles": n_tris, "processes": 1,
                                 "call": "yb200_set_devices + yb200_render (host RGBA buffer out, geometry rebuilt every step)"},
                      "rays_per_frame": rays, "build_ms_incl_peer_copies": st["build_ms"], "render_ms": st["render_ms"],
                      "kernel_ms_max_over_devices": {k: st[k] for k in ("trace_closest_ms", "trace_shadow_ms", "shade_ms", "film_ms")},
                      "image_mean": float(out[..., :3].mean())}), flush=True)


if __name__ == "__main__":
    main()
