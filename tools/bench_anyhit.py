"""the raw any-hit benchmark - yb200_trace_batch (scene_t::isShadowed -> IntersectS) on the 10 M-triangle
soup with cosine-distributed rays leaving the primary hits of the 1080p camera, AO_distance 0.1 and infinity.
    python tools/bench_anyhit.py [--tris 10000000] [--per-hit 8]
One JSON line per distance with the roofline fraction at A(N) = 48 + 128 ceil(log8(N/4)) + 192 bytes per ray."""
import argparse
import json
import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from core_b200 import scenes  # noqa: E402
from core_b200.interface import YafrayInterface  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--tris", type=int, default=10_000_000)
ap.add_argument("--per-hit", type=int, default=8)
ap.add_argument("--reps", type=int, default=5)
args = ap.parse_args()

yi = YafrayInterface(device=0)
n = scenes.triangle_soup(yi, n_tris=args.tris, width=1920, height=1080, spp=1, ao_samples=32, ao_distance=0.1)
yi.updateScene()
V, F = scenes.soup_mesh(args.tris)
prim = scenes.primary_rays((0.0, -3.2, 0.4), (0.0, 0.0, 0.0), (0.0, -3.2, 1.4), 1920, 1080, 1.2)
hits = yi.traceBatch(prim, 0)
h = hits["prim"] >= 0
tri = V.reshape(-1, 3, 3)[hits["prim"][h]]
ng = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
ng /= np.linalg.norm(ng, axis=1, keepdims=True)
ng = np.where((ng * prim[h, 4:7]).sum(1, keepdims=True) > 0, -ng, ng)           # face the camera ray
P = prim[h, 0:3] + prim[h, 4:7] * hits["t"][h, None]
dev = torch.device("cuda")
g = torch.Generator(device=dev); g.manual_seed(7)
Pn = torch.from_numpy(np.concatenate([P, ng], 1).astype(np.float32)).to(dev).repeat_interleave(args.per_hit, 0)
m = Pn.shape[0]
u1, u2 = torch.rand(m, generator=g, device=dev), torch.rand(m, generator=g, device=dev)
N = Pn[:, 3:6]
a = torch.where(N[:, 0:1].abs() > 0.9, torch.tensor([0.0, 1.0, 0.0], device=dev), torch.tensor([1.0, 0.0, 0.0], device=dev)).expand(m, 3)
T = torch.nn.functional.normalize(torch.cross(N, a, dim=1), dim=1)
B = torch.cross(N, T, dim=1)
r, phi = u1.sqrt(), 2 * math.pi * u2                                            # cosine-weighted hemisphere
d = T * (r * phi.cos())[:, None] + B * (r * phi.sin())[:, None] + N * (1 - u1).clamp_min(0).sqrt()[:, None]
A = 48 + 128 * math.ceil(math.log(max(n / 4.0, 1.0), 8)) + 192
peak = 6549.4
pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
if os.path.exists(pk):
    peak = float(json.load(open(pk))["hbm_gbs"])
ext = torch.cuda.ExternalStream(yi.cudaStream())
for dist_name, tmax in (("0.1", 0.1), ("infinity", -1.0)):
    rays = torch.zeros(m, 8, device=dev)
    rays[:, 0:3] = Pn[:, 0:3]; rays[:, 3] = 0.0005; rays[:, 4:7] = d; rays[:, 7] = tmax
    out = torch.zeros(m, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    yi.traceBatchDevice(rays.data_ptr(), m, 1, out.data_ptr())                  # warm-up
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(args.reps):
        yi.traceBatchDevice(rays.data_ptr(), m, 1, out.data_ptr())
    e1.record(ext)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.reps
    grays = m / ms / 1e6
    print(json.dumps({"benchmark": "c3 raw any-hit (yb200_trace_batch, YB200_TRACE_SHADOW)", "n_triangles": n, "rays": m, "AO_distance": dist_name,
                      "occluded_fraction": float(out.float().mean()), "ms_per_launch": ms, "Grays_per_s": grays,
                      "roofline": {"bound": "hbm", "algorithmic_bytes_per_ray": A, "achieved": grays * A, "peak": peak, "unit": "GB/s",
                                   "frac": grays * A / peak}}), flush=True)
