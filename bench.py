#!/usr/bin/env python
"""bench.py — BASELINE.json's metric (Mrays/s, s/frame of the path tracer) on BASELINE.json configs[1]:
procedural 1M-triangle icosphere field, pathtracing 4 bounces, 1920x1080, 64 spp.

  python bench.py --gpus N --steps K --warmup W            # the B200 path (one rank per GPU under torchrun for N > 1)
  python bench.py --impl reference --steps K --warmup W    # the reference's own CPU path (oracle/_ref) on the host cores

A step = one frame.  `value` is timed with the scene and BVH already resident in HBM; `e2e` goes through the
reference-facing call with host buffers: geometry re-flattened and uploaded, BVH rebuilt, frame rendered and the float
RGBA frame read back to the host, every step.  A ray = one scene_t::intersect or isShadowed call,
counted exactly on the device.  One JSON line on stdout (rank 0).
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

# stdout carries exactly one JSON line: NCCL prints its version banner there whenever NCCL_DEBUG is set
if os.environ.get("NCCL_DEBUG"):
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")

# ... and the reference library and NCCL print banners on fd 1 from native code: everything but the final line goes to stderr
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # name: (scene kwargs, cpu sample kwargs)
    "c2_field_1M_1080p_64spp": dict(n_tris=1_000_000, width=1920, height=1080, spp=64, bounces=4),
    "c2_field_small": dict(n_tris=100_000, width=480, height=270, spp=8, bounces=4),      # smoke-size, not a bench line
}
CPU_SAMPLE = dict(width=960, height=540, spp=4)      # bounded sample of the same scene for the CPU reference
# exact ray counts of the bounded CPU sample (deterministic sampler + parity => identical ray sets on both sides,
# SURVEY.md §8d); measured with the device counters of this repo by `bench.py --count-cpu-sample`, see DESIGN.md §Measurement
CPU_SAMPLE_RAYS = {"c2_field_1M_1080p_64spp": None, "c2_field_small": None}
RAYS_JSON = os.path.join(ROOT, "profiles", "cpu_sample_rays.json")


def algorithmic_bytes_per_ray(n_tris):
    """A(N) = 48 + 128*ceil(log8(N/4)) + 192"""
    return 48 + 128 * math.ceil(math.log(max(n_tris / 4.0, 1.0), 8)) + 192


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock + throttle reasons DURING the timed region (B200_PROFILING.md recipe).  NVML is polled every 10 ms from a thread
    (an nvidia-smi child needs longer to start than a multi-GPU timed region lasts); nvidia-smi -lms is the fallback."""
    REASONS = [("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4)]

    def __init__(self, index):
        self.rows, self.proc, self.index, self.nvml, self.stop = [], None, index, None, False

    def _poll_nvml(self):
        n = self.nvml
        while True:
            try:
                sm = n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)
                mx = n.nvmlDeviceGetMaxClockInfo(self.h, n.NVML_CLOCK_SM)
                try:
                    rs = n.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    rs = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.rows.append([str(sm), str(mx)] + ["Active" if rs & bit else "Not Active" for _, bit in self.REASONS])
            except Exception:
                pass
            if self.stop:
                return
            time.sleep(0.01)

    def __enter__(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.nvml = pynvml
            self.t = threading.Thread(target=self._poll_nvml, daemon=True)
            self.t.start()
            return self
        except Exception:
            self.nvml = None
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.nvml:
            self.stop = True          # the poll loop takes one more sample (GPU still warm) and returns
            self.t.join(timeout=2)
            return
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = [n for n, _ in self.REASONS]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm), "source": "nvml 10 ms poll" if self.nvml else "nvidia-smi -lms 200"}


def build_scene(yi, wl, **over):
    from core_b200 import scenes
    kw = dict(WORKLOADS[wl])
    kw.update(over)
    n = scenes.icosphere_field(yi, **kw)
    return n, kw


def cpu_reference_run(wl, steps, warmup, threads=-1):
    """The reference's own CPU implementation (oracle/_ref/libyref.so, the unmodified sources) on the host cores."""
    import refapi
    if not refapi.have_ref():
        return None
    yi = refapi.RefInterface()
    yi.setVerbosityMute()
    n_tris, kw = build_scene(yi, wl, threads=threads, **CPU_SAMPLE)
    t0 = time.time()
    yi.updateScene()                       # kd-tree build (scene.cc:782), reported separately
    kd_s = time.time() - t0
    times = []
    frame = None
    for i in range(warmup + steps):
        t = time.time()
        frame = yi.render(kw["width"], kw["height"])
        if i >= warmup:
            times.append(time.time() - t)
    rays = None
    if os.path.exists(RAYS_JSON):
        rays = json.load(open(RAYS_JSON)).get(wl)
    return dict(n_tris=n_tris, kw=kw, sec=float(np.mean(times)), kd_s=kd_s, rays=rays, cores=os.cpu_count(), frame=frame)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2_field_1M_1080p_64spp", choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--count-cpu-sample", action="store_true", help="count the rays of the bounded CPU sample on the GPU and store them")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    wl = args.workload
    full = WORKLOADS[wl]
    metric = "Mrays/s (pathtracer, rays = scene_t::intersect + isShadowed calls)"
    config = {"workload": wl, "n_triangles": None, "integrator": "pathtracing", "bounces": full["bounces"], "path_samples": 1,
              "resolution": [full["width"], full["height"]], "spp": full["spp"], "lights": "sun(4 samples)+point",
              "partition": "image tiles 32x32 round-robin over ranks, BVH replicated, film reduce to rank 0",
              "l2": "working set (geometry+BVH ~110 MB, task queues > 1 GB) exceeds the 126 MB L2; no explicit flush"}

    # ------------------------------------------------------------------ reference arm (CPU, rank 0 only)
    if args.impl == "reference":
        if rank != 0:
            return 0
        r = cpu_reference_run(wl, args.steps, args.warmup)
        if r is None:
            emit({"impl": "reference", "unavailable": "oracle/_ref/libyref.so missing (build it with make -C oracle ref)"})
            return 0
        config["n_triangles"] = r["n_tris"]
        # what this arm really renders: a bounded sample of the workload, compared per ray
        config["full_config"] = {"resolution": config["resolution"], "spp": config["spp"]}
        config["resolution"] = [r["kw"]["width"], r["kw"]["height"]]
        config["spp"] = r["kw"]["spp"]
        config["note"] = "CPU arm renders the same 1M-triangle scene / integrator at the reduced resolution and spp stated here; Mrays/s is per ray"
        sample = "same scene, %dx%d, %d spp, threads=%d (kd-tree build %.1f s not included)" % (
            r["kw"]["width"], r["kw"]["height"], r["kw"]["spp"], r["cores"], r["kd_s"])
        rays = r["rays"]
        val = (rays / r["sec"] / 1e6) if rays else None
        line = {"impl": "reference", "metric": metric, "value": val, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": r["sec"] * 1e3, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": r["cores"], "kind": "reference", "sample": sample,
                                 "rays_per_step": rays, "sec_per_step": r["sec"]},
                "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit(line)
        return 0

    # ------------------------------------------------------------------ B200 arm
    import torch
    import torch.distributed as dist
    from core_b200.interface import YafrayInterface
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the B200 path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    yi = YafrayInterface(device=local)
    n_tris, kw = build_scene(yi, wl)
    config["n_triangles"] = n_tris
    W, H = kw["width"], kw["height"]
    yi.updateScene()
    build_ms = yi.getStats()["build_ms"]
    film = torch.zeros(H, W, 5, device="cuda")
    rgba_dev = torch.zeros(H, W, 4, device="cuda")
    ext = torch.cuda.ExternalStream(yi.cudaStream()) if yi.cudaStream() else torch.cuda.current_stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def reduce_film():
        # NCCL runs on its own stream and only torch's current stream waits for it: order the library's (non-blocking)
        # stream after the reduce before anything on it reads the film
        dist.reduce(film, 0)
        ext.wait_stream(torch.cuda.current_stream())

    def frame_resident():
        film.zero_()
        torch.cuda.synchronize()
        yi.renderFilm(film, rank, world)
        if world > 1:
            reduce_film()
        if rank == 0:
            yi._call("film_resolve", film.data_ptr(), rgba_dev.data_ptr(), 1)

    def frame_e2e(host_out):
        yi.touchGeometry()                 # re-flatten + H2D of the geometry + BVH build inside the step
        if world == 1:                     # the reference-facing call itself: yafrayInterface_t::render -> yb200_render, host buffer out
            if not yi._call("render", host_out.ctypes.data, W, H, None):
                raise RuntimeError(yi.lastError())
            return
        film.zero_()
        torch.cuda.synchronize()
        yi.renderFilm(film, rank, world)
        reduce_film()
        if rank == 0:
            yi._call("film_resolve", film.data_ptr(), host_out.ctypes.data, 0)   # D2H of the float RGBA frame

    def agg(stats_list):
        keys = ["rays_closest", "rays_shadow", "kernel_launches", "trace_closest_ms", "trace_shadow_ms", "shade_ms", "film_ms", "camera_samples"]
        return {k: sum(s[k] for s in stats_list) for k in keys}

    for _ in range(args.warmup):
        frame_resident()
    barrier()
    per_step = []
    with ClockSampler(local) as clk:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        e0.record(ext)
        for _ in range(args.steps):
            frame_resident()
            per_step.append(yi.getStats())
        e1.record(ext)
        barrier()
        wall = time.time() - t0
    dev_ms = e0.elapsed_time(e1)
    t = torch.tensor([dev_ms, wall * 1e3], device="cuda", dtype=torch.float64)
    a = agg(per_step)
    cnt = torch.tensor([a["rays_closest"], a["rays_shadow"], a["kernel_launches"], a["camera_samples"]], device="cuda", dtype=torch.float64)
    ktime = torch.tensor([a["trace_closest_ms"], a["trace_shadow_ms"], a["shade_ms"], a["film_ms"]], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        dist.all_reduce(ktime, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms = float(t[0]), float(t[1])
    rays_total = float(cnt[0] + cnt[1])
    ms_per_step = dev_ms / args.steps
    value = rays_total / (dev_ms * 1e-3) / 1e6

    # e2e through the host-buffer path.  (1) step by step: every step re-flattens, uploads, rebuilds, renders and reads back
    # before the next one starts (latency of one frame);
    host_out = np.zeros((H, W, 4), np.float32)
    frame_e2e(host_out)
    barrier()
    t0 = time.time()
    e2e_steps = args.steps
    for _ in range(e2e_steps):
        frame_e2e(host_out)
    barrier()
    lat = torch.tensor([(time.time() - t0) / e2e_steps], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(lat, op=dist.ReduceOp.MAX)
    e2e_latency_s = float(lat[0])
    # (2) pipelined, the way a sequence of frames is rendered: two contexts; while one renders step i, the other one takes in the
    # geometry of step i + 1 (flatten, H2D, BVH build on its own stream) from a second host thread.  Every step still does all
    # of the work inside the timed region; only the order overlaps.
    yj = YafrayInterface(device=local)
    build_scene(yj, wl)
    yj.updateScene()
    ctxs = [yi, yj]
    exts = [ext, torch.cuda.ExternalStream(yj.cudaStream())]

    def update(c):
        c.touchGeometry()
        c.updateScene()

    def render_step(c, cext, host):
        if world == 1:
            # yb200_render: finds the scene up to date (the update ran in the other thread), renders, resolves, D2H
            if not c._call("render", host.ctypes.data, W, H, None):
                raise RuntimeError(c.lastError())
            return
        film.zero_()
        torch.cuda.synchronize()
        c.renderFilm(film, rank, world)
        dist.reduce(film, 0)
        cext.wait_stream(torch.cuda.current_stream())
        if rank == 0:
            c._call("film_resolve", film.data_ptr(), host.ctypes.data, 0)

    step_log = []

    def pipelined(steps):
        update(ctxs[0])
        for i in range(steps):
            cur, nxt = i % 2, (i + 1) % 2
            th = threading.Thread(target=update, args=(ctxs[nxt],)) if i + 1 < steps else None
            ta = time.time()
            if th:
                th.start()
            render_step(ctxs[cur], exts[cur], host_out)
            tb = time.time()
            if th:
                th.join()
            step_log.append(((tb - ta) * 1e3, (time.time() - tb) * 1e3))

    pipelined(2)
    barrier()
    t0 = time.time()
    pipelined(e2e_steps)
    barrier()
    e2e_s = torch.tensor([(time.time() - t0) / e2e_steps], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e_s[0])
    st = yi.getStats()
    rays_per_frame = rays_total / args.steps
    e2e_val = rays_per_frame / e2e_s / 1e6
    h2d = int(st["tri_bytes"]) * world
    d2h = W * H * 16

    frame_check = None
    if world > 1:
        # the reduced N-rank frame (host_out on rank 0, from the last e2e step) against a 1-rank render of the same frame
        if rank == 0:
            import cases
            film.zero_()
            torch.cuda.synchronize()
            yi.renderFilm(film, 0, 1)
            one = yi.filmResolve(film, W, H)
            rel, psnr, da = cases.compare_images(one, host_out)
            frame_check = {"rel": float(rel), "psnr": float(psnr), "max_alpha_diff": float(da),
                           "what": "frame reduced over %d ranks (NCCL) vs the same frame rendered by rank 0 alone" % world}
        barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # roofline of the dominant kernels (closest-hit + any-hit traversal; both walk the same BVH)
    A = algorithmic_bytes_per_ray(n_tris)
    trace_ms = float(ktime[0] + ktime[1])
    peak, peak_src = measured_peak_hbm()
    rays_rank = rays_total / world                       # max-over-ranks kernel time vs this rank's share (balanced tiles)
    achieved = rays_rank * A / (trace_ms * 1e-3) / 1e9 if trace_ms > 0 else None
    traffic, traffic_note = None, None
    tfiles = sorted(f for f in os.listdir(os.path.join(ROOT, "profiles")) if "traffic" in f and f.endswith(".json"))
    if tfiles and args.workload == "c2_field_1M_1080p_64spp":
        tj = json.load(open(os.path.join(ROOT, "profiles", tfiles[-1])))
        nl = len(tj["launches"])
        traffic = tj["dram_bytes"] / nl
        traffic_note = ("ncu dram__bytes_read+write per traversal launch, mean of the %d launches of one frame of this scene at the same "
                        "batch size (profiles/%s); algorithmic bytes of those launches: %.3g per launch - the 59 MB of nodes+triangles "
                        "stay in L2 (%.0f B/ray reach DRAM, %.0f B/ray reach L2)" %
                        (nl, tfiles[-1], tj["rays"] * tj["algorithmic_bytes_per_ray"] / nl, tj["dram_bytes_per_ray"], tj["l2_bytes_per_ray"]))
    roofline = {"bound": "hbm", "kernel": "k_trace_closest + k_trace_shadow (BVH8 traversal)", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": (achieved / peak) if achieved else None, "traffic": traffic, "traffic_note": traffic_note, "peak_source": peak_src,
                "algorithmic_bytes_per_ray": A, "trace_ms_per_step": trace_ms / args.steps,
                "grays_per_s_in_trace_kernels": rays_rank / (trace_ms * 1e-3) / 1e9 if trace_ms > 0 else None,
                "kernel_ms_per_step": {"trace_closest": float(ktime[0]) / args.steps, "trace_shadow": float(ktime[1]) / args.steps,
                                       "shade": float(ktime[2]) / args.steps, "film": float(ktime[3]) / args.steps}}
    # the other two bounds of SURVEY 8(d) / VERDICT: dependent-load latency (pointer chase measured now) and instruction issue
    # (thread instructions per ray from the ncu capture recorded in profiles/)
    R = math.ceil(math.log(max(n_tris / 4.0, 1.0), 8)) + 1
    try:
        t_miss = yi.probeLatency(1 << 30, 20000)
        t_hit = yi.probeLatency(16 << 20, 20000)
    except Exception:
        t_miss = t_hit = None
    clk_summary = clk.summary()
    sm_hz = (clk_summary["sm_mhz"] or 1965.0) * 1e6
    roofline["t_l2miss_ns"] = t_miss
    roofline["t_l2hit_ns"] = t_hit
    roofline["dependent_fetch_rounds"] = R
    if t_miss:
        roofline["latency_bound_grays"] = 148 * 64 * 32 / (R * t_miss * 1e-9) / 1e9          # all 64 warps per SM resident
        roofline["latency_bound_grays_at_kernel_occupancy"] = 148 * 8 * 128 / (R * t_miss * 1e-9) / 1e9   # k_trace_fast: 8 x 128 threads per SM (YB_TRACE_CTAS)
    roofline["hbm_bound_grays"] = peak / A
    ifiles = sorted(f for f in os.listdir(os.path.join(ROOT, "profiles")) if "trace_inst" in f and f.endswith(".json"))
    if ifiles and args.workload == "c2_field_1M_1080p_64spp":
        ij = json.load(open(os.path.join(ROOT, "profiles", ifiles[-1])))
        roofline["thread_inst_per_ray"] = ij["thread_inst_per_ray"]
        roofline["warp_inst_per_ray"] = ij["warp_inst_per_ray"]
        roofline["lanes_active"] = ij["lanes_active"]
        # what binds (ncu: issue slots ~70 % busy, DRAM < 5 %): 4 warp instructions per SM and clock
        roofline["issue_bound_grays"] = 148 * 4 * sm_hz / ij["warp_inst_per_ray"] / 1e9
        roofline["issue_bound_grays_all_lanes_active"] = 148 * 4 * 32 * sm_hz / ij["thread_inst_per_ray"] / 1e9
        roofline["issue_note"] = ("148 SM x 4 schedulers x SM clock / warp instructions per ray at the measured lane efficiency; "
                                  "the second figure assumes 32 of 32 lanes active (ncu, profiles/%s)" % ifiles[-1])
        roofline["binding_bound"] = "issue"
        roofline["frac_of_issue_bound"] = roofline["grays_per_s_in_trace_kernels"] / roofline["issue_bound_grays"] if roofline["grays_per_s_in_trace_kernels"] else None
    line = {"metric": metric, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "sec_per_frame": ms_per_step * 1e-3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "rays_per_frame": rays_per_frame, "camera_samples_per_frame": float(cnt[3]) / args.steps, "wall_ms_per_step": wall_ms / args.steps,
            "bvh_build_ms": build_ms, "gpu_launches": int(cnt[2]), "clocks": clk_summary,
            "e2e": {"value": e2e_val, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "sec_per_frame": e2e_s, "steps": e2e_steps,
                    "pipeline": "two contexts: the scene update (flatten + H2D + BVH build) of step i+1 runs while step i renders; every step does all of it inside the timed region",
                    "rank0_ms_render_step_then_wait_for_update": [[round(a, 2), round(b, 2)] for a, b in step_log[-e2e_steps:]],
                    "unpipelined": {"value": rays_per_frame / e2e_latency_s / 1e6, "sec_per_frame": e2e_latency_s,
                                    "note": "one step after the other, nothing overlapped (latency of a single frame)"},
                    "call": "yb200_render (yafrayInterface_t::render) with a host RGBA buffer" if world == 1 else "yb200_render_film + NCCL reduce + yb200_film_resolve to a host buffer",
                    "includes": "flatten + H2D geometry + BVH build + render + film reduce + D2H frame"},
            "roofline": roofline}
    if frame_check:
        line["frame_check"] = frame_check

    # the reference's CPU path on this box's host cores, bounded sample (rank 0, N = 1 only)
    if world == 1 and not args.no_cpu_baseline:
        if args.count_cpu_sample or not os.path.exists(RAYS_JSON):
            y2 = YafrayInterface(device=local)
            build_scene(y2, wl, **CPU_SAMPLE)
            y2.render(CPU_SAMPLE["width"], CPU_SAMPLE["height"])
            s2 = y2.getStats()
            d = json.load(open(RAYS_JSON)) if os.path.exists(RAYS_JSON) else {}
            d[wl] = int(s2["rays_closest"] + s2["rays_shadow"])
            os.makedirs(os.path.dirname(RAYS_JSON), exist_ok=True)
            json.dump(d, open(RAYS_JSON, "w"), indent=1)
            if os.path.isdir(os.path.join(ROOT, "gpurun_out")):
                json.dump(d, open(os.path.join(ROOT, "gpurun_out", "cpu_sample_rays.json"), "w"), indent=1)
            y2.close()
        r = cpu_reference_run(wl, 1, 1)
        if r is not None and r.get("frame") is not None:
            # the frame the reference just rendered against a GPU render of the same sample (after all timed regions)
            import cases
            y3 = YafrayInterface(device=local)
            build_scene(y3, wl, **CPU_SAMPLE)
            g = y3.render(CPU_SAMPLE["width"], CPU_SAMPLE["height"])
            y3.close()
            rel, psnr, da = cases.compare_images(r["frame"], g)
            line["parity"] = {"rel": float(rel), "psnr": float(psnr), "max_alpha_diff": float(da),
                              "sample": "reference frame of the cpu_baseline leg (%dx%d, %d spp, 1M triangles) vs the CUDA path on the same sample" % (
                                  r["kw"]["width"], r["kw"]["height"], r["kw"]["spp"])}
        if r is not None and r["rays"]:
            line["cpu_baseline"] = {"value": r["rays"] / r["sec"] / 1e6, "unit": "Mrays/s", "cores": r["cores"], "kind": "reference",
                                    "sample": "same scene, %dx%d, %d spp, threads=%d; kd-tree build %.1f s excluded" % (
                                        r["kw"]["width"], r["kw"]["height"], r["kw"]["spp"], r["cores"], r["kd_s"]),
                                    "rays_per_step": r["rays"], "sec_per_step": r["sec"]}
        else:
            line["cpu_baseline"] = {"value": None, "unit": "Mrays/s", "cores": os.cpu_count(), "kind": "reference", "sample": "oracle/_ref missing"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
