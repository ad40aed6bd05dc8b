"""profiles/rN_trace_inst.json (thread / warp instructions per ray of the traversal kernels; bench.py's issue bound reads it):
    ncu --metrics smsp__thread_inst_executed.sum,smsp__inst_executed.sum,gpu__time_duration.sum -k regex:k_trace_fast \
        --profile-from-start off --clock-control none --csv --log-file gpurun_out/trace_inst.csv python tools/profile_frame.py
    python tools/make_trace_inst_json.py gpurun_out/trace_inst.csv 75531780 profiles/rN_trace_inst.json
(the ray count is what tools/profile_frame.py prints for the frame)."""
import collections
import csv
import json
import sys

src, rays, dst = sys.argv[1], int(sys.argv[2]), sys.argv[3]
rows = list(open(src))
i = next(k for k, l in enumerate(rows) if l.startswith('"ID"'))
per = collections.defaultdict(lambda: collections.Counter())
n = 0
for r in csv.DictReader(rows[i:]):
    kind = "k_trace_fast<any-hit>" if r["Kernel Name"].startswith("void k_trace_fast<1") else "k_trace_fast<closest>"
    per[kind][r["Metric Name"]] += float(r["Metric Value"].replace(",", ""))
    n += r["Metric Name"] == "gpu__time_duration.sum"
ti = sum(v["smsp__thread_inst_executed.sum"] for v in per.values())
wi = sum(v["smsp__inst_executed.sum"] for v in per.values())
out = {"source": "tools/make_trace_inst_json.py on %s: ncu --metrics smsp__thread_inst_executed.sum,smsp__inst_executed.sum,gpu__time_duration.sum "
                 "-k regex:k_trace_fast --profile-from-start off --clock-control none, python tools/profile_frame.py (1M-triangle field, "
                 "1920x1080, 4 spp, %d traversal launches)" % (src, n),
       "rays": rays, "thread_inst": ti, "warp_inst": wi, "thread_inst_per_ray": ti / rays, "warp_inst_per_ray": wi / rays,
       "lanes_active": ti / wi,
       "per_kernel": {k: {"thread_inst": v["smsp__thread_inst_executed.sum"], "warp_inst": v["smsp__inst_executed.sum"],
                          "lanes_active": v["smsp__thread_inst_executed.sum"] / v["smsp__inst_executed.sum"],
                          "ms": v["gpu__time_duration.sum"] / 1e6} for k, v in per.items()}}
json.dump(out, open(dst, "w"), indent=1)
print({k: out[k] for k in ("thread_inst_per_ray", "warp_inst_per_ray", "lanes_active")})
