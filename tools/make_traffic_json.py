"""profiles/rN_traffic.json from one ncu metrics pass over tools/profile_frame.py and the wave log of the same frame:
    YB200_LOG_WAVES=1 python tools/profile_frame.py 2> gpurun_out/waves.log          # exits 0 first
    YB200_LOG_WAVES=1 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,gpu__time_duration.sum \
        --clock-control none --profile-from-start off --csv --log-file gpurun_out/traffic.csv python tools/profile_frame.py
    python tools/make_traffic_json.py gpurun_out/traffic.csv gpurun_out/waves.log profiles/rN_traffic.json
(YB200_LOG_WAVES makes the host count every wave, so that the launches of the log and of the capture correspond one to one.)"""
import collections
import csv
import json
import re
import sys

src, waves, dst = sys.argv[1:4]
rows = list(open(src))
i = next(k for k, l in enumerate(rows) if l.startswith('"ID"'))
per = collections.OrderedDict()
for r in csv.DictReader(rows[i:]):
    if "k_trace_fast" not in r["Kernel Name"]:
        continue
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"].lower()
    scale = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "ms": 1, "msecond": 1}.get(u, 1)
    per.setdefault(r["ID"], {"kernel": "k_trace_fast<any-hit>" if "k_trace_fast<1" in r["Kernel Name"] or "(bool)1" in r["Kernel Name"] else "k_trace_fast<closest>"})[r["Metric Name"]] = v * scale
w = [tuple(int(x) for x in m.groups()) for m in re.finditer(r"yb200 wave: batch (\d+) closest (\d+) shadow (\d+) children (\d+)", open(waves).read())]
w = w[len(w) // 2:]                                  # the log holds the warm-up frame and the profiled frame
counts = []
for (_, nc, ns, _) in w:
    counts.append(nc)
    if ns:
        counts.append(ns)
L = list(per.values())
assert len(L) == len(counts), (len(L), len(counts))
out = {"source": "tools/make_traffic_json.py: ncu dram__bytes_read/write.sum, lts__t_bytes.sum, gpu__time_duration.sum (--clock-control none) on "
                 "tools/profile_frame.py (1M-tri field, 1080p, 4 spp, one frame), ray counts from YB200_LOG_WAVES=1 of the same frame",
       "launches": []}
for l, n in zip(L, counts):
    out["launches"].append({"kernel": l["kernel"], "rays": n, "dram_bytes": l["dram__bytes_read.sum"] + l["dram__bytes_write.sum"],
                            "l2_bytes": l["lts__t_bytes.sum"], "ms": l["gpu__time_duration.sum"]})
out["rays"] = sum(x["rays"] for x in out["launches"])
out["dram_bytes"] = sum(x["dram_bytes"] for x in out["launches"])
out["l2_bytes"] = sum(x["l2_bytes"] for x in out["launches"])
out["dram_bytes_per_ray"] = out["dram_bytes"] / out["rays"]
out["l2_bytes_per_ray"] = out["l2_bytes"] / out["rays"]
out["algorithmic_bytes_per_ray"] = 1008
json.dump(out, open(dst, "w"), indent=1)
print({k: out[k] for k in ("rays", "dram_bytes_per_ray", "l2_bytes_per_ray")}, len(L), "launches")
