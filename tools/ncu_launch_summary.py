import csv, sys, collections
rows = list(open(sys.argv[1]))
i = next(k for k, l in enumerate(rows) if l.startswith('"ID"'))
r = csv.DictReader(rows[i:])
t = collections.defaultdict(float); n = collections.Counter()
for row in r:
    if row.get("Metric Name") != "gpu__time_duration.sum": continue
    name = row["Kernel Name"][:90]
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3}.get(u, 1e-6)
    t[name] += v; n[name] += 1
tot = sum(t.values())
for k, v in sorted(t.items(), key=lambda kv: -kv[1])[:40]:
    print(f"{v:10.3f} ms {100*v/tot:5.1f}% {n[k]:6d}  {k}")
print("total", tot)
