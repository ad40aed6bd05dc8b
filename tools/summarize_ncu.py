"""Turn ncu exports into the small summaries kept under profiles/ (the .ncu-rep files and full CSVs stay in gpurun_out/):
    python tools/summarize_ncu.py launches gpurun_out/launches.csv profiles/rN_launches_summary.csv
    python tools/summarize_ncu.py raw gpurun_out/raw.csv profiles/rN_ncu_full.csv
    python tools/summarize_ncu.py source gpurun_out/source.csv.gz profiles/rN_ncu_source_top.csv [kernel-index]
"""
import collections
import csv
import gzip
import sys

csv.field_size_limit(10 ** 9)
KEEP = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "smsp__pcsamp_warps_issue_stalled_long_scoreboard", "smsp__pcsamp_warps_issue_stalled_short_scoreboard",
        "smsp__pcsamp_warps_issue_stalled_wait", "smsp__pcsamp_warps_issue_stalled_no_instructions",
        "smsp__pcsamp_warps_issue_stalled_lg_throttle", "smsp__pcsamp_warps_issue_stalled_not_selected",
        "smsp__pcsamp_warps_issue_stalled_selected", "smsp__pcsamp_warps_issue_stalled_math_pipe_throttle",
        "smsp__pcsamp_warps_issue_stalled_branch_resolving", "smsp__pcsamp_warps_issue_stalled_barrier"]


def opener(path):
    return gzip.open(path, "rt") if path.endswith(".gz") else open(path)


def launches(src, dst):
    rows = [r for r in csv.reader(l for l in opener(src) if l.startswith('"'))]
    hdr = rows[0]
    kn, mv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    tot, cnt = collections.Counter(), collections.Counter()
    for r in rows[1:]:
        if len(r) <= mv or r[hdr.index("Metric Name")] != "gpu__time_duration.sum":
            continue
        name = r[kn].split("(")[0]
        tot[name] += float(r[mv].replace(",", "")) / 1e6      # ns -> ms
        cnt[name] += 1
    total = sum(tot.values())
    with open(dst, "w") as f:
        f.write("kernel,launches,total_ms,share_pct\n")
        for k, v in tot.most_common():
            f.write('"%s",%d,%.3f,%.1f\n' % (k, cnt[k], v, 100 * v / total))


def raw(src, dst):
    rows = list(csv.reader(opener(src)))
    hdr, units = rows[0], rows[1]
    cols = [hdr.index("Kernel Name")] + [hdr.index(k) for k in KEEP if k in hdr]
    with open(dst, "w") as f:
        w = csv.writer(f)
        w.writerow([hdr[i] + (" [%s]" % units[i] if units[i] else "") for i in cols])
        for r in rows[2:]:
            w.writerow([r[i] for i in cols])


def source(src, dst, which=1, top=60):
    kern, hdr, rows, name = 0, None, [], ""
    for row in csv.reader(opener(src)):
        if row and row[0] == "Kernel Name":
            kern += 1
            if kern == which:
                name = row[1]
            if kern > which:
                break
            continue
        if kern != which:
            continue
        if row and row[0] == "Address":
            hdr = row
            continue
        rows.append(row)
    ci = {h: i for i, h in enumerate(hdr)}
    ns = lambda r: int(r[ci["# Samples"]] or 0)
    total = sum(ns(r) for r in rows)
    ops = collections.Counter()
    for r in rows:
        t = r[1].strip().split()
        if t:
            ops[(t[1] if t[0].startswith("@") and len(t) > 1 else t[0]).split(".")[0]] += ns(r)
    with open(dst, "w") as f:
        f.write("# %s: %d SASS instructions, %d stall samples\n" % (name, len(rows), total))
        f.write("# samples by opcode: " + ", ".join("%s %.1f%%" % (k, 100.0 * v / total) for k, v in ops.most_common(14)) + "\n")
        f.write("samples,share_pct,executed,avg_threads,sass\n")
        for r in sorted(rows, key=lambda r: -ns(r))[:top]:
            f.write('%d,%.2f,%s,%s,"%s"\n' % (ns(r), 100.0 * ns(r) / total, r[ci["Instructions Executed"]],
                                              r[ci["Avg. Threads Executed"]], r[1].strip()))


if __name__ == "__main__":
    mode = sys.argv[1]
    if mode == "launches":
        launches(sys.argv[2], sys.argv[3])
    elif mode == "raw":
        raw(sys.argv[2], sys.argv[3])
    else:
        source(sys.argv[2], sys.argv[3], int(sys.argv[4]) if len(sys.argv) > 4 else 1)
