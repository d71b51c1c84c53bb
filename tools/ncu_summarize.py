#!/usr/bin/env python3
"""Turn ncu outputs brought back in gpurun_out/ into the small text summaries kept under profiles/.

    python tools/ncu_summarize.py launches gpurun_out/launches.csv profiles/rNN_launches_X.txt
    python tools/ncu_summarize.py full gpurun_out/prof.ncu-rep profiles/rNN_X_ncu_full.txt
"""
import collections
import csv
import io
import re
import subprocess
import sys

KEEP = (r"dram__bytes_(read|write)\.sum$|gpu__time_duration\.sum|dram__throughput\.avg\.pct|lts__t_bytes\.sum$|"
        r"lts__t_sector_hit_rate|l1tex__t_sector_hit_rate|sm__warps_active\.avg\.pct|launch__|"
        r"smsp__issue_active\.avg\.pct|sm__throughput\.avg\.pct|l1tex__throughput\.avg\.pct|lts__throughput\.avg\.pct|"
        r"smsp__inst_executed\.sum$|sm__pipe_(alu|fma|fmaheavy|xu|lsu)_cycles_active|sm__inst_executed_pipe_.*\.sum$|"
        r"smsp__average_warps_issue_stalled.*_per_issue_active|l1tex__t_(requests|sectors)_pipe_lsu_mem_global_op_(ld|st)\.sum$|"
        r"sm__cycles_elapsed\.max|smsp__thread_inst_executed_per_inst_executed|sm__sass_thread_inst_executed_op_.*\.sum$")


def launches(src, dst):
    rows = list(csv.reader(l for l in open(src) if l.startswith('"')))
    h = rows[0]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1.0)
        a = agg.setdefault(r[ki].split("(")[0], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    with open(dst, "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none; source {src}; {len(rows) - 1} launches\n")
        f.write("# cold-cache, serialised launches: compare SHARES, not absolutes\n")
        f.write(f"{'kernel':60s} {'launches':>9s} {'total_ms':>10s} {'avg_ms':>10s} {'share':>7s}\n")
        for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write(f"{n:60s} {c:9d} {t:10.3f} {t / c:10.4f} {t / tot:7.3f}\n")
    print(open(dst).read())


def full(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, u = rows[0], rows[1]
    with open(dst, "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on; source {src}\n")
        for r in rows[2:]:
            f.write(f"## kernel {r[h.index('Kernel Name')]} grid {r[h.index('Grid Size')]} block {r[h.index('Block Size')]}\n")
            for a, b, c in zip(h, u, r):
                if re.search(KEEP, a):
                    f.write(f"{a:86s} {c} {b}\n")
    print(open(dst).read())


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
