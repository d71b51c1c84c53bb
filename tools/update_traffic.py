#!/usr/bin/env python3
"""profiles/traffic.json from the ncu --set full summaries (tools/ncu_summarize.py full): the counters bench.py prints
under roofline.ncu / traffic, tagged with the hash of the device code they were captured from.
usage: python tools/update_traffic.py profiles/<carve summary>.txt profiles/<averaging summary>.txt"""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def parse(path):
    ks, cur = [], None
    for l in open(path):
        if l.startswith("## kernel"):
            cur = {"name": l.split("## kernel", 1)[1].strip()[:60]}
            ks.append(cur)
            continue
        m = re.match(r"(\S+)\s+([\d.,]+)\s*(\S*)", l)
        if cur is not None and m:
            v = float(m.group(2).replace(",", "")) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3}.get(m.group(3), 1)
            cur[m.group(1)] = v
    return ks


def entry(k):
    return {"kernel_ms": k["gpu__time_duration.sum"], "dram_bytes": k["dram__bytes_read.sum"] + k["dram__bytes_write.sum"],
            "issue_active_pct": k["smsp__issue_active.avg.pct_of_peak_sustained_active"],
            "l1_hit_pct": k["l1tex__t_sector_hit_rate.pct"], "l2_hit_pct": k["lts__t_sector_hit_rate.pct"],
            "warp_instructions": k["smsp__inst_executed.sum"],
            "warps_active_pct": k["sm__warps_active.avg.pct_of_peak_sustained_active"]}


def main():
    carve_txt, avg_txt = sys.argv[1], sys.argv[2]
    kh = ge.source_hashes()[1]
    c, a = parse(carve_txt), parse(avg_txt)
    items = entry(next(k for k in c if "carve_items_kernel" in k["name"]))
    segs = entry(next(k for k in c if "carve_segments_kernel" in k["name"]))
    tot = items["kernel_ms"] + segs["kernel_ms"]
    carv = {"kernel_ms": tot, "dram_bytes": items["dram_bytes"] + segs["dram_bytes"],
            "issue_active_pct": (items["issue_active_pct"] * items["kernel_ms"] + segs["issue_active_pct"] * segs["kernel_ms"]) / tot,
            "l1_hit_pct": segs["l1_hit_pct"], "l2_hit_pct": segs["l2_hit_pct"],
            "warp_instructions": items["warp_instructions"] + segs["warp_instructions"], "warps_active_pct": segs["warps_active_pct"],
            "kernel": "carve_items_kernel<false> (first 16 views) + carve_segments_kernel<false> (remaining views as segments): "
                      "sums / time-weighted means of the two launches, hit rates of the second",
            "kern_hash": kh, "source": os.path.relpath(carve_txt, ROOT), "carve_items_kernel": items, "carve_segments_kernel": segs}
    avg = dict(entry(a[0]), kernel="average_kernel<u8> with view skipping", kern_hash=kh, source=os.path.relpath(avg_txt, ROOT))
    path = os.path.join(ROOT, "profiles", "traffic.json")
    d = json.load(open(path))
    d["ncu"] = {"carving": carv, "averaging": avg}
    json.dump(d, open(path, "w"), indent=1)
    print(kh, {k: carv[k] for k in ("kernel_ms", "dram_bytes", "issue_active_pct", "warp_instructions")})
    print({k: avg[k] for k in ("kernel_ms", "dram_bytes", "issue_active_pct", "warp_instructions")})


if __name__ == "__main__":
    main()
