#!/usr/bin/env python3
"""Where do a kernel's executed instructions go?  Reads the SASS source page of an ncu report (captured with
--import-source on), finds the loops (backward branches) and attributes every instruction's "Instructions Executed" to its
innermost loop.  python tools/ncu_loops.py <report.ncu-rep> <kernel name pattern> [min share %]"""
import collections
import csv
import io
import re
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
min_share = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
blocks = re.split(r'(?m)^"Kernel Name",', txt)[1:]
for blk in blocks:
    name = blk.split("\n", 1)[0]
    if pat not in name:
        continue
    rows = list(csv.reader(io.StringIO(blk.split("\n", 1)[1])))
    hdr = rows[0]
    ia, isrc, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed")
    ins = [(int(r[ia], 16), r[isrc].strip(), int(r[iex] or 0)) for r in rows[1:] if len(r) > iex and r[ia].startswith("0x")]
    base = ins[0][0]
    addr = [a - base for a, _, _ in ins]
    loops = []
    for n, (a, s, _) in enumerate(ins):
        m = re.search(r"\bBRA(?:\.U)?\b.*?(0x[0-9a-f]+)\s*$", s)
        if m:
            t = int(m.group(1), 16)
            t = t - base if t >= base else t  # ncu prints absolute addresses, cuobjdump offsets
            if t <= a - base and t in addr:
                loops.append((addr.index(t), n))
    total = sum(e for _, _, e in ins)
    share = collections.Counter()
    for n, (_, _, e) in enumerate(ins):
        inner = sorted((l for l in loops if l[0] <= n <= l[1]), key=lambda l: l[1] - l[0])
        share[inner[0] if inner else None] += e
    print(name[:100])
    print(f"  {len(ins)} instructions, {total:.4g} executed (warp level)")
    for l, e in share.most_common():
        if 100.0 * e / total < min_share:
            continue
        if l is None:
            print(f"  outside any loop: {100.0 * e / total:5.1f} %")
            continue
        body = [s for _, s, _ in ins[l[0]:l[1] + 1]]
        ops = collections.Counter((b.split()[1] if b.startswith("@") else b.split()[0]).split(".")[0] for b in body)
        trips = ins[l[1]][2]
        hint = {k: ops[k] for k in ("MUFU", "FMNMX3", "LDG", "SHFL", "VOTE", "VOTEU", "CALL", "ATOMG") if ops[k]}
        print(f"  loop [{l[0]:4d}, {l[1]:4d}] {l[1] - l[0] + 1:4d} instr, back edge taken {trips:.4g} times: {100.0 * e / total:5.1f} %  {hint}")
