#!/usr/bin/env python3
"""List the loops (backward branches) of one kernel in libbp_b200.so with their SASS instruction count and
opcode histogram: `python tools/sass_loops.py average_kernelIhE [min_len]`.  Used to count instructions per
voxel-view before spending GPU time."""
import collections
import re
import subprocess
import sys

so = "plant-3d-vision_b200/libbp_b200.so"
pat = sys.argv[1]
min_len = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
for f in funcs[1:]:
    name = f.split("\n", 1)[0]
    if pat not in name:
        continue
    ins = re.findall(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);", f)
    addr = [int(a, 16) for a, _ in ins]
    print(name, len(ins), "instructions")
    for n, (a, s) in enumerate(ins):
        m = re.search(r"\bBRA(?:\.U)?\b.*?(0x[0-9a-f]+)", s)
        if m and int(m.group(1), 16) <= int(a, 16):
            t = int(m.group(1), 16)
            lo = addr.index(t) if t in addr else None
            if lo is None or n - lo + 1 < min_len:
                continue
            body = [x[1] for x in ins[lo:n + 1]]
            ops = collections.Counter((b.split()[1] if b.startswith("@") else b.split()[0]).split(".")[0] for b in body)
            print(f"  loop [{lo}, {n}] {len(body)} instrs:", dict(ops.most_common()))
