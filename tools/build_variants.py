#!/usr/bin/env python3
"""Compile-time A/B variants of libbp_b200.so (same sources, extra -D switches) into variants/libbp_b200_<name>.so;
select one at run time with BP_B200_LIB=<path>.  `python tools/build_variants.py name:-DX=1,-DY=2 ...`
Prints the register / spill line of the two persistent carve kernels for each variant."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

os.makedirs(os.path.join(ROOT, "variants"), exist_ok=True)
h_all, h_k = ge.source_hashes()
for spec in sys.argv[1:]:
    name, _, defs = spec.partition(":")
    defs = [d for d in defs.split(",") if d]
    out = os.path.join(ROOT, "variants", f"libbp_b200_{name}.so")
    r = subprocess.run([ge._nvcc(), *ge.NVCC_FLAGS, "-Xptxas", "-v", *defs, f'-DBP_SRC_HASH="{h_all}"', f'-DBP_KERN_HASH="{h_k}"',
                        "-I" + os.path.join(ROOT, "include"), "-I" + ge.CSRC, "-o", out,
                        *[os.path.join(ge.CSRC, f) for f in ge.SOURCES]], capture_output=True, text=True, check=True)
    lines = r.stderr.splitlines()
    print(name, defs)
    for i, l in enumerate(lines):
        m = re.search(r"Compiling entry function '(\S+)'", l)
        if m and re.search(r"carve_(items|segments)_kernelILb0", m.group(1)):
            print("   ", re.search(r"carve_\w+_kernel", m.group(1)).group(0), "|", lines[i + 2].strip(), "|", lines[i + 3].strip()[:60])
