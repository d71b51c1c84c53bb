#!/bin/bash
# Round 2, eleventh GPU call (1 GPU): bp_tail_static (no-allocation tail) - its test alone in a fresh process, then after
# the whole GPU suite; and a per-kernel launch list of the per-rank probe at world 8 (where do the segments lose 3 ms?).
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "no_alloc alone rc=$?"; tail -3 gpurun_out/pytest_no_alloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all rc=$?"; tail -4 gpurun_out/pytest_gpu.log
PROBE_SPLITS=16 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_probe_w8.csv python tools/stripe_probe.py 8 > gpurun_out/ncu_probe.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/launches_probe_w8.csv')) if len(r) > 10]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); ui = hdr.index('Metric Unit')
agg = collections.defaultdict(list)
for r in rows[1:]:
    v = float(r[vi].replace(',', '')); u = r[ui]
    ms = v / 1e6 if u in ('ns', 'nsecond') else v / 1e3 if u in ('us', 'usecond') else v if u in ('ms', 'msecond') else v * 1e3
    agg[r[ki][:60]].append(ms)
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"{k:60s} n={len(v):3d} avg_ms={sum(v)/len(v):8.4f} max_ms={max(v):8.4f}")
PY
