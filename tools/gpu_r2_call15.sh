#!/bin/bash
# Round 2, call 15 (1 GPU): lazy clear (the cull's first pass seeds the state) and 16 packed views on every handle.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
for lz in 1 0; do echo "== BP_LAZY_CLEAR=$lz"; BP_LAZY_CLEAR=$lz PROBE_SPLITS=16 python tools/stripe_probe.py 1,2,4,8 2>/dev/null; done | tee gpurun_out/probe15.txt
python bench.py --no-cpu --no-extra > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_quick.json'))
print('N1 carve ms', d['ms_per_step'], 'kernel', d['roofline']['kernel_ms_per_launch'], 'ingest', d['roofline']['ingest_ms_per_step'], 'e2e', d['e2e']['ms_per_step'], 'frac', d['roofline']['frac'], d['roofline'].get('ncu'))
a=d['averaging']; print('N1 avg ms', a['ms_per_step'], 'e2e', a['e2e']['ms_per_step'])
PY
