#!/bin/bash
# Last call of a round (1 GPU): what the driver runs at round end - the GPU suite, smoke, the default bench line.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?"; tail -2 gpurun_out/bench_default.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_default.json'))
print('N1 carve ms', d['ms_per_step'], 'kernel', d['roofline']['kernel_ms_per_launch'], 'e2e', d['e2e']['ms_per_step'], d['oracle_check'], 'frac', d['roofline']['frac'], 'ncu stale', (d['roofline'].get('ncu') or {}).get('stale'))
a=d['averaging']; print('N1 avg ms', a['ms_per_step'], 'e2e', a['e2e']['ms_per_step'], a['oracle_check'])
for k,c in d['configs'].items(): print(k, c['ms_per_step'], c['value'], c['e2e']['ms_per_step'], c['oracle_check'], (c.get('golden') or {}).get('all'), c['roofline']['frac'])
PY
