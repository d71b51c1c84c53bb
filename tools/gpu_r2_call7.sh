#!/bin/bash
# Round 2, seventh GPU call (1 GPU): view segments of the carve path - parity (whole GPU suite), the per-rank
# scheduling probe at world sizes 1 / 2 / 4 / 8, the default bench line, and the skip-block A/B of the averaging path.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
python tools/stripe_probe.py 1,2,4,8 > gpurun_out/stripe_probe.txt 2> gpurun_out/stripe_probe.err; echo "probe rc=$?"; cat gpurun_out/stripe_probe.txt; tail -3 gpurun_out/stripe_probe.err
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?"; tail -3 gpurun_out/bench_default.err
for k in 8 16 32 64; do
  BP_AVG_BLK_K=$k python bench.py --mode averaging --no-cpu --no-e2e --no-extra > gpurun_out/bench_avg_blk$k.json 2> gpurun_out/bench_avg_blk$k.err
  python -c "import json,sys; d=json.load(open(sys.argv[1])); print('avg blk_k', sys.argv[2], 'ms', d['ms_per_step'], 'evaluated', d['work']['evaluated_fraction'])" gpurun_out/bench_avg_blk$k.json $k
done
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_default.json'))
print('N1 carve ms', d['ms_per_step'], 'kernel', d['roofline']['kernel_ms_per_launch'], 'ingest', d['roofline']['ingest_ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['oracle_check'], d['work'])
a=d['averaging']; print('N1 avg ms', a['ms_per_step'], 'e2e', a['e2e']['ms_per_step'], a['oracle_check'], a['work'], 'frac', a['roofline']['frac'])
for k,c in d['configs'].items(): print(k, c['ms_per_step'], c['e2e']['ms_per_step'], c['oracle_check'], c['golden']['all'])
PY
