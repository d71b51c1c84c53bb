#!/bin/bash
# Round 2, call 19 (1 GPU): C1 (the reference's own 18-view configuration) against the cull-view and pack-head defaults;
# then the default bench line with the C2 / C4 single-GPU sub-objects.
mkdir -p gpurun_out
for cv in 3 16; do for hd in 16 -1; do
  BP_CULL_VIEWS=$cv BP_PACK_HEAD=$hd python bench.py --workload C1 --no-cpu --no-e2e --no-extra --no-averaging --steps 20 > gpurun_out/c1_$cv_$hd.json 2>/dev/null
  python -c "import json,sys; d=json.load(open(sys.argv[1])); print('C1 cull_views', sys.argv[2], 'pack_head', sys.argv[3], 'ms', round(d['ms_per_step'],4), 'kernel', round(d['roofline']['kernel_ms_per_launch'],4), 'ingest', round(d['roofline']['ingest_ms_per_step'],4))" gpurun_out/c1_$cv_$hd.json $cv $hd
done; done
for cv in 3 16; do
  BP_CULL_VIEWS=$cv python bench.py --workload C2 --no-cpu --no-e2e --no-extra --no-averaging --steps 20 > gpurun_out/c2_$cv.json 2>/dev/null
  python -c "import json,sys; d=json.load(open(sys.argv[1])); print('C2 cull_views', sys.argv[2], 'ms', round(d['ms_per_step'],4), 'kernel', round(d['roofline']['kernel_ms_per_launch'],4))" gpurun_out/c2_$cv.json $cv
done
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?"; tail -2 gpurun_out/bench_default.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_default.json'))
print('N1 carve ms', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['oracle_check'])
for k,c in d['configs'].items(): print(k, round(c['ms_per_step'],4), c['value'], round(c['e2e']['ms_per_step'],2), c['oracle_check'], (c.get('golden') or {}).get('all'), c['roofline']['frac'])
PY
