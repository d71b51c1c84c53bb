#!/bin/bash
# Round 2, thirteenth (and final-evidence) GPU call (1 GPU): whole GPU suite, smoke, default bench line, reference arm, launch list, and fresh
# --set full captures of carve_items_kernel, carve_segments_kernel and average_kernel<u8> on the final kernels.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log; tail -4 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?"; tail -3 gpurun_out/bench_default.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "bench reference rc=$?"
B="python bench.py --no-cpu --no-e2e --no-extra"
K='regex:carve_|classify_items|decode_items|cull_blocks|compact_live|tile_any|tile_sat|pack_bits|average_kernel|map_any|avg_skip|expand_state|fill_f32'
$B --steps 2 --warmup 3 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file gpurun_out/launches_default.csv \
  $B --steps 2 --warmup 3 > gpurun_out/ncu_l.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:carve_items_kernel|carve_segments_kernel' -s 6 -c 2 -o gpurun_out/prof_carve_r2c -f \
  $B --no-averaging --steps 1 --warmup 3 > gpurun_out/ncu_f.log 2>&1
echo "ncu full carve rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:average_kernel' -s 3 -c 1 -o gpurun_out/prof_avg_r2c -f \
  $B --mode averaging --steps 1 --warmup 3 > gpurun_out/ncu_fa.log 2>&1
echo "ncu full avg rc=$?"
tail -2 gpurun_out/ncu_l.log gpurun_out/ncu_f.log gpurun_out/ncu_fa.log | cut -c1-300
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_default.json'))
print('N1 carve ms', d['ms_per_step'], 'kernel', d['roofline']['kernel_ms_per_launch'], 'ingest', d['roofline']['ingest_ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['oracle_check'], d['work'], 'frac', d['roofline']['frac'])
a=d['averaging']; print('N1 avg ms', a['ms_per_step'], 'e2e', a['e2e']['ms_per_step'], a['oracle_check'], a['work'], 'frac', a['roofline']['frac'])
for k,c in d['configs'].items(): print(k, c['ms_per_step'], c['e2e']['ms_per_step'], c['oracle_check'], c['golden']['all'])
print(open('gpurun_out/bench_reference.json').read()[:600])
PY
