#!/bin/bash
# Round 2, second GPU call: the new carve path (block cull + two-tier masks).
#   1. its own parity tests, 2. A/B of the cull and of the packed head on the device-resident C5 step,
#   3. the projections actually evaluated (counting variant), 4. ncu launch list + one --set full capture of the two
#   new kernels (only after the same command exited 0 without ncu).
mkdir -p gpurun_out
python -m pytest tests/test_gpu_cull.py -m gpu -q > gpurun_out/pytest_gpu_cull.log 2>&1; echo "pytest cull rc=$?" | tee -a gpurun_out/pytest_gpu_cull.log
tail -15 gpurun_out/pytest_gpu_cull.log
B="python bench.py --no-averaging --no-cpu --no-e2e --no-extra"
for cfg in "BP_CARVE_CULL=0 BP_PACK_HEAD=-1" "BP_CARVE_CULL=1 BP_PACK_HEAD=-1" "BP_CARVE_CULL=1 BP_PACK_HEAD=8" "BP_CARVE_CULL=1 BP_PACK_HEAD=16" "BP_CARVE_CULL=1 BP_PACK_HEAD=32" "BP_CARVE_CULL=1 BP_PACK_HEAD=64" "BP_CARVE_CULL=0 BP_PACK_HEAD=32"; do
  tag=$(echo $cfg | tr ' =' '__')
  env $cfg $B > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench $cfg rc=$?"
done
for f in gpurun_out/bench_BP_*.json; do echo $f; python -c "import json,sys; d=json.load(open(sys.argv[1])); print(d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['roofline']['ingest_ms_per_step'], d.get('work'))" $f; done
# the whole default line (every leg of the new bench.py) and the reference arm, as the driver runs them
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?"; tail -3 gpurun_out/bench_default.err
python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "bench reference rc=$?"; tail -3 gpurun_out/bench_reference.err
gcc -O3 -fopenmp tools/host_bw_probe.c -o /tmp/hbw && /tmp/hbw > gpurun_out/host_bw_probe.txt 2>&1; cat gpurun_out/host_bw_probe.txt; nproc; lscpu | grep -E "Model name|Socket|NUMA node\(s\)" 
K='regex:carve_|cull_blocks|tile_any|tile_sat|pack_bits'
$B --steps 2 --warmup 3 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file gpurun_out/launches_carving.csv \
  $B --steps 2 --warmup 3 > gpurun_out/ncu_l.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:carve_items_kernel|cull_blocks_kernel' -s 6 -c 2 -o gpurun_out/prof_carve_r2 -f \
  $B --steps 1 --warmup 3 > gpurun_out/ncu_f.log 2>&1
echo "ncu full rc=$?"
tail -3 gpurun_out/ncu_l.log gpurun_out/ncu_f.log
ls -la gpurun_out
