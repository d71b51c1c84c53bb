#!/bin/bash
# Round 2, third GPU call (1 GPU): the whole GPU suite on the two-pass cull, smoke, the default bench line, the launch
# list of one carving step and of one averaging step, and fresh --set full captures of carve_items_kernel,
# cull_blocks_kernel and average_kernel<u8> (each only after the same command exited 0 without ncu).
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log; tail -8 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?"; tail -3 gpurun_out/bench_default.err
B="python bench.py --no-cpu --no-e2e --no-extra"
K='regex:carve_|cull_blocks|tile_any|tile_sat|pack_bits|average_kernel|expand_state|fill_f32'
$B --steps 2 --warmup 3 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 300 --csv --log-file gpurun_out/launches_default.csv \
  $B --steps 2 --warmup 3 > gpurun_out/ncu_l.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:carve_items_kernel|cull_blocks_kernel' -s 9 -c 3 -o gpurun_out/prof_carve_r2b -f \
  $B --no-averaging --steps 1 --warmup 3 > gpurun_out/ncu_f.log 2>&1
echo "ncu full carve rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:average_kernel' -s 3 -c 1 -o gpurun_out/prof_avg_r2 -f \
  $B --mode averaging --steps 1 --warmup 3 > gpurun_out/ncu_fa.log 2>&1
echo "ncu full avg rc=$?"
tail -2 gpurun_out/ncu_l.log gpurun_out/ncu_f.log gpurun_out/ncu_fa.log | cut -c1-300
ls -la gpurun_out
