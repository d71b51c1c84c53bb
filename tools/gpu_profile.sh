#!/bin/bash
# ncu launch lists (my kernels only; the synthetic-data rendering of bench.py is torch and is skipped by name)
# and one --set full capture of the dominant kernel per mode.  Run only after the same commands exited 0 without ncu.
mkdir -p gpurun_out
K='regex:carve_kernel|average_kernel|pack_bits|expand_state|unpad_f32|fill_f32'
for mode in carving averaging; do
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file gpurun_out/launches_$mode.csv \
  python bench.py --mode $mode --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_l_$mode.log 2>&1
done
ncu --set full --clock-control none --import-source on -k regex:carve_kernel -s 3 -c 1 -o gpurun_out/prof_carve -f \
  python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_f.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:average_kernel -s 3 -c 1 -o gpurun_out/prof_avg -f \
  python bench.py --mode averaging --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_fa.log 2>&1
ls -la gpurun_out
