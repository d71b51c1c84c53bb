#!/bin/bash
# One gpurun call: parity suite, smoke, both bench modes, ncu launch list + full capture.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
python bench.py > gpurun_out/bench_carving.json 2> gpurun_out/bench_carving.err; echo "bench carving rc=$?"
python bench.py --mode averaging > gpurun_out/bench_averaging.json 2> gpurun_out/bench_averaging.err; echo "bench averaging rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "bench ref rc=$?"
if [ "$1" = "ncu" ]; then
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_carving.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_l.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_averaging.csv \
  python bench.py --mode averaging --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_la.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:carve_kernel -s 3 -c 1 -o gpurun_out/prof_carve -f \
  python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_f.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:average_kernel -s 3 -c 1 -o gpurun_out/prof_avg -f \
  python bench.py --mode averaging --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_fa.log 2>&1
fi
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log; cat gpurun_out/bench_carving.json
