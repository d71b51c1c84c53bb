#!/bin/bash
# One gpurun call at the end of a session: full parity suite, smoke, default bench, the tools of the 8f rows, and an
# ncu launch list of those tools (kernel names + durations as evidence; numbers under ncu are not bench values).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
python tests/bench/bench_masks.py > gpurun_out/bench_masks.json 2> gpurun_out/bench_masks.err
python tests/bench/bench_score_points.py 200000 > gpurun_out/bench_score_points.json 2> gpurun_out/bench_score.err
python tests/bench/bench_task_tail.py > gpurun_out/bench_task_tail.json 2> gpurun_out/bench_tail.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_8f.csv \
  -k regex:'minmax_kernel|threshold_kernel|dilate_kernel|expand_bits|score_points|finish_scores|tail_kernel' \
  python __graft_entry__.py smoke > gpurun_out/ncu_8f.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log; cat gpurun_out/bench_default.json; cat gpurun_out/bench_masks.json gpurun_out/bench_score_points.json gpurun_out/bench_task_tail.json
