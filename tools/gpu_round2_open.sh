#!/bin/bash
# First gpurun call of the next round (the round-1 GPU minutes ran out before these could be run):
#   1. the default GPU suite + smoke (nothing regressed by the PackJob change of the view-rectangle commit),
#   2. the gated view-rectangle parity tests (BP_VIEW_RECT=1, tests/test_gpu_view_rect.py),
#   3. the default bench with and without view rectangles (A/B of the e2e leg and of the pack), and the carving
#      bench with the speculative read-back (BP_SPEC_READBACK=1, tests/test_gpu_spec_readback.py) alone and with both,
#   4. compute-sanitizer memcheck / racecheck / initcheck of one small carve + average (the fused
#      kernels stage through shared memory and shuffle survivors between lanes) - smoke() is the workload.
# Usage: gpurun --timeout 2400 -- 'bash tools/gpu_round2_open.sh'   (about 20 GPU-minutes)
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
BP_EXPERIMENTAL=1 python -m pytest tests/test_gpu_view_rect.py -m gpu -q > gpurun_out/pytest_view_rect.log 2>&1
echo "view-rect pytest rc=$?" | tee -a gpurun_out/pytest_view_rect.log
BP_EXPERIMENTAL=1 python -m pytest tests/test_gpu_spec_readback.py -m gpu -q > gpurun_out/pytest_spec_readback.log 2>&1
echo "spec-readback pytest rc=$?" | tee -a gpurun_out/pytest_spec_readback.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
BP_VIEW_RECT=1 python bench.py > gpurun_out/bench_view_rect.json 2> gpurun_out/bench_view_rect.err; echo "bench (rect) rc=$?"
BP_SPEC_READBACK=1 python bench.py --no-averaging --no-cpu > gpurun_out/bench_spec.json 2> gpurun_out/bench_spec.err; echo "bench (spec) rc=$?"
BP_SPEC_READBACK=1 BP_VIEW_RECT=1 python bench.py --no-averaging --no-cpu > gpurun_out/bench_spec_rect.json 2> gpurun_out/bench_spec_rect.err; echo "bench (spec+rect) rc=$?"
# round-2 carve path: block cull on / off, and how many leading views to bit-pack (device-resident masks)
for cfg in "BP_CARVE_CULL=0 BP_PACK_HEAD=-1" "BP_CARVE_CULL=1 BP_PACK_HEAD=-1" "BP_CARVE_CULL=1 BP_PACK_HEAD=16" "BP_CARVE_CULL=1 BP_PACK_HEAD=32" "BP_CARVE_CULL=1 BP_PACK_HEAD=64" "BP_CARVE_CULL=0 BP_PACK_HEAD=32"; do
  tag=$(echo $cfg | tr ' =' '__')
  env $cfg python bench.py --no-averaging --no-cpu --no-e2e > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench $cfg rc=$?"
done
BP_COUNT_WORK=1 python bench.py --no-averaging --no-cpu --no-e2e --steps 1 > gpurun_out/bench_count_work.json 2> gpurun_out/bench_count_work.err; echo "bench count rc=$?"
for tool in memcheck racecheck initcheck; do
  timeout 400 compute-sanitizer --tool $tool --error-exitcode 9 python __graft_entry__.py smoke \
    > gpurun_out/sanitizer_$tool.log 2>&1
  echo "compute-sanitizer $tool rc=$?" | tee -a gpurun_out/sanitizer_$tool.log
  tail -4 gpurun_out/sanitizer_$tool.log
done
tail -3 gpurun_out/pytest_gpu.log gpurun_out/pytest_view_rect.log gpurun_out/pytest_spec_readback.log; cat gpurun_out/smoke.log
cat gpurun_out/bench_default.json gpurun_out/bench_view_rect.json gpurun_out/bench_spec.json gpurun_out/bench_spec_rect.json
for f in gpurun_out/bench_BP_*.json gpurun_out/bench_count_work.json; do echo $f; python -c "import json,sys; d=json.load(open(sys.argv[1])); print(d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['roofline']['ingest_ms_per_step'], d.get('work'))" $f; done
