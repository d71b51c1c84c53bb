#!/bin/bash
# Round 2, tenth GPU call (1 GPU): view segments rebuilt (classify / segments with atomic AND / decode as separate
# kernels; carve_items_kernel back to its warp-convergent form).  Parity first, then the per-rank probe for the default
# build, two launch-shape variants and the library of commit 61f5274 (no segments) on the same box.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_cull.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/pytest_gpu_cull.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_cull.log
echo "== default"; PROBE_SPLITS=8,16,32,1000 python tools/stripe_probe.py 1,2,4,8 2> gpurun_out/probe10_default.err | tee gpurun_out/probe10_default.txt; tail -2 gpurun_out/probe10_default.err
for v in S256 I128 OLD; do
  echo "== variant $v"
  BP_B200_LIB=$PWD/variants/libbp_b200_$v.so PROBE_SPLITS=16,32 python tools/stripe_probe.py 1,2,4,8 2> gpurun_out/probe10_$v.err | tee gpurun_out/probe10_$v.txt
  tail -2 gpurun_out/probe10_$v.err
done
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all rc=$?"; tail -4 gpurun_out/pytest_gpu.log
