#!/bin/bash
# Round 2, eighth GPU call (1 GPU): compile-time A/B of the launch shape of the two persistent carve kernels
# (tools/build_variants.py; variants/ travels with the snapshot), per-rank probe at world sizes 1 / 2 / 4 / 8.
mkdir -p gpurun_out
for v in A B C D E; do
  echo "== variant $v"
  BP_B200_LIB=$PWD/variants/libbp_b200_$v.so PROBE_SPLITS=16,32,1000 python tools/stripe_probe.py 1,2,4,8 2> gpurun_out/probe_$v.err | tee gpurun_out/probe_$v.txt
  tail -2 gpurun_out/probe_$v.err
done
python -m pytest tests/test_gpu_cull.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/pytest_gpu_cull.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_cull.log
