#!/bin/bash
# Round 2, sixth GPU call (1 GPU): scheduling A/B of the carve path per rank at world sizes 1 / 2 / 4 / 8 (rank 0's
# striped handle alone), and the cull / view-skipping tests.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_cull.py -m gpu -q > gpurun_out/pytest_gpu_cull.log 2>&1; echo "pytest cull rc=$?"; tail -5 gpurun_out/pytest_gpu_cull.log
python tools/stripe_probe.py 1,2,4,8 > gpurun_out/stripe_probe.txt 2> gpurun_out/stripe_probe.err; echo "probe rc=$?"; cat gpurun_out/stripe_probe.txt; tail -3 gpurun_out/stripe_probe.err
