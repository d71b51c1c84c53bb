#!/bin/bash
# Round 2, twelfth GPU call (1 GPU): view segments drawn one per atomic from a compact list (class 1 first).
mkdir -p gpurun_out
python -m pytest tests/test_gpu_cull.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/pytest_gpu_cull.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_cull.log
echo "== default"; PROBE_SPLITS=8,16,32,64,1000 python tools/stripe_probe.py 1,2,4,8 2> gpurun_out/probe12_default.err | tee gpurun_out/probe12_default.txt; tail -2 gpurun_out/probe12_default.err
echo "== variant S256"; BP_B200_LIB=$PWD/variants/libbp_b200_S256.so PROBE_SPLITS=16,32 python tools/stripe_probe.py 1,2,4,8 2> gpurun_out/probe12_S256.err | tee gpurun_out/probe12_S256.txt
BP_COUNT_WORK=1 python - <<'PY'
# projections evaluated with and without segments (work inflation of concurrent segments)
import os, numpy as np, torch
from plant3dvision_b200 import synth
from plant3dvision_b200.cl import Backprojection
w = synth.workload("C5", 1.0)
masks = synth.render_masks(w["cams"], w["capsules"], w["H"], w["W"]).cuda()
n = len(w["cams"]); hw = np.tile(np.array([[w["H"], w["W"]]], dtype=np.int32), (n, 1))
for split in (1000, 16, 32):
    os.environ["BP_VIEW_SPLIT"] = str(split)
    bp = Backprojection(w["shape"], w["origin"], w["voxel_size"])
    bp.process_device_views(w["cams"].astype(np.float32), [masks[v].data_ptr() for v in range(n)], hw)
    bp.sync()
    print("split", split, "work_stats", bp.work_stats(), flush=True)
    bp.close()
PY
