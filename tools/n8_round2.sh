#!/bin/bash
# N-GPU carving step (masks resident in HBM) with and without view rectangles, for a few stripe blocks;
# usage (after tools/gpu_round2_open.sh is green):  gpurun --gpus 8 --timeout 900 -- 'bash tools/n8_round2.sh 8'
N=${1:-8}
mkdir -p gpurun_out
for v in "0 3" "1 3" "1 2" "1 1"; do
  set -- $v
  BP_VIEW_RECT=$1 BP_STRIPE_LG=$2 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-averaging --no-cpu \
    > gpurun_out/n${N}_rect${1}_lg${2}.json 2> gpurun_out/n${N}_rect${1}_lg${2}.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/n${N}_rect${1}_lg${2}.json").read().strip().splitlines()[-1])
e=d.get("e2e") or {}
print("rect=$1 lg=$2", round(d["ms_per_step"],3), "ms  kernel", round(d["roofline"]["kernel_ms_per_launch"],3),
      "ingest", round(d["roofline"]["ingest_ms_per_step"],3), " e2e", round(e.get("ms_per_step",0),1), "ms")
PY
done
