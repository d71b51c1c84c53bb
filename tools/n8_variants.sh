#!/bin/bash
# carving-only N-GPU step time for a few (split, stripe block) variants; usage: n8_variants.sh N
N=${1:-8}
for v in "0 3" "1 5" "0 5" "1 4"; do
  set -- $v
  BP_CARVE_SPLIT=$1 BP_STRIPE_LG=$2 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-averaging --no-e2e --no-cpu \
    > gpurun_out/var_${1}_${2}.json 2> gpurun_out/var_${1}_${2}.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/var_${1}_${2}.json").read().strip().splitlines()[-1])
print("split=$1 lg=$2", round(d["ms_per_step"],3), "ms  kernel", round(d["roofline"]["kernel_ms_per_launch"],3), "ingest", round(d["roofline"]["ingest_ms_per_step"],3))
PY
done
