#!/bin/bash
# The driver's N-GPU bench line, as the driver launches it.  usage: gpurun --gpus N --timeout 1200 -- 'bash tools/gpu_r2_scale.sh N'
N=${1:-8}
mkdir -p gpurun_out
if [ "$N" = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"; fi
$L bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"; tail -3 gpurun_out/bench_n$N.err | cut -c1-300
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n$N.json").read().strip().splitlines()[-1])
r=d["roofline"]; e=d.get("e2e") or {}
print("N=$N carve ms", round(d["ms_per_step"],3), "value", d["value"], "kernel", round(r["kernel_ms_per_launch"],3), "ingest", round(r["ingest_ms_per_step"],3),
      "e2e ms", round(e.get("ms_per_step",0),2), "e2e value", e.get("value"), d.get("oracle_check"), d.get("clocks"))
a=d.get("averaging")
if a: print("avg ms", round(a["ms_per_step"],2), "e2e", round((a.get("e2e") or {}).get("ms_per_step",0),2), a.get("oracle_check"))
v=d.get("view_sharded")
if v: print("view_sharded", json.dumps(v)[:900])
PY
