#!/bin/bash
# Round 2, fifth GPU call (2 GPUs): the whole GPU suite (view skipping of the averaging path, one-patch grabs, two-ended
# work list), the default bench line at N = 1 and N = 2, and the averaging A/B of BP_AVG_SKIP.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_2gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_2gpu.log
tail -15 gpurun_out/pytest_gpu_2gpu.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?"; tail -3 gpurun_out/bench_default.err
BP_AVG_SKIP=0 python bench.py --mode averaging --no-cpu --no-e2e --no-extra > gpurun_out/bench_avg_noskip.json 2> gpurun_out/bench_avg_noskip.err; echo "bench avg noskip rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_default.json'))
print('N1 carve ms', d['ms_per_step'], 'kernel', d['roofline']['kernel_ms_per_launch'], 'ingest', d['roofline']['ingest_ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['oracle_check'])
a=d['averaging']; print('N1 avg ms', a['ms_per_step'], 'e2e', a['e2e']['ms_per_step'], a['oracle_check'], a['work'], 'frac', a['roofline']['frac'])
for k,c in d['configs'].items(): print(k, c['ms_per_step'], c['e2e']['ms_per_step'], c['oracle_check'], c['golden']['all'], c.get('work'))
n=json.load(open('gpurun_out/bench_avg_noskip.json')); print('avg noskip ms', n['ms_per_step'])
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 \
  > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"; tail -5 gpurun_out/bench_n2.err
python - <<'PY'
import json; d=json.load(open('gpurun_out/bench_n2.json'))
print('N2 carve ms', d['ms_per_step'], 'kernel', d['roofline']['kernel_ms_per_launch'], 'e2e', d['e2e']['ms_per_step'], d['oracle_check'])
print('N2 avg ms', d['averaging']['ms_per_step'], 'e2e', d['averaging']['e2e']['ms_per_step'], d['averaging']['oracle_check'])
v=d['view_sharded']; print('N2 views C4 ms', v['ms_per_step'], 'e2e', v['e2e']['ms_per_step'], v['oracle_check'], v['allreduce'])
PY
ls -la gpurun_out
