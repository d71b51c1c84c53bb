#!/bin/bash
# Round 2, fourth GPU call (2 GPUs): the whole GPU suite including the 2-GPU NCCL tests, the work-list order and
# packed-head A/B on one GPU, and the default bench line at N = 2 as the driver launches it.
mkdir -p gpurun_out
nvidia-smi -L
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_2gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_2gpu.log
tail -15 gpurun_out/pytest_gpu_2gpu.log
B="python bench.py --no-averaging --no-cpu --no-e2e --no-extra"
for cfg in "BP_LIVE_ORDER=0 BP_PACK_HEAD=16" "BP_LIVE_ORDER=1 BP_PACK_HEAD=16" "BP_LIVE_ORDER=0 BP_PACK_HEAD=-1" "BP_LIVE_ORDER=1 BP_PACK_HEAD=-1" "BP_LIVE_ORDER=1 BP_PACK_HEAD=32"; do
  tag=$(echo $cfg | tr ' =' '__')
  env $cfg $B > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench $cfg rc=$?"
  python -c "import json,sys; d=json.load(open(sys.argv[1])); print(d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['roofline']['ingest_ms_per_step'])" gpurun_out/bench_$tag.json
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 \
  > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"; tail -5 gpurun_out/bench_n2.err
python -c "
import json; d=json.load(open('gpurun_out/bench_n2.json'))
print('N2 carve ms', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['oracle_check'])
print('N2 avg ms', d['averaging']['ms_per_step'], 'e2e', d['averaging']['e2e']['ms_per_step'], d['averaging']['oracle_check'])
v=d['view_sharded']; print('N2 views C4 ms', v['ms_per_step'], 'e2e', v['e2e']['ms_per_step'], v['oracle_check'], v['allreduce'])
"
ls -la gpurun_out
