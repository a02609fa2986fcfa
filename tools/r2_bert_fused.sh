#!/bin/bash
# BERT-base fp32 through the RayAccelerator worker path: own reducer + fused AdamW vs stock NCCL DDP + torch AdamW (N from $1).
cd "$(dirname "$0")/.."
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
A="--gpus $N --model bert_base --batch 32 --steps 10 --warmup 4 --no-allreduce-section"
for v in "fused:--no-parity" "plain:--no-parity --no-fused-opt"; do n=${v%%:*}; f=${v#*:}
  timeout 200 $TR --master-port 29595 bench.py $A $f > gpurun_out/r2_bert_f32_n${N}_b2c_$n.json 2> gpurun_out/r2_bert_f32_n${N}_b2c_$n.err
  python -c "
import json; d=json.loads(open('gpurun_out/r2_bert_f32_n${N}_b2c_$n.json').read().strip().splitlines()[-1]); print('$n', round(d['value'],1), round(d['ms_per_step'],3), d['e2e']['last_loss'], d['config']['fused_optimizer'], d['gpu_launches'])" || tail -5 gpurun_out/r2_bert_f32_n${N}_b2c_$n.err
done
timeout 200 $TR --master-port 29597 bench.py $A --impl nccl > gpurun_out/r2_bert_f32_n${N}_nccl_ddp.json 2>/dev/null
python -c "
import json; d=json.loads(open('gpurun_out/r2_bert_f32_n${N}_nccl_ddp.json').read().strip().splitlines()[-1]); print('nccl', round(d['value'],1), round(d['ms_per_step'],3), d['e2e']['last_loss'])"
