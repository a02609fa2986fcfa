TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
A="--gpus 2 --model mnist_mlp --batch 32 --steps 500 --warmup 50 --no-bucket-timing --no-parity --no-allreduce-section"
for v in "fused:" "plain:--no-fused-opt" "torchreducer:--reducer torch"; do n=${v%%:*}; f=${v#*:}; timeout 120 $TR --master-port 29591 bench.py $A $f > gpurun_out/r2_mlp_n2_b2c_$n.json 2>/dev/null; python -c "
import json,sys; d=json.loads(open('gpurun_out/r2_mlp_n2_b2c_$n.json').read().strip().splitlines()[-1]); print('$n', round(d['ms_per_step'],4), d['config']['reducer'], d['config']['fused_optimizer'], d['gpu_launches'])"; done
timeout 120 $TR --master-port 29593 bench.py --gpus 2 --impl nccl --model mnist_mlp --batch 32 --steps 500 --warmup 50 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('nccl', round(d['ms_per_step'],4))"
