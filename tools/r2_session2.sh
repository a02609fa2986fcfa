#!/bin/bash
# Round-2 2-GPU dress rehearsal of everything the 8-GPU session B runs (cheap: 2 GPUs).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
(timeout 200 $TR --master-port 29531 tools/nvlink_bytes.py --launches 100 --out gpurun_out/r2_nvlink_bytes_w2.json > gpurun_out/r2_nvlink_bytes_w2.log 2>&1; echo "nvlink_bytes rc=$?"); tail -2 gpurun_out/r2_nvlink_bytes_w2.log | cut -c1-300
(timeout 200 $TR --master-port 29533 bench.py --gpus 2 --model mnist_mlp --batch 32 --steps 400 --warmup 50 --no-bucket-timing --no-parity --no-allreduce-section > gpurun_out/r2_mlp_n2_b2c.json 2> gpurun_out/r2_mlp_n2_b2c.err; echo "mlp b2c rc=$?"); tail -2 gpurun_out/r2_mlp_n2_b2c.err | cut -c1-300
(timeout 200 $TR --master-port 29535 bench.py --gpus 2 --impl nccl --model mnist_mlp --batch 32 --steps 400 --warmup 50 > gpurun_out/r2_mlp_n2_nccl.json 2> gpurun_out/r2_mlp_n2_nccl.err; echo "mlp nccl rc=$?"); tail -2 gpurun_out/r2_mlp_n2_nccl.err | cut -c1-300
(timeout 240 $TR --master-port 29537 bench.py --gpus 2 --impl hvd --model bert_base --param-dtype bf16 --batch 32 --steps 8 --warmup 3 > gpurun_out/r2_bert_n2_b2c.json 2> gpurun_out/r2_bert_n2_b2c.err; echo "bert b2c rc=$?"); tail -2 gpurun_out/r2_bert_n2_b2c.err | cut -c1-300
(timeout 240 python tools/tune_4x2.py --trials 1 --steps 500 --out gpurun_out/r2_tune_1x2.json > gpurun_out/r2_tune_1x2.log 2>&1; echo "tune rc=$?"); tail -3 gpurun_out/r2_tune_1x2.log | cut -c1-300
(timeout 300 $TR --master-port 29539 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/r2_bench_n2_dev.json 2> gpurun_out/r2_bench_n2_dev.err; echo "bench2 rc=$?"); tail -2 gpurun_out/r2_bench_n2_dev.err | cut -c1-300
