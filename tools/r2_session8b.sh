#!/bin/bash
# Round-2 8-GPU session B: full GPU test suite, final bench line, configs #2 / #3 / #5 at their stated sizes,
# NVLink byte counters.  Every part has its own timeout and log; nothing runs under a profiler.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
(timeout 420 python -m pytest tests -m gpu -q -rs > gpurun_out/r2_pytest_gpu_n8.log 2>&1; echo "pytest n8 rc=$?"); tail -3 gpurun_out/r2_pytest_gpu_n8.log | cut -c1-200
(timeout 400 $TR --master-port 29541 bench.py --gpus 8 --steps 10 --warmup 5 > gpurun_out/r2_bench_n8_b2c.json 2> gpurun_out/r2_bench_n8_b2c.err; echo "bench8 b2c rc=$?"); tail -2 gpurun_out/r2_bench_n8_b2c.err | cut -c1-200
(timeout 200 $TR --master-port 29543 bench.py --gpus 8 --impl nccl --steps 10 --warmup 5 > gpurun_out/r2_bench_n8_nccl.json 2> gpurun_out/r2_bench_n8_nccl.err; echo "bench8 nccl rc=$?")
(timeout 150 $TR --master-port 29545 bench.py --gpus 8 --model mnist_mlp --batch 32 --steps 500 --warmup 50 --no-bucket-timing --no-parity --no-allreduce-section > gpurun_out/r2_mlp_n8_b2c.json 2> gpurun_out/r2_mlp_n8_b2c.err; echo "mlp8 b2c rc=$?"); tail -2 gpurun_out/r2_mlp_n8_b2c.err | cut -c1-200
(timeout 150 $TR --master-port 29547 bench.py --gpus 8 --impl nccl --model mnist_mlp --batch 32 --steps 500 --warmup 50 > gpurun_out/r2_mlp_n8_nccl.json 2> gpurun_out/r2_mlp_n8_nccl.err; echo "mlp8 nccl rc=$?")
(timeout 200 $TR --master-port 29549 bench.py --gpus 8 --impl hvd --model bert_base --param-dtype bf16 --batch 32 --steps 10 --warmup 3 > gpurun_out/r2_bert_n8_b2c.json 2> gpurun_out/r2_bert_n8_b2c.err; echo "bert8 b2c rc=$?"); tail -2 gpurun_out/r2_bert_n8_b2c.err | cut -c1-200
(timeout 200 $TR --master-port 29551 bench.py --gpus 8 --impl hvd --hvd-backend nccl --model bert_base --param-dtype bf16 --batch 32 --steps 10 --warmup 3 > gpurun_out/r2_bert_n8_nccl.json 2> gpurun_out/r2_bert_n8_nccl.err; echo "bert8 nccl rc=$?")
(timeout 240 python tools/tune_4x2.py --trials 4 --steps 1000 --out gpurun_out/r2_tune_4x2.json > gpurun_out/r2_tune_4x2.log 2>&1; echo "tune rc=$?"); tail -2 gpurun_out/r2_tune_4x2.log | cut -c1-300
(timeout 150 $TR --master-port 29553 tools/nvlink_bytes.py --launches 100 --out gpurun_out/r2_nvlink_bytes_w8.json > gpurun_out/r2_nvlink_bytes_w8.log 2>&1; echo "nvlink_bytes rc=$?"); tail -2 gpurun_out/r2_nvlink_bytes_w8.log | cut -c1-300
