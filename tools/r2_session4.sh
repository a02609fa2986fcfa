#!/bin/bash
# Round-2 4-GPU session: bench line with parity at N=4, MLP config b2c vs stock DDP, W=4 fixed-cost probe.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
(timeout 300 $TR --master-port 29561 bench.py --gpus 4 --steps 10 --warmup 5 > gpurun_out/r2_bench_n4_b2c.json 2> gpurun_out/r2_bench_n4_b2c.err; echo "bench4 rc=$?"); tail -2 gpurun_out/r2_bench_n4_b2c.err | cut -c1-200
(timeout 120 $TR --master-port 29563 bench.py --gpus 4 --model mnist_mlp --batch 32 --steps 500 --warmup 50 --no-bucket-timing --no-parity --no-allreduce-section > gpurun_out/r2_mlp_n4_b2c.json 2> gpurun_out/r2_mlp_n4_b2c.err; echo "mlp4 b2c rc=$?")
(timeout 120 $TR --master-port 29565 bench.py --gpus 4 --impl nccl --model mnist_mlp --batch 32 --steps 500 --warmup 50 > gpurun_out/r2_mlp_n4_nccl.json 2> gpurun_out/r2_mlp_n4_nccl.err; echo "mlp4 nccl rc=$?")
(timeout 200 $TR --master-port 29567 tools/r2_probe.py --iters 10 --sizes 4096,111528,1625000,12415000,31500000 --bcast 212992,102441040 --out gpurun_out/r2_probe_w4.json > gpurun_out/r2_probe_w4.log 2>&1; echo "probe4 rc=$?"); tail -2 gpurun_out/r2_probe_w4.log | cut -c1-200
