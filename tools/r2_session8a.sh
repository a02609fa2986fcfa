#!/bin/bash
# Round-2 8-GPU session A: cross-GPU parity at W=8, fixed-cost probe, one short bench line.  Every part has its
# own timeout and log; nothing here runs under a profiler.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2_gpus8.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
(timeout 600 python -m pytest tests -m gpu -q -rs > gpurun_out/r2_pytest_gpu_n8.log 2>&1; echo "pytest n8 rc=$?")
tail -4 gpurun_out/r2_pytest_gpu_n8.log
(timeout 300 $TR --master-port 29521 tools/r2_probe.py --iters 10 --sizes 4096,1625000,12415000,27021000,31500000 \
   --bcast 212992,102441040 --out gpurun_out/r2_probe_w8.json > gpurun_out/r2_probe_w8.log 2>&1; echo "probe rc=$?")
tail -4 gpurun_out/r2_probe_w8.log
(timeout 400 $TR --master-port 29523 bench.py --gpus 8 --steps 6 --warmup 4 > gpurun_out/r2_bench_n8_dev.json 2> gpurun_out/r2_bench_n8_dev.err; echo "bench8 rc=$?")
tail -3 gpurun_out/r2_bench_n8_dev.err
