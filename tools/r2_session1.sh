#!/bin/bash
# Round-2 1-GPU session: bench line at N=1, ncu launch list of the same command, ncu --set full of the fused step.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-allreduce-section"
(timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench1 rc=$?")
$CMD > gpurun_out/r2_plain_n1.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file gpurun_out/r2_launches_n1.csv $CMD > gpurun_out/r2_ncu_launches.log 2>&1; echo "launch list rc=$?"
python tools/ncu_target.py fused > gpurun_out/r2_target_fused.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_fused_step_local -s 1 -c 2 -o gpurun_out/r2_prof_fused python tools/ncu_target.py fused > gpurun_out/r2_ncu_fused.log 2>&1; echo "ncu fused rc=$?"
python tools/ncu_target.py fusedzc > gpurun_out/r2_target_fusedzc.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_fused_step_zc -s 1 -c 1 -o gpurun_out/r2_prof_fusedzc python tools/ncu_target.py fusedzc > gpurun_out/r2_ncu_fusedzc.log 2>&1; echo "ncu fusedzc rc=$?"
ls -la gpurun_out/*.ncu-rep 2>/dev/null | tail -3
