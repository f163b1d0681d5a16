#!/bin/bash
# One-GPU evidence run: GPU test suite, bench lines, ncu launch lists and full captures (see /opt/skills/guides/B200_PROFILING.md).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/r02_pytest_gpu.log 2>&1; tail -3 gpurun_out/r02_pytest_gpu.log
timeout 300 python bench.py > gpurun_out/r02_bench_ngp.json 2> gpurun_out/r02_bench_ngp.err; tail -c 400 gpurun_out/r02_bench_ngp.json
B="python bench.py --steps 20 --warmup 5 --no-render --no-cpu-baseline"
timeout 200 $B > gpurun_out/plain_ngp.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 700 --csv --log-file gpurun_out/r02_launches_ngp.csv $B > gpurun_out/ncu1.log 2>&1
echo "launch list ngp rc=$?"
timeout 600 ncu --set full --clock-control none -s 300 -c 24 -o gpurun_out/r02_ngp_full $B > gpurun_out/ncu2.log 2>&1
echo "full ngp rc=$?"
N="python bench.py --config nerf --steps 3 --warmup 3 --no-cpu-baseline"
timeout 200 $N > gpurun_out/plain_nerf.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 300 --csv --log-file gpurun_out/r02_launches_nerf.csv $N > gpurun_out/ncu3.log 2>&1
echo "launch list nerf rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dense_ -s 18 -c 6 -o gpurun_out/r02_nerf_full $N > gpurun_out/ncu4.log 2>&1
echo "full nerf rc=$?"
ls -la gpurun_out/*.ncu-rep
