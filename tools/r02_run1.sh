#!/bin/bash
# round-2 GPU pass: tests, bench, launch list, one full ncu capture of the fused kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt
timeout 120 python tools/tma_probe.py --n 48 --nf 4 > gpurun_out/probe1.log 2>&1; echo rc=$? >> gpurun_out/probe1.log; tail -4 gpurun_out/probe1.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo rc=$? >> gpurun_out/pytest_gpu.log; tail -15 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo rc=$?; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-extras > gpurun_out/ncu_launch.log 2>&1; tail -2 gpurun_out/ncu_launch.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pressure_coarsen_tma -s 4 -c 1 -o gpurun_out/r02_tma_g python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-extras > gpurun_out/ncu_tma.log 2>&1; tail -3 gpurun_out/ncu_tma.log
