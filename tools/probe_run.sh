#!/bin/bash
# dev helper: TMA kernel probes under timeouts (a wrong barrier protocol hangs)
mkdir -p gpurun_out
timeout 120 python tools/tma_probe.py --n 48 --nf 4 > gpurun_out/probe1.log 2>&1; echo rc=$? >> gpurun_out/probe1.log; tail -8 gpurun_out/probe1.log
if grep -q "rc=0" gpurun_out/probe1.log; then
  timeout 120 python tools/tma_probe.py --n 40 --nf 3 > gpurun_out/probe2.log 2>&1; echo rc=$? >> gpurun_out/probe2.log; tail -6 gpurun_out/probe2.log
  timeout 120 python tools/tma_probe.py --n 64 --nf 1 --rough --km 9 > gpurun_out/probe3.log 2>&1; echo rc=$? >> gpurun_out/probe3.log; tail -5 gpurun_out/probe3.log
  timeout 200 python tools/tma_probe.py --time > gpurun_out/probe4.log 2>&1; tail -1 gpurun_out/probe4.log
  B200CS_TMA_CTAS=0 timeout 200 python tools/tma_probe.py --time > gpurun_out/probe6.log 2>&1; tail -1 gpurun_out/probe6.log
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_tma.json 2> gpurun_out/bench_tma.err; tail -c 600 gpurun_out/bench_tma.json; tail -3 gpurun_out/bench_tma.err
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:pressure_coarsen_tma -s 4 -c 1 -o gpurun_out/tma_v1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_tma.log 2>&1; tail -3 gpurun_out/ncu_tma.log
fi
