#!/bin/bash
# dev helper: TMA kernel probes under timeouts (a wrong barrier protocol hangs)
mkdir -p gpurun_out
timeout 120 python tools/tma_probe.py --n 48 --nf 4 > gpurun_out/probe1.log 2>&1; echo rc=$? >> gpurun_out/probe1.log; tail -8 gpurun_out/probe1.log
if grep -q "rc=0" gpurun_out/probe1.log; then
  timeout 120 python tools/tma_probe.py --n 40 --nf 3 > gpurun_out/probe2.log 2>&1; echo rc=$? >> gpurun_out/probe2.log; tail -6 gpurun_out/probe2.log
  timeout 120 python tools/tma_probe.py --n 64 --nf 1 --rough > gpurun_out/probe3.log 2>&1; echo rc=$? >> gpurun_out/probe3.log; tail -5 gpurun_out/probe3.log
  timeout 200 python tools/tma_probe.py --time > gpurun_out/probe4.log 2>&1; tail -2 gpurun_out/probe4.log
  B200CS_TMA_CTAS=2 timeout 200 python tools/tma_probe.py --time > gpurun_out/probe5.log 2>&1; tail -2 gpurun_out/probe5.log
  B200CS_TMA_CTAS=0 timeout 200 python tools/tma_probe.py --time > gpurun_out/probe6.log 2>&1; tail -2 gpurun_out/probe6.log
fi
