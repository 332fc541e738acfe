#!/bin/bash
# dev helper: bring-up of the warp-specialised kernel under timeouts (a wrong barrier protocol hangs)
mkdir -p gpurun_out
export B200CS_DEV_WS=1
timeout 60 python tools/tma_probe.py --n 16 --nf 4 > gpurun_out/ws1.log 2>&1; echo "rc=$?" >> gpurun_out/ws1.log; tail -7 gpurun_out/ws1.log
if grep -q "rc=0" gpurun_out/ws1.log; then
  timeout 60 python tools/tma_probe.py --n 48 --nf 4 > gpurun_out/ws2.log 2>&1; echo "rc=$?" >> gpurun_out/ws2.log; tail -6 gpurun_out/ws2.log
  timeout 60 python tools/tma_probe.py --n 40 --nf 3 > gpurun_out/ws3.log 2>&1; echo "rc=$?" >> gpurun_out/ws3.log; tail -5 gpurun_out/ws3.log
  timeout 60 python tools/tma_probe.py --n 64 --nf 1 --rough --km 9 > gpurun_out/ws4.log 2>&1; echo "rc=$?" >> gpurun_out/ws4.log; tail -4 gpurun_out/ws4.log
  timeout 120 python tools/tma_probe.py --time > gpurun_out/ws5.log 2>&1; tail -2 gpurun_out/ws5.log
fi
