#!/bin/bash
export B200CS_DEV_WS=1
timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-extras > gpurun_out/bench_ws.json 2> gpurun_out/bench_ws.err; python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/bench_ws.json").read().strip().splitlines()[-1])
    print("kernel_ms", d["roofline"]["kernel_ms"], "frac", d["roofline"]["frac"], "parity", d["parity_slab"]["max_rel_err"])
except Exception as e:
    print("bench failed", e, open("gpurun_out/bench_ws.err").read()[-1500:])
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pressure_coarsen_ws -s 4 -c 1 -o gpurun_out/$1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-extras > gpurun_out/ncu_$1.log 2>&1; tail -2 gpurun_out/ncu_$1.log
