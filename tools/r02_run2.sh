#!/bin/bash
# round-2 GPU pass: tests, short bench, optional ncu of the fused kernel (arg 1 = report name)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline --no-e2e --no-extras > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; echo "bench rc=$?"; python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/bench_short.json").read().strip().splitlines()[-1])
    print("ms_per_step", d["ms_per_step"], "frac", d["roofline"]["frac"], "kernel_ms", d["roofline"].get("kernel_ms"), "parity", d.get("parity_slab", {}).get("max_rel_err"), "other", d.get("other_terrain"))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/bench_short.err").read()[-1500:])
PY
if [ -n "$1" ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:pressure_coarsen_tma -s 4 -c 1 -o gpurun_out/$1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-extras > gpurun_out/ncu_$1.log 2>&1; tail -2 gpurun_out/ncu_$1.log
fi
