#!/bin/bash
# round-2 final GPU pass, what the driver runs at round end: the GPU tests, smoke(), the default
# bench line and the reference arm.  Outputs under gpurun_out/ (copy what is to be judged to profiles/).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python - <<'PY'
import json
for name in ("bench", "bench_ref"):
    try:
        d = json.loads(open(f"gpurun_out/{name}.json").read().strip().splitlines()[-1])
        r = d.get("roofline", {})
        print(name, "value", d.get("value"), "ms", d.get("ms_per_step"), "frac", r.get("frac"), "kernel_ms", r.get("kernel_ms"),
              "e2e", (d.get("e2e") or {}).get("value"), "parity", (d.get("parity_slab") or {}).get("max_rel_err"))
    except Exception as e:
        print(name, "parse failed", e); print(open(f"gpurun_out/{name}.err").read()[-1500:])
PY
