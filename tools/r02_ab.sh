#!/bin/bash
# dev helper: time the fused kernel for the .so variants found in gpurun_out/variants/*.so (each copied over the
# product library in turn); prints kernel ms per variant
mkdir -p gpurun_out
cp fv3net_b200/libb200cs.so /tmp/orig.so
for v in gpurun_variants/*.so; do
  cp $v fv3net_b200/libb200cs.so
  timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-extras --steps 30 > /tmp/b.json 2> /tmp/b.err
  python - "$v" <<'PY'
import json, sys
try:
    d = json.loads(open("/tmp/b.json").read().strip().splitlines()[-1])
    print(sys.argv[1], "kernel_ms", round(d["roofline"]["kernel_ms"], 4), "parity", d["parity_slab"]["max_rel_err"])
except Exception as e:
    print(sys.argv[1], "failed", e, open("/tmp/b.err").read()[-800:])
PY
done
cp /tmp/orig.so fv3net_b200/libb200cs.so
