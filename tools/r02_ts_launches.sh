#!/bin/bash
# dev helper: per-kernel durations of one whole C384 timestep (FAST) under ncu
mkdir -p gpurun_out
timeout 300 python tools/restart_bw.py --modes fast --reps 5 > gpurun_out/restart_bw_fast.json 2>&1; tail -c 400 gpurun_out/restart_bw_fast.json; echo
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/ts_launches.csv python tools/restart_bw.py --modes fast --reps 1 --warm 1 > gpurun_out/ts_ncu.log 2>&1; tail -2 gpurun_out/ts_ncu.log
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/ts_launches.csv")) if len(r) > 5]
hdr = rows[0]; ik = hdr.index("Kernel Name"); iv = hdr.index("Metric Value"); iu = hdr.index("Metric Unit")
ks = [(r[ik], float(r[iv].replace(",", "")) * (1e-3 if r[iu] in ("us", "usecond") else (1e-6 if r[iu] in ("ns", "nsecond") else 1.0))) for r in rows[1:]
      if "b200cs::" in r[ik]]   # torch's kernels are the synthetic data being generated
half = ks[len(ks) // 2:]   # second timestep (after the warm-up one)
tot = sum(t for _, t in half)
agg = collections.OrderedDict()
for k, t in half:
    k = k[:70]
    agg.setdefault(k, [0, 0.0]); agg[k][0] += 1; agg[k][1] += t
print("launches in one timestep:", len(half), "sum of kernel time %.3f ms" % tot)
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%3d x %-70s %8.3f ms" % (n, k, t))
PY
