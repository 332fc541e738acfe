#!/bin/bash
# dev helper: tools/nfield_time.py for each library variant in gpurun_variants/*.so
cp fv3net_b200/libb200cs.so /tmp/orig.so
for v in gpurun_variants/*.so; do
  cp $v fv3net_b200/libb200cs.so; echo "== $v"
  timeout 300 python tools/nfield_time.py 2>&1 | tail -6
done
cp /tmp/orig.so fv3net_b200/libb200cs.so
