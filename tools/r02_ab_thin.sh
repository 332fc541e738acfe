#!/bin/bash
# dev helper: tools/kernel_bw.py (thin kernels) for each library variant in gpurun_variants/*.so
cp fv3net_b200/libb200cs.so /tmp/orig.so
for v in gpurun_variants/*.so; do
  cp $v fv3net_b200/libb200cs.so; echo "== $v"
  timeout 300 python tools/kernel_bw.py 2>&1 | grep "^K1\|^K3\|^K5"
done
cp /tmp/orig.so fv3net_b200/libb200cs.so
