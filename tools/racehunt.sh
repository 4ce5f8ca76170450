#!/bin/bash
# Development aid: run threaded decodes repeatedly and report any pass that differs from the
# golden MD5s (E2E_FAST=2 keeps every frame and describes where the first bad frames differ).
n=0
for mode in "" "EAV1_SYNC_EACH=1"; do
for i in 1 2 3 4 5 6; do
  env E2E_FAST=2 $mode EASYAV1_THREADS=16 timeout 200 python tools/e2e_repeat.py c2_1080p_bench 4 > gpurun_out/hunt_$n.out 2> gpurun_out/hunt_$n.err
  if grep -q "bad \[[0-9]" gpurun_out/hunt_$n.out; then echo "FOUND in run $n ($mode)"; cat gpurun_out/hunt_$n.out; tail -3 gpurun_out/hunt_$n.err; exit 1; fi
  n=$((n+1))
done
done
echo "no bad pass in $n runs"
