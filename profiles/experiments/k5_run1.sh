#!/bin/bash
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/k5_smi.txt
for sz in "24 7 0 1" "24 7 2 1" "8 5 0 1" "124 11 3 1" "4096 100 30 1" "3840 2160 30 2" "1000 333 29 1"; do echo "== $sz"; timeout 120 ./build/exp/k5 $sz; done > $O/k5_check.log 2>&1
timeout 300 ./build/exp/k5 > $O/k5_time.log 2>&1
echo rc=$?
tail -5 $O/k5_check.log; grep -c MISMATCH $O/k5_check.log $O/k5_time.log
