#!/bin/bash
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
for sz in "24 7 0 1 128" "8 5 0 1" "124 11 3 1 128" "3840 2160 30 2 128" "1000 333 29 1 128"; do echo "== $sz"; timeout 120 ./build/exp/k5 $sz; done > $O/k5_check3.log 2>&1
grep -c MISMATCH $O/k5_check3.log
timeout 300 ./build/exp/k5 3840 2160 0 0 > $O/k5_time3.log 2>&1
grep -E "^----|fused|hist|blur|sobel|copy|read" $O/k5_time3.log | grep -v exact
timeout 100 ./build/exp/k5 3840 2160 30 3 128 > $O/k5_prof_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k5strip|histBytes' -s 10 -c 5 -f -o $O/k5_prof3 ./build/exp/k5 3840 2160 30 3 128 > $O/k5_ncu.log 2>&1
echo ncu rc=$?
