#!/bin/bash
# round-1 profiling pass (1 GPU): every ncu command runs only after the same command exited 0 without ncu
set -u
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
run() { echo "== $*" >> $O/prof.log; "$@" >> $O/prof.log 2>&1; echo "rc=$?" >> $O/prof.log; }
: > $O/prof.log
for b in 16 32 64; do
  timeout 300 python bench.py --no-all --steps 5 --warmup 3 --e2e-blobs $b > $O/e2e_blobs_$b.json 2>> $O/prof.log; echo "blobs $b rc=$?" >> $O/prof.log
done
timeout 300 python bench.py --only nbody --steps 3 --warmup 3 > $O/p_nbody.json 2>> $O/prof.log && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:nBody3Dw --launch-skip 4 -c 1 -f -o $O/prof_nbody python bench.py --only nbody --steps 3 --warmup 3 > $O/ncu_nbody.log 2>&1
timeout 300 python bench.py --only saxpy --steps 3 --warmup 3 > $O/p_saxpy.json 2>> $O/prof.log && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:saxpy4 -c 24 -f -o $O/prof_saxpy python bench.py --only saxpy --steps 3 --warmup 3 > $O/ncu_saxpy.log 2>&1
timeout 300 python bench.py --steps 2 --warmup 3 > $O/p_default.json 2>> $O/prof.log && \
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 > $O/ncu_list.log 2>&1
for r in nbody saxpy; do
  ncu -i $O/prof_$r.ncu-rep --page raw --csv > $O/ncu_${r}_raw.csv 2>> $O/prof.log
done
ls -la $O >> $O/prof.log
