#!/bin/bash
# round 2 profiling pass (1 GPU): launch list of the default bench command, and nBody3Dw after the unit-size change
# (every ncu command runs only after the same command exited 0 without ncu)
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 400 python bench.py --steps 2 --warmup 3 > $O/r2_p_default.json 2> $O/r2_p_default.err && \
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r2_launches.csv python bench.py --steps 2 --warmup 3 > $O/r2_ncu_list.log 2>&1
echo "list rc=$?"
timeout 300 python bench.py --only nbody --steps 3 --warmup 3 > $O/r2_p_nbody.json 2> $O/r2_p_nbody.err && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:nBody3Dw --launch-skip 4 -c 1 -f -o $O/r2_ncu_nbody python bench.py --only nbody --steps 3 --warmup 3 > $O/r2_ncu_nbody.log 2>&1
echo "nbody rc=$?"
ncu -i $O/r2_ncu_nbody.ncu-rep --page raw --csv > $O/r2_ncu_nbody_raw.csv 2>/dev/null
