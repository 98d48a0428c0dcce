#!/bin/bash
# round-1 profiling pass 2 (1 GPU): tests, resident SAXPY capture, full launch list.  ncu only after the plain run exited 0.
set -u
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
: > $O/prof.log
timeout 900 python -m pytest tests -x -q -m gpu > $O/t_all.log 2>&1; echo "pytest rc=$?"; tail -3 $O/t_all.log
timeout 300 python bench.py --only saxpy --steps 3 --warmup 3 > $O/p_saxpy.json 2>> $O/prof.log && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:saxpy4 --launch-skip 97 -c 1 -f -o $O/prof_saxpy python bench.py --only saxpy --steps 3 --warmup 3 > $O/ncu_saxpy.log 2>&1
echo "saxpy ncu rc=$?"
timeout 300 python bench.py --steps 2 --warmup 3 > $O/p_default.json 2>> $O/prof.log && \
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 > $O/ncu_list.log 2>&1
echo "launch list rc=$?"
ncu -i $O/prof_saxpy.ncu-rep --page raw --csv > $O/ncu_saxpy_raw.csv 2>> $O/prof.log
python profiles/aggregate_launches.py $O/launches.csv > $O/launches_by_kernel.csv; cat $O/launches_by_kernel.csv
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/p_reference.json 2>> $O/prof.log; echo "reference rc=$?"; head -c 600 $O/p_reference.json
