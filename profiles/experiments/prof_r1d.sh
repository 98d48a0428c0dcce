#!/bin/bash
# final 1-GPU pass of round 1: full GPU tests, smoke, default bench, resident SAXPY capture (after the plain run)
set -u
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > $O/t_all.log 2>&1; echo "pytest rc=$?"; tail -2 $O/t_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
timeout 400 python bench.py > $O/p_default.json 2> $O/p_default.err; echo "bench rc=$?"
timeout 300 python bench.py --only saxpy --steps 3 --warmup 3 > $O/p_saxpy.json 2>> $O/prof.log && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:saxpy4 --launch-skip 193 -c 1 -f -o $O/prof_saxpy python bench.py --only saxpy --steps 3 --warmup 3 > $O/ncu_saxpy.log 2>&1
echo "saxpy ncu rc=$?"
ncu -i $O/prof_saxpy.ncu-rep --page raw --csv > $O/ncu_saxpy_raw.csv 2>> $O/prof.log
python profiles/make_summary.py saxpy=$O/ncu_saxpy_raw.csv | python -c "import json,sys; d=json.load(sys.stdin)['saxpy']; print(d['grid'], d['time_s'], d['traffic_B'], d['dram_GBps'], d['dram_throughput_pct'])"
