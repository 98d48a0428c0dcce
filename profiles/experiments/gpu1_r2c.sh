#!/bin/bash
# round 2, one GPU: full GPU tests (chained pipeline, strip kernels, raw ABI ...), smoke, remaining ncu captures, default bench
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > $O/r2_t_all.log 2>&1; echo "pytest rc=$?"; tail -6 $O/r2_t_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/r2_smoke.log
bash profiles/experiments/prof_r2b.sh
timeout 600 python bench.py > $O/r2_bench_default.json 2> $O/r2_bench_default.err; echo "bench rc=$?"; tail -3 $O/r2_bench_default.err
