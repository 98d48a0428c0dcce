#!/bin/bash
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r2_t_all_final.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2_t_all_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/r2_smoke.log
timeout 600 python bench.py > $O/r2_bench_default.json 2> $O/r2_bench_default.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('$O/r2_bench_default.json')); print(d['value'], d['roofline']['frac'], d['e2e']['value'])
for k,v in d['workloads']['pipeline_c5']['kernels_resident'].items():
    if isinstance(v,dict) and 'us_per_frame' in v: print(k, round(v['us_per_frame'],2), round(v['us_per_frame_back_to_back'],2), round(v['roofline']['frac'],3), v['roofline']['traffic'])
print(d['workloads']['pipeline_c5']['kernels_resident']['parity'])"
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > $O/r2_bench_reference.json 2> $O/r2_bench_reference.err; echo "ref rc=$?"; head -c 400 $O/r2_bench_reference.json
