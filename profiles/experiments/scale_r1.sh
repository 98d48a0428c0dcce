#!/bin/bash
# round-1 scaling pass on an 8-GPU box: full GPU tests, then bench at N = 1, 2, 4, 8 (one process per GPU), then the
# in-process 8-device variant (one process driving all devices: NCCL broadcast path of whole-array reads)
set -u
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 600 python -m pytest tests/test_kernels_parity.py tests/test_modes.py tests/test_pipeline_d2d.py tests/test_device_pool.py tests/test_next_rows.py tests/test_zz_no_alloc.py -x -q -m gpu > $O/t_all8.log 2>&1; echo "pytest rc=$?"; tail -3 $O/t_all8.log
timeout 400 python bench.py --gpus 1 --steps 5 --warmup 3 > $O/scale_1.json 2> $O/scale_1.err; echo "N=1 rc=$?"
port=29600
for n in 2 4 8; do
  port=$((port+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps 5 --warmup 3 > $O/scale_$n.json 2> $O/scale_$n.err; echo "N=$n rc=$?"
done
timeout 600 python bench.py --gpus 8 --steps 5 --warmup 3 > $O/bench_inproc8.json 2> $O/bench_inproc8.err; echo "inproc8 rc=$?"
python - <<'PY'
import json
for n in (1, 2, 4, 8, "inproc8"):
    f = f"gpurun_out/scale_{n}.json" if n != "inproc8" else "gpurun_out/bench_inproc8.json"
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        nb = d.get("workloads", {}).get("nbody_c4", {})
        print(n, "mandel", round(d["value"]), "e2e", round(d["e2e"]["value"]), "frac", round(d["roofline"]["frac"], 3),
              "conv", d["balancer"]["iterations_to_converge"], "shares", d["balancer"]["final_shares"],
              "| nbody kernel", round(nb.get("GInter_per_s_kernel", 0)), "dev", round(nb.get("GInter_per_s", 0)), "e2e", round(nb.get("GInter_per_s_e2e", 0)),
              "frac", round(nb.get("roofline", {}).get("frac", 0), 3))
    except Exception as e:
        print(n, "failed", e)
PY
