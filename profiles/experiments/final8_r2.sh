#!/bin/bash
# round 2, final state on one 8-GPU box: the whole GPU suite, then the default torchrun bench at 8
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > $O/r2_t_final8.log 2>&1; echo "pytest8 rc=$?"; tail -3 $O/r2_t_final8.log
timeout 900 python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8 --master-port 29808 bench.py --gpus 8 --steps 20 --warmup 5 > $O/r2_final_spmd_8.json 2> $O/r2_final_spmd_8.err; echo "spmd8 rc=$?"
python -c "
import json; d=json.load(open('$O/r2_final_spmd_8.json')); print(d['value'], d['roofline']['frac'], d['e2e']['value'], d['parity'])"
