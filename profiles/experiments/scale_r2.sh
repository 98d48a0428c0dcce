#!/bin/bash
# round 2, one 8-GPU box: multi-GPU tests, 1/2/4/8 scaling (one process per GPU), balancer traces, N-GPU host link
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
nvidia-smi topo -m > $O/r2_topo8.txt 2>&1
timeout 1800 python -m pytest tests -x -q -m gpu > $O/r2_t_all8.log 2>&1; echo "pytest8 rc=$?"; tail -4 $O/r2_t_all8.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 8 4 2; do
  timeout 900 $TR --nproc-per-node $n --master-port $((29600+n)) bench.py --gpus $n --steps 20 --warmup 5 > $O/r2_scale_spmd_$n.json 2> $O/r2_scale_spmd_$n.err; echo "spmd $n rc=$?"
done
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/r2_scale_spmd_1.json 2> $O/r2_scale_spmd_1.err; echo "1 rc=$?"
timeout 600 $TR --nproc-per-node 8 --master-port 29650 bench.py --gpus 8 --steps 25 --warmup 3 --no-all --balancer reference --max-settle 0 > $O/r2_balancer_reference_smooth_8.json 2> $O/r2_bal_ref_s.err; echo "ref smooth rc=$?"
timeout 600 $TR --nproc-per-node 8 --master-port 29651 bench.py --gpus 8 --steps 25 --warmup 3 --no-all --balancer reference --no-smooth --max-settle 0 > $O/r2_balancer_reference_nosmooth_8.json 2> $O/r2_bal_ref_n.err; echo "ref nosmooth rc=$?"
timeout 600 $TR --nproc-per-node 8 --master-port 29652 bench.py --gpus 8 --steps 25 --warmup 3 --no-all --balancer profile --max-settle 0 > $O/r2_balancer_profile_8.json 2> $O/r2_bal_prof.err; echo "profile rc=$?"
timeout 600 python bench.py --only hostlink --gpus 8 > $O/r2_hostlink_inproc_8.json 2> $O/r2_hostlink8.err; echo "hostlink8 rc=$?"
timeout 900 python bench.py --gpus 8 --steps 10 --warmup 5 > $O/r2_scale_inproc_8.json 2> $O/r2_scale_inproc_8.err; echo "inproc8 rc=$?"
