#!/bin/bash
# 8-GPU box: topology, the no-alloc test after the multi-GPU tests, SPMD N=8 with and without NUMA-local host memory
set -u
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
nvidia-smi topo -m > $O/topo.txt 2>&1
(lscpu | grep -i -E "numa|socket|model name|^CPU\(s\)"; cat /sys/devices/system/node/node*/cpulist; for d in /sys/bus/pci/devices/*; do if [ "$(cat $d/vendor)" = "0x10de" ] && [[ "$(cat $d/class)" == 0x0302* ]]; then echo "$d numa $(cat $d/numa_node)"; fi; done) >> $O/topo.txt 2>&1
timeout 600 python -m pytest tests/test_kernels_parity.py tests/test_modes.py tests/test_device_pool.py tests/test_zz_no_alloc.py -x -q -m gpu > $O/t_all8.log 2>&1; echo "pytest rc=$?"; tail -3 $O/t_all8.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 8 --steps 5 --warmup 3 > $O/scale_8.json 2> $O/scale_8.err; echo "N=8 numa rc=$?"
CK_NUMA=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29702 bench.py --gpus 8 --steps 5 --warmup 3 --no-all > $O/scale_8_nonuma.json 2> $O/scale_8_nonuma.err; echo "N=8 default policy rc=$?"
python - <<'PY'
import json
for f in ("scale_8", "scale_8_nonuma"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        e = d["e2e"]
        print(f, "mandel", round(d["value"]), "e2e", round(e["value"]), "ms", round(e["ms_per_step"], 2), "settle", e["balancer_settle_steps"],
              "shares", e["balancer"]["final_shares"], d["config"].get("host_memory_policy_rank0"))
        nb = d.get("workloads", {}).get("nbody_c4")
        if nb:
            print("   nbody kernel", round(nb["GInter_per_s_kernel"]), "dev", round(nb["GInter_per_s"]), "e2e", round(nb["GInter_per_s_e2e"]))
    except Exception as ex:
        print(f, "failed", ex)
PY
head -30 $O/topo.txt
