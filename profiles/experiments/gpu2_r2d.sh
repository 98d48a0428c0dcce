#!/bin/bash
# round 2, two GPUs: the tests that follow the raw-ABI file in collection order (the 8-GPU run stopped there), incl. NCCL/IPC and the second-device no-alloc test
cd "$GRAFT_REPO_ROOT"
timeout 900 python -m pytest tests/test_raw_abi_replay.py tests/test_single_gpu_pipeline.py tests/test_spmd_nccl.py tests/test_zz_no_alloc.py tests/test_pipeline_d2d.py tests/test_device_pool.py -x -q -m gpu > gpurun_out/r2_t_tail2.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_t_tail2.log
