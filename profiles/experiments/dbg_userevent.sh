#!/bin/bash
cd "$GRAFT_REPO_ROOT"
timeout 60 python profiles/experiments/dbg_ue.py > gpurun_out/dbg_ue.log 2>&1; echo rc=$?; tail -40 gpurun_out/dbg_ue.log
