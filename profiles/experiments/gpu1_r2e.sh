#!/bin/bash
# round 2, one GPU: whole GPU suite once more after the last changes, C1 with the host-only per-call cost, histogramStrips group counts
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r2_t_all_final.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2_t_all_final.log
timeout 300 python bench.py --only c1 > $O/r2_c1c.json 2> $O/r2_c1c.err; echo "c1 rc=$?"; python -c "
import json; d=json.load(open('$O/r2_c1c.json')); print(d['host_us_per_call_enqueue_mode'], d['plain'], d['resident_kernel_only']['us_per_call_wall'])"
for g in 148 296 444 592 888 1184; do
  CK_HIST_GROUPS=$g timeout 300 python bench.py --only image > $O/r2_image_hg$g.json 2> $O/r2_image_hg$g.err
  python -c "
import json; d=json.load(open('$O/r2_image_hg$g.json'))['histogramStrips']; print($g, round(d['us_per_frame'],2), round(d['us_per_frame_back_to_back'],2))"
done
