#!/bin/bash
# round 2: ncu --set full of the three strip kernels not yet captured (each ncu run after the same command exited 0 without ncu)
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 300 python bench.py --only image > $O/r2_image3.json 2> $O/r2_image3.err || exit 1
for k in sobelStrips histogramStrips blurSobelHistogram; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^$k\$" --launch-skip 8 -c 2 -f -o $O/r2_ncu_$k python bench.py --only image > $O/r2_ncu_$k.log 2>&1; echo "ncu $k rc=$?"
  ncu -i $O/r2_ncu_$k.ncu-rep --page raw --csv > $O/r2_ncu_${k}_raw.csv 2>/dev/null
done
timeout 300 python bench.py --only c1 > $O/r2_c1b.json 2> $O/r2_c1b.err; echo "c1 rc=$?"
