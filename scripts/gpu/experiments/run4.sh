#!/bin/bash
# ncu source-level profiles with a fine sampling interval: block-1 shapes (h=13) and block-3 shapes (h=3)
set -o pipefail
for h in 13 3; do
python scripts/ncu_small.py $h > gpurun_out/plain_small_$h.log 2>&1 &&
ncu --set full --clock-control none --import-source on --warp-sampling-interval 1 -k regex:conv_ -s 6 -c 6 -f -o gpurun_out/prof_h$h python scripts/ncu_small.py $h > gpurun_out/ncu_h$h.log 2>&1
echo "ncu h=$h rc=$?"; tail -2 gpurun_out/ncu_h$h.log
done
