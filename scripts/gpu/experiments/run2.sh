#!/bin/bash
# ncu: source-level profile of the small dense-layer kernels + launch list of one DenseNet-121 step
set -o pipefail
python scripts/ncu_small.py > gpurun_out/plain_small.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_ -s 6 -c 6 -f -o gpurun_out/prof_small python scripts/ncu_small.py > gpurun_out/ncu_small.log 2>&1
echo "ncu small rc=$?"; tail -3 gpurun_out/ncu_small.log
python scripts/profile_step.py --steps 1 > gpurun_out/plain_step.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_r2.csv python scripts/profile_step.py --steps 1 > gpurun_out/ncu_step.log 2>&1
echo "ncu step rc=$?"; tail -2 gpurun_out/ncu_step.log
