#!/bin/bash
for m in mobilenet_v2 vgg16 densenet121; do
python scripts/profile_step.py --model $m --steps 1 > gpurun_out/plain_step_$m.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$m.csv python scripts/profile_step.py --model $m --steps 1 > gpurun_out/ncu_step_$m.log 2>&1
echo "ncu $m rc=$?"
done
