#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_layerwise.py -m gpu -x -q -s 2>&1 | grep -E "log-loss|passed|failed|Error" | cut -c1-400
python scripts/profile_step.py --steps 1 > gpurun_out/plain_step.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_r2b.csv python scripts/profile_step.py --steps 1 > gpurun_out/ncu_step.log 2>&1
echo "ncu step rc=$?"
