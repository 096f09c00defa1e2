#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_models.py -m gpu -x -q -k "depthwise or mobilenet" 2>&1 | grep -E "passed|failed"
run() { env "$@" timeout 300 python bench.py --model $M --steps 30 --warmup 5 --batch-per-gpu $B 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$M b$B $*', d['ms_per_step'], d['value'], d['gpu_launches'])"; }
M=mobilenet_v2
for B in 128 32; do
for t in 1 2 4 8 1000; do run IDCB_DW_SMEM_MIN_ITEMS=$t; done; done
