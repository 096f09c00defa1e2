#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_models.py tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -3
run() { env "$@" timeout 300 python bench.py --model $M --steps 30 --warmup 5 --batch-per-gpu $B 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$M b$B $*', d['ms_per_step'], d['value'], d['gpu_launches'])"; }
M=mobilenet_v2
for B in 128 32; do
for cfg in "2 1" "4 2" "8 4" "16 8"; do set -- $cfg; run IDCB_DW_BLOCKS_PER_SM=$1 IDCB_DW_WGRAD_BLOCKS_PER_SM=$2; done; done
B=128; run IDCB_BN_TAIL=0
M=vgg16; run IDCB_PDL=1
M=densenet121; run IDCB_PDL=1
