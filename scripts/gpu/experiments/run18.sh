#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_models.py tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | grep -E "passed|failed"
run() { env "$@" timeout 300 python bench.py --model $M --steps 30 --warmup 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$M $*', d['ms_per_step'], d['value'], d['e2e']['value'], d['gpu_launches'])"; }
for M in densenet121 mobilenet_v2 vgg16; do run IDCB_CONV_TMA_TILED_1X1=1; run IDCB_CONV_TMA_TILED_1X1=0; done
