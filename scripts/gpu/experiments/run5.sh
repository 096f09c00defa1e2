#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_models.py tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -4
echo "== conv phases"; python scripts/conv_phases.py 2>&1 | grep -E "us/launch|trivial" | cut -c1-220
run() { echo "== $M $*"; env "$@" python bench.py --model $M --steps 20 --warmup 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['gpu_launches'])"; }
for M in densenet121 vgg16 mobilenet_v2; do run IDCB_PDL=1; done
