#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_models.py -m gpu -x -q 2>&1 | tail -2
echo "== conv phases"; timeout 300 python scripts/conv_phases.py 2>&1 | grep -E "^fprop" | cut -c1-70
run() { echo "== $M $*"; env "$@" timeout 300 python bench.py --model $M --steps 20 --warmup 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['gpu_launches'])"; }
for M in densenet121 vgg16 mobilenet_v2; do run IDCB_PDL=1; done
