#!/bin/bash
# TMA im2col activation operand: numerics, per-kernel timings with/without, end-to-end effect
timeout 600 python -m pytest tests/test_gpu_conv.py tests/test_gpu_models.py -m gpu -x -q 2>&1 | tail -8
echo "== conv phases TMA=1"; python scripts/conv_phases.py 2>&1 | grep -E "us/launch|trivial" | cut -c1-200
echo "== conv phases TMA=0"; IDCB_CONV_TMA=0 python scripts/conv_phases.py 2>&1 | grep -E "us/launch|trivial" | cut -c1-64
run() { echo "== $M $*"; env "$@" python bench.py --model $M --steps 20 --warmup 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['gpu_launches'])"; }
for M in densenet121 vgg16 mobilenet_v2; do run IDCB_CONV_TMA=1; run IDCB_CONV_TMA=0; done
