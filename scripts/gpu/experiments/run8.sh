#!/bin/bash
run() { echo "== $M $*"; env "$@" timeout 300 python bench.py --model $M --steps 20 --warmup 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['gpu_launches'])"; }
M=densenet121
for k in 2 3 4 5 6; do run IDCB_CONV_KSPLIT_MINKB=$k; done
M=mobilenet_v2
for k in 2 4 6; do run IDCB_CONV_KSPLIT_MINKB=$k; done
M=vgg16
for k in 2 4; do run IDCB_CONV_KSPLIT_MINKB=$k; done
