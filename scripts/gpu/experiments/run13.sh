#!/bin/bash
run() { env "$@" timeout 300 python bench.py --model $M --steps 30 --warmup 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$M $*', d['ms_per_step'], d['value'], d['e2e']['value'], d['gpu_launches'])"; }
for M in densenet121 mobilenet_v2; do run IDCB_CONV_TWO_PER_SM=1; run IDCB_CONV_TWO_PER_SM=0; done
