#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_models.py -m gpu -x -q 2>&1 | grep -E "passed|failed"
run() { env "$@" timeout 300 python bench.py --model $M --steps 30 --warmup 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$M $*', d['ms_per_step'], d['value'], d['e2e']['value'], d['gpu_launches'])"; }
M=densenet121; run IDCB_PDL=1; run IDCB_PDL=1
