#!/bin/bash
# round-2 first GPU pass: full GPU test suite, then A/B of programmatic dependent launch and the BN tail
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
run() { echo "== $*"; env "$@" python bench.py --model $M --steps 20 --warmup 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['gpu_launches'])"; }
M=densenet121
run IDCB_PDL=1 IDCB_BN_TAIL=1
run IDCB_PDL=0 IDCB_BN_TAIL=1
run IDCB_PDL=1 IDCB_BN_TAIL=0
run IDCB_PDL=0 IDCB_BN_TAIL=0
for M in vgg16 mobilenet_v2; do run IDCB_PDL=1; run IDCB_PDL=0; done
echo "== conv phases (PDL on)"; python scripts/conv_phases.py 2>&1 | tail -40
echo "== conv phases (PDL off)"; IDCB_PDL=0 python scripts/conv_phases.py 2>&1 | tail -40 | grep -E "us/launch|trivial" | cut -c1-60
