#!/bin/bash
N=${1:-2}
timeout 1500 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "fused_allreduce_matches" 2>&1 | grep -vE "Warning|warn|^$" | tail -8 | cut -c1-400
for model in mobilenet_v2 vgg16; do for comm in fused nccl; do
  echo "== bench $model N=$N comm=$comm"
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --model $model --steps 20 --warmup 5 --comm $comm 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['replicas_identical'], d['allreduce_optim_launches_per_step'])"
done; done
