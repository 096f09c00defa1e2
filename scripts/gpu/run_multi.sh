#!/bin/bash
# multi-GPU correctness + scaling check (N = number of GPUs of the lease)
N=${1:-2}
timeout 1500 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -s 2>&1 | grep -vE "Warning|warn|^$" | tail -25 | cut -c1-600 | tee gpurun_out/multi_gpu_tests_$N.log
for comm in fused nccl; do
  echo "== bench N=$N comm=$comm"
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --comm $comm 2>&1 | tail -1 | cut -c1-1800 | tee gpurun_out/bench_${N}_$comm.json
done
echo "== fused all-reduce kernel"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 scripts/bench_allreduce.py 2>&1 | grep '^{' | tee gpurun_out/allreduce_$N.jsonl
