#!/bin/bash
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514"
echo "== bench_fed all"; timeout 900 $TR scripts/bench_fed.py --mode all 2>&1 | grep '^{' | tee gpurun_out/bench_fed_$N.jsonl
echo "== secure masked sum at MobileNetV2 size"; timeout 600 $TR scripts/bench_fed.py --mode secure_size 2>&1 | grep -E '^\{|Error' | tee -a gpurun_out/bench_fed_$N.jsonl
echo "== fed_model.py, BASELINE config 4: MobileNetV2, 10 clients (8 train / 2 test), 50 rounds x 1 local epoch"
mkdir -p /tmp/fedrun
timeout 1500 $TR fed_model.py /tmp/fedrun 50 iid --synthetic --arch mobilenet_v2 --clients 10 --pretrain-epochs 1 --steps-per-epoch 30 --graphs 2>&1 | grep -vE "Warning|warn" | tail -70 | tee gpurun_out/fed_model_50rounds_$N.log
