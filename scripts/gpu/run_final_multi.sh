#!/bin/bash
# final multi-GPU measurements of the round (N = GPUs of the lease)
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
out=gpurun_out/final_multi_$N.jsonl; : > $out
for model in densenet121 vgg16 mobilenet_v2; do for comm in fused nccl; do
  timeout 600 $TR --master-port 29512 bench.py --gpus $N --model $model --steps 20 --warmup 5 --comm $comm 2>&1 | tail -1 | cut -c1-2000 >> $out
done; done
python - <<PY
import json
for l in open("$out"):
    try: d=json.loads(l)
    except Exception: print("BAD", l[:200]); continue
    print(d["config"]["model"], d["config"]["comm"], d["n_gpus"], d["ms_per_step"], d["value"], d["e2e"]["value"], d["replicas_identical"], d["allreduce_optim_launches_per_step"], d["vs_baseline"])
PY
if [ "$N" = "8" ]; then
echo "== bench_fed"; timeout 900 $TR --master-port 29514 scripts/bench_fed.py --mode all 2>&1 | grep '^{' | tee gpurun_out/bench_fed_$N.jsonl | cut -c1-420
timeout 600 $TR --master-port 29515 scripts/bench_fed.py --mode secure_size 2>&1 | grep -E '^\{' | tee -a gpurun_out/bench_fed_$N.jsonl | cut -c1-600
echo "== library FedAvg baseline"; for m in graph eager; do timeout 600 $TR --master-port 29516 baseline/fed_baseline.py --mode $m 2>&1 | grep -E '^\{' | tee -a gpurun_out/fed_baseline_$N.jsonl; done
echo "== DDP baseline, graph arm"; timeout 600 $TR --master-port 29517 baseline/ddp_baseline.py --gpus $N --steps 20 --warmup 5 --mode graph 2>&1 | grep -E '^\{|Error' | tail -2 | tee gpurun_out/ddp_graph_$N.jsonl
fi
