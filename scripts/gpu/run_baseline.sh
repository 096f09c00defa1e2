#!/bin/bash
# library baseline arms (PyTorch + cuDNN, none of our kernels): eager, whole-step CUDA graph, torch.compile
for m in densenet121 vgg16 mobilenet_v2; do
  for mode in eager graph; do
    timeout 600 python baseline/ddp_baseline.py --gpus 1 --steps 20 --warmup 5 --model $m --mode $mode 2>&1 | grep '^{' | tee -a gpurun_out/baseline_arms.jsonl
  done
done
timeout 900 python baseline/ddp_baseline.py --gpus 1 --steps 20 --warmup 5 --model densenet121 --mode compile 2>&1 | grep -E '^\{|Error|error' | tail -3 | tee -a gpurun_out/baseline_arms.jsonl
