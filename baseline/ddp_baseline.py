#!/usr/bin/env python
"""Comparison baseline of BASELINE.md section 3: PyTorch DDP + NCCL + cuDNN, torchvision model definitions
(weights=None), channels-last, bf16 autocast, binary (1-logit) head, RMSprop, synthetic 50x50x3 patches,
device-timed with CUDA events, max over ranks, same weak-scaling batch (128 / GPU) and the same L2 flush
between timed steps as bench.py. None of this repository's kernels are on this path.

    python baseline/ddp_baseline.py --gpus 1 --steps 20 --warmup 5 [--model densenet121|vgg16|mobilenet_v2]
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 baseline/ddp_baseline.py --gpus N
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--model", default="densenet121")
    ap.add_argument("--batch-per-gpu", type=int, default=128)
    ap.add_argument("--graph", action="store_true", help="capture fwd+bwd+opt in a CUDA graph (single GPU only)")
    a = ap.parse_args()
    import torchvision
    size = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    lrank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lrank)
    dev = torch.device("cuda", lrank)
    if size > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29555")
        dist.init_process_group("nccl", rank=rank, world_size=size, device_id=dev)
    torch.manual_seed(0)
    ctor = {"densenet121": torchvision.models.densenet121, "vgg16": torchvision.models.vgg16,
            "mobilenet_v2": torchvision.models.mobilenet_v2}[a.model]
    net = ctor(weights=None)
    if a.model == "densenet121":
        net.classifier = nn.Linear(1024, 1)
    elif a.model == "vgg16":
        net.avgpool = nn.AdaptiveAvgPool2d(1)
        net.classifier = nn.Linear(512, 1)
    else:
        net.classifier = nn.Sequential(nn.Linear(1280, 1))
    net = net.to(dev).to(memory_format=torch.channels_last)
    model = nn.parallel.DistributedDataParallel(net, device_ids=[lrank]) if size > 1 else net
    opt = torch.optim.RMSprop(model.parameters(), lr=1e-4, alpha=0.9, eps=1e-7)
    lossf = nn.BCEWithLogitsLoss()
    per = a.batch_per_gpu
    x = torch.rand(per, 3, 50, 50, device=dev).to(memory_format=torch.channels_last)
    y = (torch.rand(per, 1, device=dev) > 0.5).float()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def step():
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = model(x)
        loss = lossf(out.float(), y)
        loss.backward()
        opt.step()
        return loss

    for _ in range(max(a.warmup, 3)):
        step()
    torch.cuda.synchronize()
    if size > 1:
        dist.barrier(device_ids=[lrank])
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps)]
    for i in range(a.steps):
        flush.fill_(i & 0xFF)
        starts[i].record()
        step()
        ends[i].record()
    torch.cuda.synchronize()
    ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    if rank == 0:
        gb = per * size
        print(json.dumps({"impl": "pytorch_ddp_cudnn_baseline", "metric": "IDC patches/sec (training step)",
                          "value": round(gb * a.steps / (ms / 1e3), 1), "unit": "patches/s", "n_gpus": size,
                          "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": round(ms / a.steps, 4),
                          "dtype": "bf16 autocast", "data": "synthetic",
                          "config": {"model": a.model, "global_batch": gb, "parallelism": f"ddp{size}",
                                     "torch": torch.__version__, "cudnn": torch.backends.cudnn.version()}}))
    if size > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
