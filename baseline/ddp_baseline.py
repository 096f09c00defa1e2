#!/usr/bin/env python
"""Comparison baseline of BASELINE.md section 3: PyTorch DDP + NCCL + cuDNN, torchvision model definitions
(weights=None), channels-last, bf16 autocast, binary (1-logit) head, RMSprop, synthetic 50x50x3 patches,
device-timed with CUDA events, max over ranks, same weak-scaling batch (128 / GPU) and the same L2 flush
between timed steps as bench.py. None of this repository's kernels are on this path.

Three arms, so that the library stack is given its best shot (an eager step of a 121-layer model is bound by
host launch cost, which says nothing about cuDNN):

* ``--mode eager``   : plain PyTorch, one Python-issued launch per op;
* ``--mode graph``   : the whole step (zero_grad, forward, backward, optimizer with ``capturable=True``; under
                       DDP also its NCCL all-reduces) captured once in a CUDA graph and replayed;
* ``--mode compile`` : ``torch.compile`` of the model (``--compile-mode reduce-overhead`` = Inductor + CUDA graphs,
                       or ``max-autotune``), eager optimizer with ``foreach`` kernels.

    python baseline/ddp_baseline.py --gpus 1 --steps 20 --warmup 5 [--model densenet121|vgg16|mobilenet_v2] [--mode graph]
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
    ap.add_argument("--mode", default="eager", choices=["eager", "graph", "compile"])
    ap.add_argument("--graph", action="store_true", help="same as --mode graph")
    ap.add_argument("--compile-mode", default="reduce-overhead")
    a = ap.parse_args()
    if a.graph:
        a.mode = "graph"
    import torchvision
    size = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    lrank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lrank)
    dev = torch.device("cuda", lrank)
    if size > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29555")
        if a.mode == "graph":
            os.environ["TORCH_NCCL_ASYNC_ERROR_HANDLING"] = "0"     # the watchdog must not poll inside a capture
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
    side = torch.cuda.Stream(device=dev)
    if size > 1:
        # DDP built (and warmed up) on the stream that will later be captured
        with torch.cuda.stream(side):
            model = nn.parallel.DistributedDataParallel(net, device_ids=[lrank])
    else:
        model = net
    opt = torch.optim.RMSprop(model.parameters(), lr=1e-4, alpha=0.9, eps=1e-7, capturable=(a.mode == "graph"))
    lossf = nn.BCEWithLogitsLoss()
    per = a.batch_per_gpu
    x = torch.rand(per, 3, 50, 50, device=dev).to(memory_format=torch.channels_last)
    y = (torch.rand(per, 1, device=dev) > 0.5).float()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    fwd = torch.compile(model, mode=a.compile_mode) if a.mode == "compile" else model

    def eager_step():
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = fwd(x)
        loss = lossf(out.float(), y)
        loss.backward()
        opt.step()
        return loss

    step = eager_step
    if a.mode == "graph":
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(12):                 # DDP needs >= 11 eager iterations before it can be captured
                eager_step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        opt.zero_grad(set_to_none=True)
        with torch.cuda.graph(graph, stream=side):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = model(x)
            static_loss = lossf(out.float(), y)
            static_loss.backward()
            opt.step()                          # .grad was None at capture: every replay overwrites the same buffers

        def step():
            graph.replay()
            return static_loss

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
        print(json.dumps({"impl": "pytorch_ddp_cudnn_baseline", "mode": a.mode, "metric": "IDC patches/sec (training step)",
                          "value": round(gb * a.steps / (ms / 1e3), 1), "unit": "patches/s", "n_gpus": size,
                          "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": round(ms / a.steps, 4),
                          "dtype": "bf16 autocast", "data": "synthetic",
                          "config": {"model": a.model, "global_batch": gb, "parallelism": f"ddp{size}",
                                     "torch": torch.__version__, "cudnn": torch.backends.cudnn.version()}}))
    if size > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
