#!/usr/bin/env python
"""Library baseline for the FedAvg round latency (BASELINE.json config 4): one client per GPU, each trains a
torchvision MobileNetV2 (weights=None, channels-last, bf16 autocast, RMSprop) for one local epoch over its own
synthetic 50x50x3 patches, then the server model is the example-weighted mean of the client models, computed with an
NCCL all-reduce of the flattened parameters. None of this repository's kernels are on this path.

Two arms: ``--mode eager`` (plain PyTorch) and ``--mode graph`` (the local training step captured once in a CUDA graph
and replayed, capturable optimizer) — the library stack at its best for a launch-bound batch-32 step.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 baseline/fed_baseline.py --mode graph
"""
import argparse
import json
import os

import torch
import torch.distributed as dist
import torch.nn as nn


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="eager", choices=["eager", "graph"])
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--client-size", type=int, default=3000)
    ap.add_argument("--batch-size", type=int, default=32)
    a = ap.parse_args()
    import torchvision
    size = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    lrank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lrank)
    dev = torch.device("cuda", lrank)
    if size > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29556")
        dist.init_process_group("nccl", rank=rank, world_size=size, device_id=dev)
    torch.manual_seed(0)
    net = torchvision.models.mobilenet_v2(weights=None)
    net.classifier = nn.Sequential(nn.Linear(1280, 1))
    net = net.to(dev).to(memory_format=torch.channels_last)
    lossf = nn.BCEWithLogitsLoss()
    B = a.batch_size
    steps = -(-a.client_size // B)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    data_x = torch.rand(steps * B, 3, 50, 50, device=dev, generator=g).to(memory_format=torch.channels_last)
    data_y = (torch.rand(steps * B, 1, device=dev, generator=g) > 0.5).float()
    params = [p for p in net.parameters()]
    server = torch.cat([p.detach().reshape(-1) for p in params]).clone()
    x = torch.zeros(B, 3, 50, 50, device=dev).to(memory_format=torch.channels_last)
    y = torch.zeros(B, 1, device=dev)

    def make_opt():
        return torch.optim.RMSprop(net.parameters(), lr=1e-4, alpha=0.9, eps=1e-7, capturable=(a.mode == "graph"))

    opt = make_opt()

    def train_step():
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = net(x)
        loss = lossf(out.float(), y)
        loss.backward()
        opt.step()

    graph = None
    if a.mode == "graph":
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                train_step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        opt.zero_grad(set_to_none=True)
        with torch.cuda.graph(graph):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = net(x)
            lossf(out.float(), y).backward()
            opt.step()

    def load(flat):
        off = 0
        with torch.no_grad():
            for p in params:
                n = p.numel()
                p.copy_(flat[off:off + n].view_as(p))
                off += n

    def one_round():
        nonlocal server
        load(server)
        # fresh optimizer state per client per round (TFF semantics): zero the state in place (keeps graph addresses)
        for st in opt.state.values():
            for v in st.values():
                if torch.is_tensor(v):
                    v.zero_()
        for s in range(steps):
            x.copy_(data_x[s * B:(s + 1) * B])
            y.copy_(data_y[s * B:(s + 1) * B])
            if graph is not None:
                graph.replay()
            else:
                train_step()
        flat = torch.cat([p.detach().reshape(-1) for p in params])
        w = float(steps * B)
        acc = (flat - server) * w
        tot = torch.tensor([w], device=dev)
        if size > 1:
            dist.all_reduce(acc)
            dist.all_reduce(tot)
        server = server + acc / tot

    one_round()
    torch.cuda.synchronize()
    if size > 1:
        dist.barrier(device_ids=[lrank])
    ms = 0.0
    for _ in range(a.rounds):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        s.record()
        one_round()
        e.record()
        torch.cuda.synchronize()
        ms += s.elapsed_time(e)
    t = torch.tensor([ms / a.rounds], dtype=torch.float64, device=dev)
    if size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"bench": "fedavg_round", "impl": "pytorch_cudnn_nccl_baseline", "mode": a.mode, "arch": "mobilenet_v2",
                          "n_gpus": size, "clients": size, "client_size": a.client_size, "batch": B, "local_steps": steps,
                          "round_ms": round(float(t.item()), 3), "dtype": "bf16 autocast"}))
    if size > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
