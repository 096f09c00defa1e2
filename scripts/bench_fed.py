#!/usr/bin/env python
"""Round latency of the two federated workloads, fused kernels vs. the library (NCCL) baseline.

BASELINE.json configs 4 and 5:
* ``fedavg``  FedAvg (fed_model), MobileNetV2, one client per GPU, 1 local epoch per round;
* ``secure``  secure-aggregation FedAvg (secure_fed_model), 8 clients per GPU (64 over 8 GPUs), the 10x10 CNN,
              5 local epochs per round, pairwise-mask masked sum.

    python scripts/bench_fed.py --mode all                                     # 1 GPU (aggregation degenerates)
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/bench_fed.py --mode all

Every number is device time (CUDA events on the launching stream, synchronize on both sides), max over ranks.
Rank 0 prints one JSON line per (mode, comm).
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from idc_models_b200.parallel.world import World  # noqa: E402


def max_over_ranks(world, ms: float) -> float:
    t = torch.tensor([ms], dtype=torch.float64, device=world.device)
    if world.is_distributed:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def timed(world, fn, reps: int, warm: int = 2) -> float:
    """mean device milliseconds of fn() over `reps` calls, max over ranks"""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    world.barrier()
    tot = 0.0
    for _ in range(reps):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        world.barrier()
        torch.cuda.synchronize()
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        tot += s.elapsed_time(e)
    return max_over_ranks(world, tot / reps)


def bench_fedavg(world, args, comm: str) -> dict:
    from idc_models_b200.fed import fedavg as F
    from idc_models_b200.models.applications import build_transfer_model
    from idc_models_b200.optim.optimizers import RMSprop
    from idc_models_b200.parallel.strategy import FederatedStrategy
    from idc_models_b200.train.losses import BinaryCrossentropy
    from idc_models_b200.train.metrics import BinaryAccuracy
    dev = world.device
    strategy = FederatedStrategy(world)

    def model_fn():
        with strategy.scope():
            m = build_transfer_model(args.arch, (50, 50, 3), 1, device=dev, seed=0)
            m.base.trainable = True
        return F.from_keras_model(m, loss=BinaryCrossentropy(from_logits=True), metrics=[BinaryAccuracy()])

    fed = F.build_federated_averaging_process(model_fn, client_optimizer_fn=lambda: RMSprop(lr=1e-4), comm=comm)
    n_clients = world.size * args.clients_per_gpu
    g = torch.Generator(device=dev).manual_seed(100 + world.rank)
    B = args.batch_size
    data = []
    for cid in range(n_clients):
        if cid % world.size != world.rank:
            data.append([])          # other ranks' clients: never iterated here
            continue
        n = args.client_size
        x = torch.rand(n, 50, 50, 3, device=dev, generator=g)
        y = (torch.rand(n, device=dev, generator=g) > 0.5).float()
        data.append([(x[s:s + B], y[s:s + B]) for s in range(0, n, B)])
    state = fed.initialize()
    fed._work().model.executor.use_graphs = not args.no_graph    # replay the local fwd+bwd as a CUDA graph
    holder = {"state": state}

    def one_round():
        holder["state"], _ = fed.next(holder["state"], data)

    round_ms = timed(world, one_round, args.rounds, warm=1)
    # aggregation alone, on the buffers the last round left behind
    net = fed._work().model
    n = net.store.master.numel()
    k_local = args.clients_per_gpu
    clients, weights, gbuf, gshadow = fed._client_block(net, k_local)
    ranges = net.store.ranges(net.trainable_layer_indices())
    agg_ms = timed(world, lambda: fed._aggregate(net, holder["state"], clients, weights, gbuf, gshadow, k_local, ranges),
                   args.agg_reps, warm=3)
    n_tr = sum(e - b for b, e in ranges)
    N = world.size
    # bytes that must enter each GPU: its 1/N shard of every remote client + the broadcast of the other shards
    inbound = (N - 1) * k_local * (n_tr / N) * 4 + (N - 1) / N * n_tr * (4 + 2)
    roof_us = inbound / 900e9 * 1e6 if N > 1 else 0.0
    return {"bench": "fedavg_round", "comm": comm, "arch": args.arch, "n_gpus": N, "clients": n_clients,
            "client_size": args.client_size, "batch": B, "local_steps": -(-args.client_size // B),
            "round_ms": round(round_ms, 3), "aggregate_us": round(agg_ms * 1e3, 2), "trainable_floats": n_tr,
            "nvlink_roofline_us": round(roof_us, 2),
            "roofline_fraction": round(roof_us / (agg_ms * 1e3), 3) if N > 1 else None}


def bench_secure(world, args, comm: str) -> dict:
    from idc_models_b200.fed.secure_fused import FusedSecureFederation
    from idc_models_b200.workloads import create_small_cnn
    dev = world.device
    template = create_small_cnn(str(dev), seed=0, lr=1e-3)
    n_clients = world.size * args.secure_clients_per_gpu
    fed = FusedSecureFederation(world, n_clients, args.percent, template=template, lr=1e-3, batch_size=args.batch_size,
                                seed=0, comm=comm)
    g = torch.Generator(device=dev).manual_seed(200 + world.rank)
    for j in range(fed.k_local):
        x = torch.rand(args.secure_client_size, 10, 10, 3, device=dev, generator=g)
        y = (torch.rand(args.secure_client_size, device=dev, generator=g) > 0.5).float()
        fed.set_client_data(j, x, y)

    def one_round():
        fed.client_fit(args.secure_epochs)
        fed.aggregate()
        fed.client_update()

    round_ms = timed(world, one_round, args.rounds, warm=1)
    fit_ms = timed(world, lambda: fed.client_fit(args.secure_epochs), max(1, args.rounds // 2), warm=0)
    agg_ms = timed(world, fed.aggregate, args.agg_reps, warm=3)
    N = world.size
    pe = fed.protected_end
    inbound = (N - 1) * (pe / N) * 4 + (N - 1) / N * fed.stride * 4
    roof_us = inbound / 900e9 * 1e6 if N > 1 else 0.0
    # Philox work: every element of every resident client draws one word per other client
    words = fed.k_local * (n_clients - 1) * pe
    return {"bench": "secure_round", "comm": comm, "n_gpus": N, "clients": n_clients, "clients_per_gpu": fed.k_local,
            "client_size": args.secure_client_size, "epochs": args.secure_epochs, "percent": args.percent,
            "protected_floats": pe, "round_ms": round(round_ms, 3), "local_fit_ms": round(fit_ms, 3),
            "aggregate_us": round(agg_ms * 1e3, 2), "philox_words_per_gpu": words,
            "nvlink_roofline_us": round(roof_us, 3)}


def bench_secure_size(world, args) -> dict:
    """The masked-sum kernel alone at model size (default: MobileNetV2's 2.26 M floats, 8 clients per GPU): the regime
    where the pairwise Philox masks, not the NVLink traffic, are the cost. Needs >= 2 GPUs (pairs on
    one GPU cancel and generate nothing)."""
    import ctypes as C
    from idc_models_b200.ops import _lib
    from idc_models_b200.parallel.symm import SymmAllocator
    dev, N = world.device, world.size
    lib = _lib.cuda_lib()
    alloc = SymmAllocator(world)
    n = -(-args.secure_floats // 4) * 4
    k = args.secure_clients_per_gpu
    clients = alloc.alloc(k * n, torch.float32)
    clients.tensor.normal_(0, 0.05)
    ybuf = alloc.alloc(n, torch.int32)
    out = alloc.alloc(n, torch.float32)
    cw = alloc.comm_world()
    vp = C.c_void_p

    def launch():
        _lib.check(lib.idcb_secagg_masked_sum(C.byref(cw), clients.ptrs(), ybuf.ptrs(), ybuf.mc(), out.ptrs(), out.mc(),
                                              C.c_int(k), C.c_longlong(n), C.c_longlong(0), C.c_longlong(n), C.c_int(16),
                                              C.c_ulonglong(1234), _lib.stream()))

    ms = timed(world, launch, args.agg_reps, warm=3)
    want = clients.tensor.view(k, n).double().sum(0)
    if world.is_distributed:
        dist.all_reduce(want)
    err = (out.tensor[:n].double() - want / (k * N)).abs().max().item()
    words = k * (N - 1) * k * n                       # 32-bit mask words generated per GPU
    philox_calls = words / 4
    # Philox4x32-10: 10 rounds x (2 wide multiplies + 4 xor + 2 key adds) ~ 100 integer operations per call; INT32 peak of a
    # B200 ~ 148 SMs x 128 lanes x 1.965 GHz
    int_ops = philox_calls * 100
    peak = 148 * 128 * 1.965e9
    return {"bench": "secure_masked_sum_at_model_size", "n_gpus": N, "clients_per_gpu": k, "clients": k * N, "floats": n,
            "kernel_us": round(ms * 1e3, 1), "max_abs_err_vs_float_mean": err, "mask_words_per_gpu": words,
            "mask_words_per_s": round(words / (ms * 1e-3), 1), "int32_ops_estimate": int_ops,
            "fraction_of_int32_peak_estimate": round(int_ops / (ms * 1e-3) / peak, 3),
            "nvlink_bytes_in_per_gpu": (N - 1) / N * n * 8}


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="all", choices=["fedavg", "secure", "secure_size", "all"])
    ap.add_argument("--secure-floats", type=int, default=2257984, help="secure_size: protected floats (MobileNetV2)")
    ap.add_argument("--comm", default="both", choices=["fused", "nccl", "both"])
    ap.add_argument("--arch", default="mobilenet_v2")
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--agg-reps", type=int, default=20)
    ap.add_argument("--clients-per-gpu", type=int, default=1)
    ap.add_argument("--client-size", type=int, default=3000)
    ap.add_argument("--batch-size", type=int, default=32)
    ap.add_argument("--secure-clients-per-gpu", type=int, default=8)
    ap.add_argument("--secure-client-size", type=int, default=300)
    ap.add_argument("--secure-epochs", type=int, default=5)
    ap.add_argument("--percent", type=float, default=1.0)
    ap.add_argument("--no-graph", action="store_true")
    args = ap.parse_args()
    world = World.from_env()
    if world.device.type != "cuda":
        print("bench_fed.py needs CUDA devices", file=sys.stderr)
        return 2
    comms = ["fused", "nccl"] if args.comm == "both" else [args.comm]
    modes = ["fedavg", "secure"] if args.mode == "all" else [args.mode]
    if args.mode == "secure_size":
        res = bench_secure_size(world, args)
        if world.rank == 0:
            print(json.dumps(res), flush=True)
        modes = []
    for mode in modes:
        for comm in comms:
            res = (bench_fedavg if mode == "fedavg" else bench_secure)(world, args, comm)
            if world.rank == 0:
                print(json.dumps(res), flush=True)
    world.barrier()
    if world.is_distributed:
        world.shutdown()
    return 0


if __name__ == "__main__":
    sys.exit(main())
