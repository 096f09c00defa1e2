"""Federated averaging with the TensorFlow-Federated surface the reference drives
(/root/reference/fed_model.py:196-229):

    fed_avg = build_federated_averaging_process(model_fn, client_optimizer_fn=lambda: RMSprop(lr=1e-4))
    evaluation = build_federated_evaluation(model_fn)
    state = fed_avg.initialize()
    state = state_with_new_model_weights(state, trainable_weights, non_trainable_weights)
    state, train_metrics = fed_avg.next(state, federated_train_data)   # -> (binary_accuracy, loss)
    test_metrics = evaluation(state.model, federated_test_data)

Round semantics (TFF): broadcast the server model; each client starts from it with a FRESH optimizer,
makes ONE pass over its dataset, and reports ``delta = w_after - w_before`` on the trainable variables
weighted by its example count (a client with a non-finite delta is dropped: zero delta, weight 0); the
server takes the example-weighted mean of the deltas and applies it with SGD(lr=1.0), i.e. the weighted
average of the client models. Non-trainable variables (frozen layers, BatchNorm moving statistics) are not
aggregated. Metrics are example-weighted sums over clients.

Clients are mapped round-robin to ranks (one process per GPU); the aggregation runs in ONE kernel over all
clients resident on all GPUs (csrc/comm.cu:fedavg_reduce_kernel) which also broadcasts the new server
model into every rank's buffers — i.e. the next round's broadcast is free. On CPU the same algebra runs in
PyTorch (and is the oracle for the kernel).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from ..ops import _lib
from ..optim.optimizers import SGD
from ..train import engine


@dataclass
class ModelWeights:
    trainable: list
    non_trainable: list


class ServerState:
    """The FL server state: the flat master weights + moving statistics of the global model."""

    def __init__(self, master: torch.Tensor, stats: torch.Tensor, net):
        self.master, self.stats, self._net = master, stats, net

    @property
    def model(self) -> "ServerState":
        return self

    def model_weights(self) -> ModelWeights:
        net = self._net
        saved_m, saved_s = net.store.master.clone(), net.store.stats.clone()
        net.store.master.copy_(self.master)
        net.store.stats.copy_(self.stats)
        mw = ModelWeights([t.clone() for t in net.trainable_weights], [t.clone() for t in net.non_trainable_weights])
        net.store.master.copy_(saved_m)
        net.store.stats.copy_(saved_s)
        return mw


class _Wrapped:
    def __init__(self, model, loss, metrics):
        self.model, self.loss, self.metrics = model, loss, metrics


def from_keras_model(model, dummy_batch=None, loss=None, metrics=None) -> _Wrapped:
    """``tff.learning.from_keras_model`` (fed_model.py:201-205): bundle a model with its loss/metrics."""
    return _Wrapped(model, loss, metrics or [])


def state_with_new_model_weights(state: ServerState, trainable_weights, non_trainable_weights) -> ServerState:
    """Inject pre-trained weights into the server state (fed_model.py:219-223)."""
    net = state._net
    tw, ntw = net.trainable_weights, net.non_trainable_weights
    if len(tw) != len(trainable_weights) or len(ntw) != len(non_trainable_weights):
        raise ValueError("weight lists do not match the model structure")
    saved_m, saved_s = net.store.master.clone(), net.store.stats.clone()
    net.store.master.copy_(state.master)
    net.store.stats.copy_(state.stats)
    for dst, src in list(zip(tw, trainable_weights)) + list(zip(ntw, non_trainable_weights)):
        dst.copy_(torch.as_tensor(src).to(dst.device, dst.dtype).reshape(dst.shape))
    new = ServerState(net.store.master.clone(), net.store.stats.clone(), net)
    net.store.master.copy_(saved_m)
    net.store.stats.copy_(saved_s)
    return new


def weighted_delta_mean(w0: torch.Tensor, clients: torch.Tensor, weights: torch.Tensor, server_lr: float = 1.0,
                        ranges=None) -> torch.Tensor:
    """PyTorch statement of the aggregation (oracle): w0 + lr * sum_k n_k (w_k - w0) / sum_k n_k with
    non-finite clients dropped. ``clients`` is [K, n]; only ``ranges`` (list of (b, e)) are aggregated."""
    out = w0.clone()
    ranges = ranges or [(0, w0.numel())]
    wts = weights.clone().to(torch.float64)
    for k in range(clients.shape[0]):
        if not all(torch.isfinite(clients[k, b:e]).all() for b, e in ranges):
            wts[k] = 0.0
    tot = wts.sum()
    if tot <= 0:
        return out
    for b, e in ranges:
        acc = torch.zeros(e - b, dtype=torch.float64, device=w0.device)
        for k in range(clients.shape[0]):
            if wts[k] > 0:
                acc += wts[k] * (clients[k, b:e].double() - w0[b:e].double())
        out[b:e] = (w0[b:e].double() + server_lr * acc / tot).float()
    return out


def _local_clients(federated_data: list, world, placement: str) -> list:
    """The client datasets this rank trains. ``round_robin``: client i -> rank i % N. ``balanced``: longest-
    processing-time-first on the clients' batch counts (fed/scheduler.py, C++), for clients of unequal size; every
    rank must then hold every client's dataset object (its length is the cost)."""
    if not world.is_distributed:
        return list(federated_data)
    if placement == "balanced":
        from .scheduler import clients_of_rank
        costs = [len(ds) if hasattr(ds, "__len__") else 1 for ds in federated_data]
        return [federated_data[i] for i in clients_of_rank(costs, world.size, world.rank)]
    return federated_data[world.rank::world.size]


def _sum_over_ranks(world, t: torch.Tensor, fused: bool = True) -> torch.Tensor:
    """Federated metric aggregation: on GPUs the three scalars are summed by the peer-memory
    kernel (parallel/symm.py:ScalarReducer), elsewhere by the library collective."""
    if not world.is_distributed:
        return t
    if fused and t.is_cuda:
        from ..parallel.symm import ScalarReducer
        return ScalarReducer.for_world(world).sum(t)
    return world.all_reduce_sum(t)


class FederatedAveragingProcess:
    def __init__(self, model_fn, client_optimizer_fn, server_optimizer_fn=None, strategy=None, comm: str = "fused",
                 placement: str = "round_robin"):
        if placement not in ("round_robin", "balanced"):
            raise ValueError("placement must be 'round_robin' or 'balanced'")
        self.placement = placement
        if comm not in ("fused", "nccl"):
            raise ValueError("comm must be 'fused' (one kernel over peer memory) or 'nccl' (library all-reduce baseline)")
        self.comm = comm
        self.model_fn = model_fn
        self.client_optimizer_fn = client_optimizer_fn
        sopt = server_optimizer_fn() if server_optimizer_fn is not None else SGD(learning_rate=1.0)
        if not isinstance(sopt, SGD) or sopt.momentum != 0.0:
            raise NotImplementedError("the server optimizer is SGD without momentum (TFF default, lr=1.0)")
        self.server_lr = sopt.learning_rate
        self._wrapped = None
        self._block = None
        self._symm = None

    # ---- lazily built working model (one per process, reused for every local client) -----------------
    def _work(self):
        if self._wrapped is None:
            w = self.model_fn()
            self._wrapped = w if isinstance(w, _Wrapped) else _Wrapped(w, None, [])
        return self._wrapped

    def initialize(self) -> ServerState:
        net = self._work().model
        net.strategy.sync_variables(net.store)
        return ServerState(net.store.master.clone(), net.store.stats.clone(), net)

    def _client_block(self, net, k_local: int):
        n = net.store.master.numel()
        world = net.strategy.world
        if self._block is None or self._block[0].tensor.numel() < k_local * n:
            if net.device.type == "cuda" and self.comm == "fused":
                from ..parallel.symm import SymmAllocator
                self._symm = SymmAllocator(world)
                self._block = (self._symm.alloc(k_local * n, torch.float32), self._symm.alloc(max(k_local, 4), torch.float32),
                               self._symm.alloc(n, torch.float32), self._symm.alloc(n, torch.bfloat16))
            else:
                from ..parallel.symm import SymmBuffer
                mk = lambda m, dt: SymmBuffer(torch.zeros(m, dtype=dt, device=net.device), [0])  # noqa: E731
                self._block = (mk(k_local * n, torch.float32), mk(max(k_local, 4), torch.float32),
                               mk(n, torch.float32), mk(n, torch.bfloat16))
        return self._block

    def next(self, state: ServerState, federated_data: list):
        """One round over ``federated_data`` (list of per-client datasets). Returns
        ``(new_state, (binary_accuracy, loss))`` — TFF's metric order (fed_model.py:229)."""
        wr = self._work()
        net = wr.model
        world = net.strategy.world
        n = net.store.master.numel()
        mine = _local_clients(federated_data, world, self.placement)
        k_counts = world.all_gather_object(len(mine))
        k_local = max(k_counts)
        clients, weights, gbuf, gshadow = self._client_block(net, k_local)
        weights.tensor.zero_()
        msum = torch.zeros(3, dtype=torch.float64, device=net.device)   # accumulated on the device: no sync per step
        for k, ds in enumerate(mine):
            net.store.master.copy_(state.master)
            net.store.stats.copy_(state.stats)
            net.store.refresh_shadow()
            # fresh optimizer state per client per round (TFF); the executor plans are kept
            if net.optimizer is None or net._dirty:
                net.compile(optimizer=self.client_optimizer_fn(), loss=wr.loss, metrics=wr.metrics)
            else:
                engine.rebind_optimizer(net, self.client_optimizer_fn())
            n_ex = 0
            for x, y in ds:
                stats, _ = net.executor.train_batch(x, y, x.shape[0])
                net.optimizer.apply(1.0)
                msum += stats.detach()[:3].to(torch.float64)
                n_ex += int(x.shape[0])
            clients.tensor[k * n:(k + 1) * n].copy_(net.store.master)
            weights.tensor[k] = float(n_ex)
        ranges = net.store.ranges(net.trainable_layer_indices())
        new_master = self._aggregate(net, state, clients, weights, gbuf, gshadow, k_local, ranges)
        msum = _sum_over_ranks(world, msum, fused=self.comm == "fused").cpu()
        cnt = max(float(msum[2]), 1.0)
        metrics = (float(msum[1]) / cnt, float(msum[0]) / cnt)
        return ServerState(new_master, state.stats.clone(), net), metrics

    def _aggregate(self, net, state, clients, weights, gbuf, gshadow, k_local, ranges) -> torch.Tensor:
        n = net.store.master.numel()
        world = net.strategy.world
        if net.device.type == "cuda" and self.comm == "nccl":
            return self._aggregate_library(net, state, clients, weights, k_local, ranges)
        if net.device.type != "cuda":
            cl = clients.tensor[:k_local * n].view(k_local, n)
            wt = weights.tensor[:k_local]
            if world.is_distributed:   # gloo: gather every rank's block
                blocks = world.all_gather_object((cl.clone(), wt.clone()))
                cl = torch.cat([b[0] for b in blocks]); wt = torch.cat([b[1] for b in blocks])
            return weighted_delta_mean(state.master, cl, wt, self.server_lr, ranges)
        lib = _lib.cuda_lib()
        base = state.master
        # everything that does not change from round to round is marshalled once (the launch below is the whole
        # aggregation, so the Python in front of it is what the GPU waits for)
        key = (tuple(sorted(ranges)), n, k_local, id(clients), id(weights), id(gbuf), id(gshadow))
        ra = getattr(self, "_round_args", None)
        if ra is None or ra["key"] != key:
            rs = sorted(ranges)
            ra = self._round_args = {
                "key": key, "keep": (clients, weights, gbuf, gshadow),
                "ok": len(rs) <= 16 and all(b % 4 == 0 and (e - b) % 4 == 0 for b, e in rs),
                "cw": self._symm.comm_world(use_barrier=True),     # a world of one GPU still needs the grid barrier
                "n_ranges": C.c_int(len(rs)),
                "offs": (C.c_longlong * max(len(rs), 1))(*[b for b, _ in rs]),
                "lens": (C.c_longlong * max(len(rs), 1))(*[e - b for b, e in rs]),
                "ptrs": (clients.ptrs(), weights.ptrs(), gbuf.ptrs(), gshadow.ptrs(), gbuf.mc(), gshadow.mc()),
            }
        if ra["ok"] and base.dtype == torch.float32 and base.is_contiguous() and base.is_cuda:
            # one launch: finite scan of the local clients, every range's weighted mean (multicast into every rank),
            # and this rank's private copy of the complete new model
            out = torch.empty_like(base)
            cp, wp, gp, sp, gmc, smc = ra["ptrs"]
            _lib.check(lib.idcb_fedavg_round(C.byref(ra["cw"]), cp, wp, gp, sp, gmc, smc, C.c_int(k_local),
                                             C.c_longlong(n), ra["n_ranges"], ra["offs"], ra["lens"],
                                             C.c_float(self.server_lr), C.c_void_p(base.data_ptr()),
                                             C.c_void_p(out.data_ptr()), C.c_longlong(n), _lib.stream()))
            _lib.count_launch()
            return out
        cw = self._symm.comm_world()
        gbuf.tensor[:n].copy_(base)
        for b, e in ranges:
            _lib.check(lib.idcb_fedavg_finite_flags(C.c_void_p(clients.tensor.data_ptr()), C.c_longlong(n),
                                                    C.c_longlong(b), C.c_longlong(e - b), C.c_int(k_local),
                                                    C.c_void_p(weights.tensor.data_ptr()), _lib.stream()))
            _lib.count_launch()
        for b, e in ranges:
            _lib.check(lib.idcb_fedavg_reduce(C.byref(cw), clients.ptrs(), weights.ptrs(), gbuf.ptrs(), gshadow.ptrs(),
                                              gbuf.mc(), gshadow.mc(), C.c_int(k_local), C.c_longlong(n),
                                              C.c_longlong(b), C.c_longlong(e - b), C.c_float(self.server_lr),
                                              _lib.stream()))
            _lib.count_launch()
        return gbuf.tensor[:n].clone()


    def _aggregate_library(self, net, state, clients, weights, k_local, ranges) -> torch.Tensor:
        """The baseline the fused kernel is measured against: PyTorch elementwise kernels for the local
        weighted delta sum + NCCL all-reduce of (sum, weight) + PyTorch kernels for the server update."""
        n = net.store.master.numel()
        world = net.strategy.world
        cl = clients.tensor[:k_local * n].view(k_local, n)
        wt = weights.tensor[:k_local].clone()
        new = state.master.clone()
        for b, e in ranges:
            finite = torch.isfinite(cl[:, b:e]).all(dim=1)
            wt = torch.where(finite, wt, torch.zeros_like(wt))
        tot = wt.sum().reshape(1)
        if world.is_distributed:
            world.all_reduce_sum(tot)
        for b, e in ranges:
            d = torch.where(wt[:, None] > 0, cl[:, b:e] - state.master[None, b:e], torch.zeros((), device=cl.device))
            acc = (wt[:, None] * d).sum(dim=0)
            if world.is_distributed:
                world.all_reduce_sum(acc)
            new[b:e] = torch.where(tot > 0, state.master[b:e] + self.server_lr * acc / tot.clamp_min(1e-30), state.master[b:e])
        return new


def build_federated_averaging_process(model_fn, client_optimizer_fn, server_optimizer_fn=None, comm: str = "fused",
                                      placement: str = "round_robin"):
    return FederatedAveragingProcess(model_fn, client_optimizer_fn, server_optimizer_fn, comm=comm, placement=placement)


def build_federated_evaluation(model_fn, placement: str = "round_robin"):
    """-> evaluation(server_model, federated_data) -> (binary_accuracy, loss), example-weighted over the
    clients' full datasets (fed_model.py:210,224,228)."""
    holder = {}

    def evaluation(server_model: ServerState, federated_data: list):
        if "w" not in holder:
            w = model_fn()
            holder["w"] = w if isinstance(w, _Wrapped) else _Wrapped(w, None, [])
        wr = holder["w"]
        net = wr.model
        world = net.strategy.world
        net.store.master.copy_(server_model.master)
        net.store.stats.copy_(server_model.stats)
        net.store.refresh_shadow()
        if net.optimizer is None:
            net.compile(optimizer=SGD(0.0), loss=wr.loss, metrics=wr.metrics)
        mine = _local_clients(federated_data, world, placement)
        msum = torch.zeros(3, dtype=torch.float64, device=net.device)   # stays on the device: no sync per batch
        for ds in mine:
            for x, y in ds:
                stats, _ = net.executor.eval_batch(x, y)
                msum += stats.detach()[:3].to(torch.float64)
        msum = _sum_over_ranks(world, msum).cpu()
        cnt = max(float(msum[2]), 1.0)
        return (float(msum[1]) / cnt, float(msum[0]) / cnt)
    return evaluation
