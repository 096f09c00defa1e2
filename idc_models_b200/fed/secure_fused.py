"""Secure-aggregation FedAvg with the clients mapped onto GPUs (BASELINE.json config 5: 64 clients over
8 x B200; SURVEY.md 2.D "Federated", 2.E C9/C10, 8.6, 8.14).

The reference trains its clients one after another in one process, encrypts each weight tensor scalar by
scalar, and averages ciphertexts on the host (/root/reference/secure_fed_model.py:212-236). Here

* client ``c`` lives on rank ``c % world`` (local slot ``c // world``); all clients resident on a GPU train
  **concurrently, one kernel launch per step** (csrc/small_cnn.cu, ``blockIdx.y`` = client) out of one
  ``[k_local][stride]`` parameter block, followed by one RMSprop launch over the whole block (optimizer state
  persists across rounds, as the reference's long-lived model objects do, secure_fed_model.py:103,133);
* the aggregation of the protected tensor prefix (``int(n_tensors * percent)`` tensors, the CLI meaning of
  ``percent``, secure_fed_model.py:115-121) is ONE kernel over every client of every GPU
  (csrc/comm.cu:secagg_kernel): quantise, add pairwise Philox masks generated in registers, local ring sum,
  cross-GPU ring sum (``multimem.ld_reduce.add.u32`` when the fabric has NVLS), de-quantise, unweighted mean
  (``Server.aggregate``, secure_fed_model.py:160-168), broadcast into every rank;
* the unprotected remainder is averaged in float by the FedAvg kernel (csrc/comm.cu:fedavg_reduce_kernel)
  with equal client weights;
* ``client_update`` is a broadcast of the averaged block into every resident client's slot.

``comm='nccl'`` keeps a library-only aggregation (masks materialised by PyTorch kernels, NCCL integer
all-reduce) as the baseline the fused path is measured against.
"""
from __future__ import annotations

import ctypes as C

import torch

from ..ops import _lib
from ..optim.optimizers import KIND
from .secure import FRAC_BITS

vp = C.c_void_p


class FusedSecureFederation:
    def __init__(self, world, n_clients: int, percent: float, *, template, lr: float = 1e-3, rho: float = 0.9,
                 eps: float = 1e-7, batch_size: int = 32, seed: int = 0, frac_bits: int = FRAC_BITS,
                 comm: str = "fused", use_multicast: bool = True):
        """``template``: a compiled small-CNN :class:`Network` (workloads.create_small_cnn) whose flat layout
        and initial weights every client starts from."""
        if world.device.type != "cuda":
            raise RuntimeError("FusedSecureFederation needs CUDA devices (the host Client/Server classes run anywhere)")
        if n_clients % world.size:
            raise ValueError(f"{n_clients} clients cannot be spread evenly over {world.size} GPUs")
        if comm not in ("fused", "nccl"):
            raise ValueError("comm must be 'fused' or 'nccl'")
        from ..parallel.symm import SymmAllocator
        self.world, self.n_clients, self.percent = world, n_clients, percent
        self.k_local = n_clients // world.size
        self.comm = comm
        self.lr, self.rho, self.eps = lr, rho, eps
        self.batch_size, self.seed, self.frac_bits = batch_size, seed, frac_bits
        self.template = template
        st = template.store
        self.stride = st.master.numel()
        names = [p.name for l in template.layers for p in l.spec.params]
        self.names = names
        self.offsets = (C.c_int * 6)(*[st.slots[n].offset for n in names])
        self.cstride = st.slots[names[0]].storage_shape[-1]
        self.rates = [l.spec.attrs["rate"] for l in template.layers if l.spec.kind == "dropout"]
        npro = int(len(names) * percent)
        # the protected prefix is a contiguous range of the flat block (slots are laid out in tensor order)
        self.n_protected_tensors = npro
        self.protected_end = 0 if npro == 0 else (st.slots[names[npro]].offset if npro < len(names) else self.stride)
        dev = world.device
        self.alloc = SymmAllocator(world, use_multicast)
        k, n = self.k_local, self.stride
        self.params = self.alloc.alloc(k * n, torch.float32)
        self.wts = self.alloc.alloc(max(k, 4), torch.float32)
        self.ybuf = self.alloc.alloc(max(self.protected_end, 4), torch.int32)
        self.out = self.alloc.alloc(n, torch.float32)
        self.out_shadow = self.alloc.alloc(n, torch.bfloat16)
        self.grads = torch.zeros(k * n, dtype=torch.float32, device=dev)
        self.ms = torch.zeros(k * n, dtype=torch.float32, device=dev)
        self.acc = torch.zeros(k, 4, dtype=torch.float32, device=dev)
        self.params.tensor[:k * n].view(k, n).copy_(st.master[None, :].expand(k, n))
        self.wts.tensor[:k] = 1.0
        self.step = 0
        self.round = 0
        self.data = [None] * k
        self._logits = None
        torch.cuda.synchronize()
        world.barrier()

    # ---- data ------------------------------------------------------------------------------------------
    def client_ids(self) -> list:
        return [self.world.rank + self.world.size * j for j in range(self.k_local)]

    def set_client_data(self, local_slot: int, x: torch.Tensor, y: torch.Tensor) -> None:
        """x [n,10,10,3] float32 in [0,1], y [n] 0/1. Every resident client needs the same example count
        (the reference gives every client ``CLIENT_SIZE`` examples, secure_fed_model.py:46-48)."""
        dev = self.world.device
        self.data[local_slot] = (x.to(dev, torch.float32).contiguous(), y.to(dev, torch.float32).reshape(-1).contiguous())

    # ---- local training: all resident clients per launch -------------------------------------------------
    def _launch(self, x, y, batch: int, training: bool):
        lib = _lib.cuda_lib()
        k = self.k_local
        if self._logits is None or self._logits.numel() < k * batch:
            self._logits = torch.zeros(k * batch, dtype=torch.float32, device=self.world.device)
        _lib.check(lib.idcb_small_cnn_step(
            vp(x.data_ptr()), vp(y.data_ptr()), vp(self.params.tensor.data_ptr()), vp(self.grads.data_ptr()),
            vp(self._logits.data_ptr()), vp(self.acc.data_ptr()), C.c_int(k), C.c_int(batch), C.c_int(self.stride),
            self.offsets, C.c_int(self.cstride), C.c_int(1 if training else 0), C.c_float(1.0 / max(batch, 1)),
            C.c_float(self.rates[0]), C.c_float(self.rates[1]), C.c_ulonglong(self.seed), C.c_uint(self.step),
            C.c_float(self.template.acc_threshold), C.c_int(1 if self.template.acc_on_prob else 0), _lib.stream()))
        _lib.count_launch()

    def _launch_indexed(self, X, Y, perm, col0: int, n_ex: int, batch: int) -> None:
        lib = _lib.cuda_lib()
        k = self.k_local
        if self._logits is None or self._logits.numel() < k * batch:
            self._logits = torch.zeros(k * batch, dtype=torch.float32, device=self.world.device)
        _lib.check(lib.idcb_small_cnn_step_indexed(
            vp(X.data_ptr()), vp(Y.data_ptr()), vp(perm.data_ptr() + 8 * col0), C.c_longlong(perm.shape[1]), C.c_int(n_ex),
            vp(self.params.tensor.data_ptr()), vp(self.grads.data_ptr()), vp(self._logits.data_ptr()),
            vp(self.acc.data_ptr()), C.c_int(k), C.c_int(batch), C.c_int(self.stride), self.offsets,
            C.c_int(self.cstride), C.c_int(1), C.c_float(1.0 / max(batch, 1)), C.c_float(self.rates[0]),
            C.c_float(self.rates[1]), C.c_ulonglong(self.seed), C.c_uint(self.step),
            C.c_float(self.template.acc_threshold), C.c_int(1 if self.template.acc_on_prob else 0), _lib.stream()))
        _lib.count_launch()

    def client_fit(self, epochs: int = 5) -> torch.Tensor:
        """Every resident client trains ``epochs`` passes over its shard. Returns the per-client
        [k_local, 3] (loss sum, correct, count) accumulators of the last epoch (device tensor)."""
        if any(d is None for d in self.data):
            raise RuntimeError("set_client_data() has not been called for every resident client")
        n_ex = {d[0].shape[0] for d in self.data}
        if len(n_ex) != 1:
            raise ValueError("resident clients must hold the same number of examples")
        n_ex = n_ex.pop()
        k, B = self.k_local, self.batch_size
        lib = _lib.cuda_lib()
        X = torch.stack([d[0] for d in self.data])           # [k, n, 10, 10, 3]
        Y = torch.stack([d[1] for d in self.data])
        gen = torch.Generator(device=X.device).manual_seed(self.seed * 7919 + self.round * 104729 + self.world.rank)
        for _ in range(epochs):
            self.acc.zero_()
            perm = torch.stack([torch.randperm(n_ex, device=X.device, generator=gen) for _ in range(k)])   # int64 [k, n_ex]
            for s in range(0, n_ex, B):
                b = min(B, n_ex - s)
                self.step += 1
                # the batch is gathered inside the kernel from the resident shards (no index-gather launches per step)
                self._launch_indexed(X, Y, perm, s, n_ex, b)
                _lib.check(lib.idcb_optim_step(C.c_int(KIND["rmsprop"]), vp(self.params.tensor.data_ptr()),
                                               vp(self.grads.data_ptr()), vp(self.ms.data_ptr()), vp(0), vp(0),
                                               C.c_longlong(k * self.stride), C.c_float(self.lr), C.c_float(self.rho),
                                               C.c_float(0.0), C.c_float(self.eps), C.c_float(1.0), C.c_int(1),
                                               _lib.stream()))
                _lib.count_launch()
        self.round += 1
        return self.acc[:, :3].clone()

    # ---- aggregation -------------------------------------------------------------------------------------
    def _round_seed(self) -> int:
        return (self.seed * 1000003 + self.round) & 0xFFFFFFFFFFFFFFFF

    def aggregate(self) -> torch.Tensor:
        """-> this rank's copy of the averaged parameter block (every rank receives the same values)."""
        if self.comm == "nccl":
            return self._aggregate_library()
        lib = _lib.cuda_lib()
        cw = self.alloc.comm_world()
        k, n, pe = self.k_local, self.stride, self.protected_end
        if pe > 0:
            _lib.check(lib.idcb_secagg_masked_sum(C.byref(cw), self.params.ptrs(), self.ybuf.ptrs(), self.ybuf.mc(),
                                                  self.out.ptrs(), self.out.mc(), C.c_int(k), C.c_longlong(n),
                                                  C.c_longlong(0), C.c_longlong(pe), C.c_int(self.frac_bits),
                                                  C.c_ulonglong(self._round_seed()), _lib.stream()))
            _lib.count_launch()
        if pe < n:
            # unprotected tensors: plain float mean (equal weights, base 0 -> the kernel's weighted mean)
            # (no host barrier: the kernel reads only THIS rank's base range, which this stream has just zeroed, and
            # no peer can store into it before the kernel's opening cross-GPU barrier, i.e. before this rank's kernel —
            # hence its memset — has started)
            self.out.tensor[pe:n].zero_()
            _lib.check(lib.idcb_fedavg_reduce(C.byref(cw), self.params.ptrs(), self.wts.ptrs(), self.out.ptrs(),
                                              self.out_shadow.ptrs(), self.out.mc(), self.out_shadow.mc(), C.c_int(k),
                                              C.c_longlong(n), C.c_longlong(pe), C.c_longlong(n - pe), C.c_float(1.0),
                                              _lib.stream()))
            _lib.count_launch()
        return self.out.tensor[:n]

    def _aggregate_library(self) -> torch.Tensor:
        """Baseline: pairwise masks materialised in HBM by PyTorch's Philox generator, one tensor per pair,
        ring arithmetic in int64 kernels, NCCL all-reduce of the masked partial sums."""
        k, n, pe = self.k_local, self.stride, self.protected_end
        dev = self.world.device
        P = self.params.tensor[:k * n].view(k, n)
        out = self.out.tensor[:n]
        if pe > 0:
            q = torch.round(P[:, :pe].double() * (1 << self.frac_bits)).to(torch.int64)
            gen = torch.Generator(device=dev)
            total = torch.zeros(pe, dtype=torch.int64, device=dev)
            for j, c in enumerate(self.client_ids()):
                y = q[j].clone()
                for other in range(self.n_clients):
                    if other == c:
                        continue
                    lo, hi = (c, other) if c < other else (other, c)
                    gen.manual_seed((self._round_seed() * 4099 + lo * 65537 + hi) & 0x7FFFFFFFFFFFFFFF)
                    m = torch.randint(0, 1 << 32, (pe,), dtype=torch.int64, device=dev, generator=gen)
                    y = y + m if c < other else y - m
                total += y
            total &= 0xFFFFFFFF
            if self.world.is_distributed:
                self.world.all_reduce_sum(total)
            total &= 0xFFFFFFFF
            signed = torch.where(total >= (1 << 31), total - (1 << 32), total)
            out[:pe] = (signed.double() / ((1 << self.frac_bits) * self.n_clients)).float()
        if pe < n:
            s = P[:, pe:].sum(dim=0)
            if self.world.is_distributed:
                self.world.all_reduce_sum(s)
            out[pe:] = s / self.n_clients
        return out

    def client_update(self) -> None:
        """Every resident client adopts the averaged model (secure_fed_model.py:143-149, 232-233)."""
        k, n = self.k_local, self.stride
        self.params.tensor[:k * n].view(k, n).copy_(self.out.tensor[None, :n].expand(k, n))

    # ---- evaluation (client 0's view, secure_fed_model.py:235) --------------------------------------------
    def evaluate(self, x: torch.Tensor, y: torch.Tensor, steps: int | None = 20, with_auc: bool = False):
        """loss, accuracy (and, with ``with_auc``, the reference's per-batch sklearn AUROC averaged over batches,
        secure_fed_model.py:81-82) of local slot 0's model over ``steps`` batches of (x, y)."""
        dev = self.world.device
        x = x.to(dev, torch.float32)
        y = y.to(dev, torch.float32).reshape(-1)
        B = self.batch_size
        tot = torch.zeros(4, dtype=torch.float64)
        saved_k = self.k_local
        nb = 0
        aucs = []
        try:
            self.k_local = 1
            for s in range(0, x.shape[0], B):
                if steps is not None and nb >= steps:
                    break
                xb, yb = x[s:s + B].contiguous(), y[s:s + B].contiguous()
                self.acc.zero_()
                self._launch(xb, yb, xb.shape[0], False)
                tot += self.acc[0].double().cpu()
                nb += 1
                if with_auc:
                    yt = yb.cpu().numpy()
                    if 0 < yt.sum() < len(yt):       # sklearn raises on a single-class batch; Keras would skip it too
                        from sklearn.metrics import roc_auc_score
                        aucs.append(float(roc_auc_score(yt, self._logits[:xb.shape[0]].cpu().numpy())))
        finally:
            self.k_local = saved_k
        cnt = max(float(tot[2]), 1.0)
        if with_auc:
            return float(tot[0]) / cnt, float(tot[1]) / cnt, (sum(aucs) / len(aucs) if aucs else 0.0)
        return float(tot[0]) / cnt, float(tot[1]) / cnt

    def round_trip(self, epochs: int = 5):
        """One full round: fit -> aggregate -> update (secure_fed_model.py:226-233)."""
        stats = self.client_fit(epochs)
        self.aggregate()
        self.client_update()
        return stats
