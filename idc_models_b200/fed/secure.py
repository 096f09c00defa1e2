"""Secure-aggregation federated averaging with the ``Client`` / ``Server`` structure of the reference
(/root/reference/secure_fed_model.py:101-168,212-236).

The reference protects the first ``int(n_tensors * percent)`` weight tensors by element-wise Paillier
encryption and lets the server average ciphertexts. Here the same tensors are protected by **pairwise
additive masks in a 2^32 fixed-point ring** (Bonawitz et al. style): client k sends

    y_k = Q(w_k) + sum_{j>k} PRG(k, j) - sum_{j<k} PRG(j, k)        (mod 2^32)

so the server only ever sees masked vectors, while ``sum_k y_k = sum_k Q(w_k)`` holds EXACTLY (integer
arithmetic) and the unweighted mean the reference's ``Server.aggregate`` computes is recovered by the
clients (``dec_model``: dequantise the ring sum, divide by the client count). The unprotected tensors are
averaged in float like the reference does. ``percent`` keeps its CLI meaning.

Two execution paths share the arithmetic (same Philox4x32-10 header on host and device):
* host objects below (``Client.enc_model`` -> ``Server.aggregate`` -> ``Client.dec_model``), masks from the
  C++ restatement in ``libidcb200_host.so`` — the reference-shaped path and the oracle;
* :func:`fused_secure_mean` — ONE kernel for all clients resident on all GPUs
  (csrc/comm.cu:secagg_kernel): masks are generated in registers and never exist in memory, partial ring
  sums cross NVLink once (``multimem.ld_reduce.add.u32`` when the fabric offers it).

Full participation is assumed (as in the reference: every client every round); a missing contribution
would leave un-cancelled masks, so ``Server.aggregate`` refuses a short client list.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from ..ops import _lib
from ..utils.timer import Timer

FRAC_BITS = 16


def _host():
    lib = _lib.host_lib()
    return lib


def mask_vectors(w: np.ndarray, seed: int, frac_bits: int = FRAC_BITS) -> np.ndarray:
    """[K, n] float32 -> [K, n] uint32 masked ring vectors (n padded to a multiple of 4 internally)."""
    w = np.ascontiguousarray(w, dtype=np.float32)
    k, n = w.shape
    n4 = -(-n // 4) * 4
    wp = np.zeros((k, n4), dtype=np.float32)
    wp[:, :n] = w
    y = np.zeros((k, n4), dtype=np.uint32)
    rc = _host().idcbh_secagg_mask(wp.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p), C.c_int(k),
                                   C.c_longlong(n4), C.c_int(frac_bits), C.c_ulonglong(seed))
    if rc != 0:
        raise RuntimeError(_host().idcbh_last_error().decode())
    return y[:, :n]


def ring_sum(ys: list) -> np.ndarray:
    """Server side: sum of masked vectors modulo 2^32."""
    s = np.zeros_like(ys[0], dtype=np.uint32)
    for y in ys:
        s = (s + y.astype(np.uint32)).astype(np.uint32)
    return s


def dequantise_mean(s: np.ndarray, n_clients: int, frac_bits: int = FRAC_BITS) -> np.ndarray:
    return (s.astype(np.uint32).view(np.int32).astype(np.float64) / ((1 << frac_bits) * n_clients)).astype(np.float32)


class MaskedTensor:
    """A protected weight tensor in flight: uint32 ring values + the shape to restore."""

    def __init__(self, values: np.ndarray, shape, n_clients: int = 1, is_sum: bool = False):
        self.values, self.shape, self.n_clients, self.is_sum = values, tuple(shape), n_clients, is_sum


class Client:
    """secure_fed_model.py:101-154. ``model_fn`` builds the client's model; ``data`` is its shard."""

    def __init__(self, data, num: int, p: float, *, model_fn, n_clients: int, batch_size: int = 32,
                 train_size: int | None = None, test_size: int | None = None, session_seed: int = 0,
                 frac_bits: int = FRAC_BITS, prepare=None, protect: str = "mask", keypair=None):
        """``protect='mask'`` (default): pairwise masks in the fixed-point ring. ``protect='paillier'``: the
        reference's own mechanism — element-wise Paillier ciphertexts under one shared ``keypair``
        (secure_fed_model.py:79,109-121) — kept as the parity path; it is orders of magnitude slower."""
        if protect not in ("mask", "paillier"):
            raise ValueError("protect must be 'mask' or 'paillier'")
        if protect == "paillier" and keypair is None:
            raise ValueError("protect='paillier' needs the shared (public_key, private_key) pair")
        self.protect, self.keypair = protect, keypair
        from ..data.dataset import prepare_for_training
        self.model = model_fn()
        prep = prepare or (lambda ds: prepare_for_training(ds, batch_size))
        if train_size is not None:
            self.train = prep(data.take(train_size))
            self.validation = prep(data.skip(train_size).take(test_size)) if test_size else None
        else:
            self.train, self.validation = prep(data), None
        self.id = num
        self.percent = p
        self.n_clients = n_clients
        self.session_seed = session_seed
        self.frac_bits = frac_bits
        self.round = 0

    # ---- protection of the first int(n_tensors * percent) tensors ------------------------------------
    def n_protected(self, n_tensors: int) -> int:
        return int(n_tensors * self.percent)

    def enc_model(self, weights: list) -> list:
        """Quantise + mask the protected prefix. Needs only this client's id, the client count and the
        per-round session seed (pairwise seeds are derived from it inside the generator)."""
        out = list(weights)
        npro = self.n_protected(len(weights))
        if npro == 0:
            return out
        if self.protect == "paillier":
            enc = np.vectorize(self.keypair[0].encrypt, otypes=[object])     # secure_fed_model.py:109-110,117
            for i in range(npro):
                out[i] = enc(np.asarray(weights[i], np.float64))
            return out
        flat = np.concatenate([np.asarray(w, np.float32).reshape(-1) for w in weights[:npro]])
        # the mask routine works on the full client matrix; build this client's row only
        mat = np.zeros((self.n_clients, flat.size), np.float32)
        mat[self.id] = flat
        y = mask_vectors(mat, self._round_seed(), self.frac_bits)[self.id]
        # rows of absent clients quantise 0 and do not affect this client's own masked row
        off = 0
        for i in range(npro):
            n = int(np.prod(np.shape(weights[i]))) if np.ndim(weights[i]) else 1
            out[i] = MaskedTensor(y[off:off + n].copy(), np.shape(weights[i]))
            off += n
        return out

    def dec_model(self, weights: list) -> list:
        """Recover the mean of the protected tensors from the server's ring sums."""
        out = []
        for w in weights:
            if isinstance(w, np.ndarray) and w.dtype == object:
                dec = np.vectorize(self.keypair[1].decrypt, otypes=[np.float64])   # secure_fed_model.py:112-113,121
                out.append(dec(w).astype(np.float32))
            elif isinstance(w, MaskedTensor):
                if not w.is_sum:
                    raise ValueError("dec_model expects the server's aggregated ring sums")
                out.append(dequantise_mean(w.values, w.n_clients, self.frac_bits).reshape(w.shape))
            else:
                out.append(np.asarray(w, dtype=np.float32))
        return out

    def _round_seed(self) -> int:
        return (self.session_seed * 1000003 + self.round) & 0xFFFFFFFFFFFFFFFF

    # ---- the reference's client API ----------------------------------------------------------------------
    def client_fit(self, epochs: int = 10, verbose: int = 1):
        with Timer("Training for client " + str(self.id)):
            history = self.model.fit(self.train, epochs=epochs, validation_data=self.validation, verbose=verbose)
        self.round += 1
        if self.percent > 0:
            with Timer("Encryption for client " + str(self.id)):
                weights = self.enc_model(self.model.get_weights())
                return weights, history
        return self.model.get_weights(), history

    def client_update(self, weights: list) -> None:
        if self.percent > 0:
            with Timer("Decryption for client " + str(self.id)):
                new_weights = self.dec_model(weights)
            self.model.set_weights(new_weights)
        else:
            self.model.set_weights([np.asarray(w, np.float32) for w in weights])

    def evaluate(self, test_batches, steps: int = 20):
        res = self.model.evaluate(test_batches, steps=steps)
        return tuple(res)


class Server:
    """secure_fed_model.py:156-168: unweighted aggregation; never sees a protected tensor in the clear."""

    def __init__(self, n_clients: int, model_fn=None):
        self.n_clients = n_clients
        self.model = model_fn() if model_fn is not None else None

    def aggregate(self, client_weights: list) -> list:
        if len(client_weights) != self.n_clients:
            raise RuntimeError(f"secure aggregation needs all {self.n_clients} clients (got {len(client_weights)}): "
                               "pairwise masks of a missing client would not cancel")
        if self.n_clients == 1:
            w = client_weights[0]
            return [MaskedTensor(t.values, t.shape, 1, True) if isinstance(t, MaskedTensor) else t for t in w]
        ave = []
        for tensors in zip(*client_weights):
            if isinstance(tensors[0], MaskedTensor):
                ave.append(MaskedTensor(ring_sum([t.values for t in tensors]), tensors[0].shape, self.n_clients, True))
            elif isinstance(tensors[0], np.ndarray) and tensors[0].dtype == object:
                # ciphertext arrays: the reference's np.array(w).mean(axis=0) works because EncryptedNumber has
                # + and scalar / (secure_fed_model.py:160-168); the server holds no key
                acc = tensors[0]
                for t in tensors[1:]:
                    acc = acc + t
                ave.append(acc / float(self.n_clients))
            else:
                ave.append(np.mean(np.stack([np.asarray(t, np.float64) for t in tensors]), axis=0).astype(np.float32))
        return ave


# -------------------------------------------------------------------------------------------------------
# Fused path: all clients of all GPUs in one kernel
# -------------------------------------------------------------------------------------------------------
def fused_secure_mean(alloc, clients_buf, ybuf, out_buf, k_local: int, stride: int, offset: int, n: int,
                      session_seed: int, frac_bits: int = FRAC_BITS, comm_world=None) -> None:
    """out[offset:offset+n] (every rank) = mean over all world*k_local clients of the protected range,
    computed through masked ring sums (csrc/comm.cu:secagg_kernel). Buffers are SymmBuffers."""
    lib = _lib.cuda_lib()
    cw = comm_world if comm_world is not None else alloc.comm_world()
    _lib.check(lib.idcb_secagg_masked_sum(C.byref(cw), clients_buf.ptrs(), ybuf.ptrs(), ybuf.mc(), out_buf.ptrs(),
                                          out_buf.mc(), C.c_int(k_local), C.c_longlong(stride), C.c_longlong(offset),
                                          C.c_longlong(n), C.c_int(frac_bits), C.c_ulonglong(session_seed),
                                          _lib.stream()))
    _lib.count_launch()
