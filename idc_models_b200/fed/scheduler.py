"""Client -> rank placement (csrc/host/scheduler.cpp): longest-processing-time-first on the clients' example
counts, so that a federated round — which lasts as long as its slowest GPU — is balanced when clients differ in
size (the reference's non-IID split keeps sizes equal, fed_model.py:178-180, but ``ClientData`` allows any).
Every rank computes the same plan from the same counts."""
from __future__ import annotations

import ctypes as C

from ..ops import _lib


def assign_clients(costs, n_ranks: int) -> list:
    """-> rank index for every client. ``costs``: per-client work (e.g. number of examples)."""
    n = len(costs)
    if n_ranks <= 1 or n == 0:
        return [0] * n
    lib = _lib.host_lib()
    arr = (C.c_longlong * n)(*[int(c) for c in costs])
    out = (C.c_int * n)()
    load = (C.c_longlong * n_ranks)()
    if lib.idcbh_schedule_clients(C.c_int(n), arr, C.c_int(n_ranks), out, load) != 0:
        raise RuntimeError(lib.idcbh_last_error().decode())
    return list(out)


def clients_of_rank(costs, n_ranks: int, rank: int) -> list:
    plan = assign_clients(costs, n_ranks)
    return [i for i, r in enumerate(plan) if r == rank]
