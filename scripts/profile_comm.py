#!/usr/bin/env python
"""Single-GPU driver for `ncu` captures of the three fused communication kernels: N ranks are emulated by
one process (buffers of all ranks on one device, barriers off), which exercises exactly the device code
that runs across NVLink (peer pointers are simply local here)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import ctypes as C
import numpy as np
import torch
from idc_models_b200.fed import secure
from idc_models_b200.ops import _lib
from idc_models_b200.parallel.fused_comm import allreduce_apply_emulated
from idc_models_b200.parallel.symm import EmulatedWorld

world = 8
emu = EmulatedWorld(world)
n = 6955008      # DenseNet-121 trainable floats
bufs = {"grad": emu.alloc(n, torch.float32), "master": emu.alloc(n, torch.float32), "shadow": emu.alloc(n, torch.bfloat16)}
for r in range(world):
    bufs["grad"][r].tensor.normal_()
states = [(torch.zeros(n, device="cuda"), None) for _ in range(world)]
for _ in range(2):
    allreduce_apply_emulated(emu, bufs, states, n, "rmsprop", 1e-4, 0.9, 0.0, 1e-7, 1.0)
torch.cuda.synchronize()
# FedAvg: 8 GPUs x 1 client, MobileNetV2-sized model
lib = _lib.cuda_lib()
nm = 2225408
clients = emu.alloc(nm, torch.float32); weights = emu.alloc(4, torch.float32)
glob = emu.alloc(nm, torch.float32); shadow = emu.alloc(nm, torch.bfloat16)
for r in range(world):
    clients[r].tensor.normal_(); weights[r].tensor[0] = 3000.0
for r in range(world):
    cw = emu.comm_world(r)
    _lib.check(lib.idcb_fedavg_reduce(C.byref(cw), clients[r].ptrs(), weights[r].ptrs(), glob[r].ptrs(), shadow[r].ptrs(),
                                      None, None, C.c_int(1), C.c_longlong(nm), C.c_longlong(0), C.c_longlong(nm),
                                      C.c_float(1.0), _lib.stream()))
torch.cuda.synchronize()
# secure aggregation: 64 clients over 8 GPUs, MobileNetV2-sized protected range
k_local = 8
cl = emu.alloc(k_local * nm, torch.float32); yb = emu.alloc(nm, torch.int32); out = emu.alloc(nm, torch.float32)
for r in range(world):
    cl[r].tensor.normal_()
for r in range(2):
    secure.fused_secure_mean(None, cl[r], yb[r], out[r], k_local, nm, 0, nm, 1234, comm_world=emu.comm_world(r))
torch.cuda.synchronize()
print("ok")
