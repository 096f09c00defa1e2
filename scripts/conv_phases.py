#!/usr/bin/env python
"""Per-phase clock stamps of one conv CTA (profiling aid) + graph-replayed GPU-only timings."""
import ctypes, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from idc_models_b200.ops import conv as C, _lib

dev = "cuda"
lib = _lib.cuda_lib()
dbg = torch.zeros(16, dtype=torch.int64, device=dev)
names = ["entry", "setup done", "A0 loaded", "producer done", "mma first full", "mma issued", "acc ready", "pre final sync", "post final sync", "dealloc done"]
for (b, h, cin, cout, k) in [(128, 1, 992, 128, 1), (128, 1, 128, 32, 3), (128, 13, 128, 128, 1)]:
    g = C.geom(h, h, cin, cout, k)
    x = torch.randn(b, h, h, cin, device=dev).to(torch.bfloat16)
    w = (torch.randn(cout, k, k, cin, device=dev) * 0.05).to(torch.bfloat16)
    y = torch.zeros(b, h, h, cout, device=dev, dtype=torch.bfloat16)
    for rep in range(3):
        lib.idcb_conv_set_debug(ctypes.c_void_p(dbg.data_ptr()))
        C.conv_fprop_raw(g, x, w, y)
        torch.cuda.synchronize()
        lib.idcb_conv_set_debug(ctypes.c_void_p(0))
        t = dbg.tolist()
        print(f"fprop {b}x{h}x{h} {cin}->{cout} k{k} rep{rep}: " + ", ".join(f"{names[i]}={t[i]-t[0]}" for i in range(1, 10)))
    # GPU-only time via graph replay of 20 launches
    gph = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        C.conv_fprop_raw(g, x, w, y)
        torch.cuda.synchronize()
        with torch.cuda.graph(gph, stream=s):
            for _ in range(20):
                C.conv_fprop_raw(g, x, w, y)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    gph.replay(); torch.cuda.synchronize()
    e0.record(); gph.replay(); e1.record(); torch.cuda.synchronize()
    print(f"   graph replay: {e0.elapsed_time(e1) * 1000 / 20:.2f} us per launch")
# trivial kernel for reference
from idc_models_b200.ops import kernels as K
a = torch.zeros(64, device=dev); bb = torch.zeros(64, device=dev)
gph = torch.cuda.CUDAGraph()
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    K.bn_fold(a, a, a, a + 1, 1e-3, bb, bb.clone())
    torch.cuda.synchronize()
    with torch.cuda.graph(gph, stream=s):
        for _ in range(20):
            K.bn_fold(a, a, a, a + 1, 1e-3, bb, bb)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
gph.replay(); torch.cuda.synchronize()
e0.record(); gph.replay(); e1.record(); torch.cuda.synchronize()
print(f"bn_fold graph replay: {e0.elapsed_time(e1) * 1000 / 20:.2f} us per launch")
