#!/usr/bin/env python
"""Regenerate profiles/sass_mnemonic_counts.txt: per-kernel counts of the SASS mnemonics that prove which hardware
paths a kernel uses (tcgen05 / TMEM / TMA / clusters / multimem / griddepcontrol), from ``cuobjdump -sass`` of the
built library. Runs without a GPU.

    python scripts/sass_counts.py [path/to/libidcb200.so] > profiles/sass_mnemonic_counts.txt
    python scripts/sass_counts.py --comm [path/to/libidcb200.so] > profiles/sass_comm_kernels.txt

``--comm``: the SASS of the communication kernels (csrc/comm.cu) trimmed to the instructions that touch peer /
multicast memory or order it (system-scope loads and stores, multimem reductions, atomics of the barrier, fences).
"""
import re
import subprocess
import sys
from collections import OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
KEYS = ["UTCHMMA", "UTMALDG.2D", "UTMALDG.4D.IM2COL", "UTMAPF", "LDTM", "UTCBAR", "UTCATOMSWS", "LDGSTS", "SYNCS",
        "ACQBULK", "PREEXIT", "UCGABAR", "MAPA", "LDGMC", "REDG", "MEMBAR", "FENCE.VIEW.ASYNC", "HMMA"]
HEADER = """# SASS mnemonic counts per kernel of idc_models_b200/_native/libidcb200.so (cuobjdump -sass, sm_100a)
# UTCHMMA = tcgen05.mma, UTMALDG.2D = TMA tiled load, UTMALDG.4D.IM2COL = TMA im2col load, UTMAPF = tensormap prefetch, \
LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UTCATOMSWS = tcgen05.alloc/dealloc, LDGSTS = cp.async, SYNCS = mbarrier, \
ACQBULK = griddepcontrol.wait, PREEXIT = griddepcontrol.launch_dependents, UCGABAR = barrier.cluster, MAPA = mapa (DSMEM \
address), LDGMC = multimem.ld_reduce, REDG = red.global, MEMBAR = membar/fence, FENCE.VIEW.ASYNC = fence.proxy.async, \
HMMA = legacy mma.sync
# regenerate with: python scripts/sass_counts.py > profiles/sass_mnemonic_counts.txt
"""


def demangle(names: list) -> list:
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout
    return out.splitlines()


def short(sig: str) -> str:
    """'void idcb::k<32, true, 0>(Args)' -> 'idcb::k<32, true, 0>'"""
    sig = re.sub(r"^void\s+", "", sig)
    depth, cut = 0, len(sig)
    for i, ch in enumerate(sig):
        if ch == "<":
            depth += 1
        elif ch == ">":
            depth -= 1
        elif ch == "(" and depth == 0:
            cut = i
            break
    return sig[:cut]


COMM_KERNELS = ("secagg_kernel", "fedavg_reduce_kernel", "fedavg_round_kernel", "stat_mean_kernel", "allreduce_optim_kernel")
COMM_HEADER = """# SASS evidence for the fused communication kernels (cuobjdump -sass libidcb200.so, sm_100a), trimmed to memory / \
barrier instructions
# LDGMC = multimem.ld_reduce (NVLS in-switch reduction); STG/LDG ...STRONG.SYS = system-scope peer stores/loads \
(multimem.st, st/ld.relaxed|release|acquire.sys)
# regenerate with: python scripts/sass_counts.py --comm > profiles/sass_comm_kernels.txt
"""


def comm_listing(sass: str) -> int:
    keep = re.compile(r"LDGMC|STRONG\.SYS|MEMBAR|ATOMG|ERRBAR")
    sys.stdout.write(COMM_HEADER)
    on = False
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            nice = demangle([m.group(1)])[0]
            on = any(k in nice for k in COMM_KERNELS)
            if on:
                sys.stdout.write("\t\tFunction : " + re.sub(r"^void\s+", "", nice) + "\n")
            continue
        if on and re.match(r"^\s*/\*[0-9a-f]+\*/", line) and keep.search(line):
            sys.stdout.write(line + "\n")
    return 0


def main() -> int:
    args = [a for a in sys.argv[1:] if a != "--comm"]
    lib = Path(args[0]) if args else ROOT / "idc_models_b200" / "_native" / "libidcb200.so"
    sass = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True, check=True).stdout
    if "--comm" in sys.argv[1:]:
        return comm_listing(sass)
    counts: "OrderedDict[str, dict]" = OrderedDict()
    cur = None
    op = re.compile(r"^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P[0-9T]+\s+)?([A-Z][A-Z0-9_.]*)")
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = counts.setdefault(m.group(1), {k: 0 for k in KEYS})
            continue
        if cur is None:
            continue
        m = op.match(line)
        if not m:
            continue
        mn = m.group(1)
        for k in KEYS:
            if mn == k or mn.startswith(k + ".") or mn.startswith(k + "_"):
                # whole-mnemonic prefixes only: UTMALDG.4D.IM2COL is not also a UTMALDG.2D, HMMA does not match UTCHMMA
                cur[k] += 1
    names = demangle(list(counts))
    sys.stdout.write(HEADER + "\n")
    for raw, nice in zip(counts, names):
        c = counts[raw]
        cells = " ".join(f"{k}:{c[k]}" for k in KEYS if c[k])
        sys.stdout.write(f"{short(nice):<96}{cells}\n")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
