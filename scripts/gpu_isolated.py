#!/usr/bin/env python
"""Run each test FUNCTION of the given pytest files in its own process (with a timeout), so a kernel
that traps or hangs only takes its own group down. Writes a summary to gpurun_out/isolated.log.

usage: python scripts/gpu_isolated.py tests/test_gpu_conv.py [more files] [--timeout 180]
"""
import re
import subprocess
import sys
import time
from pathlib import Path

args = [a for a in sys.argv[1:] if not a.startswith("--")]
timeout = 180
if "--timeout" in sys.argv:
    timeout = int(sys.argv[sys.argv.index("--timeout") + 1])
    args = [a for a in args if a != str(timeout)]
out = Path("gpurun_out")
out.mkdir(exist_ok=True)
log = open(out / "isolated.log", "a")
ok = True
for f in args:
    funcs = re.findall(r"^def (test_\w+)", Path(f).read_text(), flags=re.M)
    for fn in funcs:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, "-m", "pytest", f, "-q", "-m", "gpu", "-k", fn, "-x",
                                "--no-header", "-p", "no:cacheprovider"],
                               capture_output=True, text=True, timeout=timeout)
            tail = (r.stdout + r.stderr)[-3000:]
            status = "PASS" if r.returncode == 0 else f"FAIL({r.returncode})"
        except subprocess.TimeoutExpired as e:
            tail = ((e.stdout or b"").decode(errors="replace") + (e.stderr or b"").decode(errors="replace"))[-3000:]
            status = "TIMEOUT"
        if status != "PASS":
            ok = False
        line = f"=== {f}::{fn}: {status} ({time.time() - t0:.1f}s)"
        print(line)
        log.write(line + "\n")
        if status != "PASS":
            print(tail)
            log.write(tail + "\n")
        log.flush()
sys.exit(0 if ok else 1)
