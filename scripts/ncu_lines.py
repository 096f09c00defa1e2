#!/usr/bin/env python
"""Per-source-line stall samples of every kernel launch in an .ncu-rep (needs -lineinfo and --import-source on).
usage: python scripts/ncu_lines.py gpurun_out/prof.ncu-rep [top_n]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"],
                     capture_output=True, text=True).stdout
launches = []          # one Counter per kernel launch (a launch's sections start again at common.cuh)
cur = None; fpath = None; hdr = None; seen_files = set(); fn = None
for row in csv.reader(io.StringIO(raw)):
    if not row: continue
    if row[0] == "File Path":
        f = row[1].split("/")[-1]
        if cur is None or f in seen_files:
            cur = {"samples": collections.Counter(), "instr": collections.Counter(), "stalls": collections.Counter(), "fn": None}
            launches.append(cur); seen_files = set()
        seen_files.add(f); fpath = f; continue
    if row[0] == "Function Name":
        cur["fn"] = row[1]; continue
    if row[0] == "Line No": hdr = row; continue
    if hdr is None or cur is None or not row[0].isdigit(): continue
    d = dict(zip(hdr, row))
    try:
        s = int(d.get("# Samples") or 0); n = int(d.get("Instructions Executed") or 0)
    except ValueError:
        continue
    key = (fpath, int(row[0]), row[1][:100])
    cur["samples"][key] += s; cur["instr"][key] += n
    for k2 in hdr:
        if k2.startswith("stall_") and "Not Issued" not in k2:
            try: cur["stalls"][(key, k2)] += int(d.get(k2) or 0)
            except ValueError: pass
for i, L in enumerate(launches):
    tot = sum(L["samples"].values())
    print(f"\n=== launch {i}: {L['fn'][:100]}  total samples {tot}")
    for key, s in L["samples"].most_common(top):
        if s == 0: break
        st = sorted(((v, k2) for (k1, k2), v in L["stalls"].items() if k1 == key and v), reverse=True)[:2]
        sts = " ".join(f"{k2[6:]}={v}" for v, k2 in st)
        print(f"{100*s/max(tot,1):5.1f}% {s:5d} ins={L['instr'][key]:7d} {key[0]}:{key[1]}: {key[2].strip()[:80]}  [{sts}]")
