#!/usr/bin/env python
"""Turn an .ncu-rep into a small markdown table of the metrics the B200 profiling recipe asks for.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep profiles/out.md "title" """
import csv
import io
import subprocess
import sys

rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = [("Kernel Name", "kernel"), ("Grid Size", "grid"), ("Block Size", "block"),
        ("gpu__time_duration.sum", "duration"), ("launch__registers_per_thread", "regs"),
        ("launch__shared_mem_per_block_dynamic", "dyn smem"), ("launch__waves_per_multiprocessor", "waves/SM"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
        ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
        ("lts__t_bytes.sum", "L2 bytes"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_scoreboard"),
        ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier"),
        ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait"),
        ("smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "stall no_instruction"),
        ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short_scoreboard"),
        ("smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "stall membar")]
idx = [(hdr.index(k), lab) for k, lab in want if k in hdr]
lines = [f"# {title}", "", f"source: `{rep}` (`ncu --set full --clock-control none --import-source on`, one B200)", ""]
for r in rows[2:]:
    lines.append("| metric | value |")
    lines.append("|---|---|")
    for i, lab in idx:
        v = r[i]
        if lab == "kernel":
            v = "`" + v[:110] + "`"
        lines.append(f"| {lab} | {v} {units[i] if lab not in ('kernel', 'grid', 'block') else ''} |")
    lines.append("")
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines[:60]))
