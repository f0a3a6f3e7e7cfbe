#!/usr/bin/env python3
"""Summarise an .ncu-rep (read here, no GPU needed): key raw metrics + top stall lines.

usage: scripts/ncu_summary.py gpurun_out/x.ncu-rep profiles/x.txt ["note"]"""
import csv
import io
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
note = sys.argv[3] if len(sys.argv) > 3 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
WANT = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size",
        "launch__cluster", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "smsp__sass_inst_executed_op_tmem_ldt.sum"]
with open(out, "w") as f:
    f.write(f"ncu --set full --clock-control none, report {rep.split('/')[-1]}\n{note}\n\n")
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        f.write(f"kernel: {name}\n")
        for h, u, v in zip(hdr, units, r):
            if any(h == w or h.endswith(w) for w in WANT):
                f.write(f"  {h} [{u}] = {v}\n")
        stalls = sorted(((float(v), h) for h, v in zip(hdr, r)
                         if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio") and v),
                        reverse=True)[:6]
        f.write("  top stall reasons (warps per issue-active cycle):\n")
        for v, h in stalls:
            f.write(f"    {h.split('stalled_')[1].split('_per_issue')[0]:28s} {v:.3f}\n")
        f.write("\n")
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    srows = list(csv.reader(io.StringIO(src)))
    if len(srows) > 2 and "# Samples" in srows[1]:
        h2 = srows[1]
        si, so = h2.index("# Samples"), h2.index("Source")
        data = []
        for r in srows[2:]:
            try:
                data.append((int(r[si]), r[so].strip()))
            except (ValueError, IndexError):
                pass
        tot = sum(d[0] for d in data) or 1
        f.write("hottest SASS lines by warp-state samples:\n")
        for n, line in sorted(data, reverse=True)[:14]:
            f.write(f"  {100.0 * n / tot:5.1f}%  {line[:100]}\n")
print(open(out).read())
