#!/usr/bin/env python3
"""Summarise an ncu report offline: key raw metrics + instruction mix + stall samples per opcode.
usage: ncu_summary.py <report.ncu-rep> [cells_per_launch]"""
import collections, csv, io, re, subprocess, sys

rep = sys.argv[1]
cells = float(sys.argv[2]) if len(sys.argv) > 2 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_sector_hit_rate.pct", "sm__cycles_elapsed.avg", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed"]
for w in want:
    if w in hdr:
        i = hdr.index(w)
        print(f"{w:80s} {units[i]:16s} {[r[i] for r in data]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
blocks, cur, h = [], None, None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; blocks.append(cur); continue
    if r and r[0] == "Address":
        h = r; continue
    if cur is not None and r:
        cur["rows"].append(r)
b = blocks[0]
iS, iE, iSamp = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
ops, samp = collections.Counter(), collections.Counter()
tot = 0
for r in b["rows"]:
    e = int(r[iE]); tot += e
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[iS].strip())
    op = m.group(2).split(".")[0] if m else r[iS]
    ops[op] += e; samp[op] += int(r[iSamp])
print(b["name"], "static", len(b["rows"]), "dynamic warp instr", tot)
for op, c in ops.most_common(28):
    per = f"{c * 32 / cells:7.2f}/cell" if cells else ""
    print(f"  {op:10s} {c:12d} {per}  samples {samp[op]}")
if cells:
    print("  thread-instr per cell:", round(tot * 32 / cells, 2))
agg = collections.Counter()
for r in b["rows"]:
    for i, name in enumerate(h):
        if name.startswith("stall_") and "Not Issued" not in name:
            try: agg[name] += int(r[i])
            except ValueError: pass
print("stalls:", agg.most_common(10))
