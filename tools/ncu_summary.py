"""Summarise an .ncu-rep (first kernel): key raw metrics, opcode mix, stall samples per opcode, hottest SASS lines.
usage: python tools/ncu_summary.py report.ncu-rep [n_hot]"""
import csv, io, re, subprocess, sys
from collections import Counter

rep = sys.argv[1]
nhot = int(sys.argv[2]) if len(sys.argv) > 2 else 20
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = [r"^gpu__time_duration.sum$", r"^sm__cycles_elapsed.avg$", r"^sm__inst_executed_pipe_(alu|fma|xu|lsu|fp64|uniform)\.avg\.pct_of_peak_sustained_active$",
        r"^sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_(active|elapsed)$", r"^sm__pipe_tensor_subpipe_.mma_cycles_active.avg.pct_of_peak_sustained_active$",
        r"^smsp__issue_active.avg.pct_of_peak_sustained_active$", r"^sm__warps_active.avg.pct_of_peak_sustained_active$",
        r"^smsp__inst_executed.sum$", r"^dram__bytes_(read|write).sum$", r"^launch__(registers_per_thread|grid_size|block_size|occupancy_limit_[a-z_]+)$",
        r"^sm__throughput.avg.pct_of_peak_sustained_elapsed$", r"^l1tex__data_pipe_lsu_wavefronts_mem_shared.sum$",
        r"^l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum$", r"^smsp__pcsamp_warps_issue_stalled_[a-z_]+(?<!not_issued)$",
        r"^lts__t_bytes.sum$", r"^sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active$"]
print("kernel:", vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?")
for h, u, v in zip(hdr, units, vals):
    if any(re.search(w, h) for w in want):
        print(f"  {h} [{u}] = {v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]
isrc, ins, isamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
# multi-kernel reports repeat the header row per kernel: keep numeric rows only
data = [r for r in rows[h + 1:] if len(r) == len(hdr) and r[ins].strip().isdigit() and r[isamp].strip().isdigit()]
tot = sum(int(r[ins]) for r in data) or 1
tots = sum(int(r[isamp]) for r in data) or 1
c, cs = Counter(), Counter()
for r in data:
    t = r[isrc].split()
    op = t[1] if t[0].startswith("@") else t[0]
    c[op] += int(r[ins]); cs[op] += int(r[isamp])
print(f"warp instructions {tot}, samples {tots}")
for op, n in c.most_common(18):
    print(f"  {op:30s} {100*n/tot:5.1f}% inst  {100*cs[op]/tots:5.1f}% samples")
print("hottest SASS lines (index, samples, executed):")
for i, r in sorted(enumerate(data), key=lambda x: -int(x[1][isamp]))[:nhot]:
    print(f"  {i:5d} {int(r[isamp]):6d} {int(r[ins]):9d}  {r[isrc].strip()[:90]}")
