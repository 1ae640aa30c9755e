"""Text summary of an .ncu-rep for profiles/: python tools/ncu_summary.py <rep> > profiles/<name>.txt
Key section metrics per kernel, the stall mix and the hottest source lines (needs -lineinfo + --import-source on)."""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
KEYS = ("Duration", "Executed Ipc Active", "Executed Ipc Elapsed", "Issue Slots Busy", "Active Warps Per Scheduler", "Eligible Warps Per Scheduler",
        "Registers Per Thread", "Dynamic Shared Memory Per Block", "Theoretical Active Warps per SM", "Achieved Active Warps Per SM",
        "Avg. Active Threads Per Warp", "Executed Instructions ", "DRAM Throughput", "Memory Throughput", "L1/TEX Hit Rate", "Waves Per SM",
        "Grid Size", "Block Size")
det = subprocess.run(["ncu", "-i", rep, "--page", "details"], capture_output=True, text=True).stdout
for line in det.splitlines():
    s = line.strip()
    if s.endswith("CC 10.0") or any(s.startswith(k) for k in KEYS):
        print(line.rstrip())
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
if rows:
    hdr = rows[0]
    for name in ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "smsp__inst_executed.sum",
                 "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum", "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum",
                 "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum",
                 "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_xu.sum",
                 "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
                 "sm__cycles_active.avg", "sm__cycles_elapsed.avg.per_second"):
        if name in hdr:
            i = hdr.index(name)
            print(f"raw {name}: " + ", ".join(r[i] for r in rows[2:]) + f" [{rows[1][i]}]")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Line No"]
per_kernel = collections.defaultdict(lambda: {"stall": collections.Counter(), "lines": []})
for k, h in enumerate(hi):
    fn = rows[h - 1][1][:60]
    f = rows[h - 2][1].split("/")[-1]
    hdr = rows[h]
    cols = {n: i for i, n in enumerate(hdr)}
    end = hi[k + 1] - 2 if k + 1 < len(hi) else len(rows)
    for r in rows[h + 1:end]:
        if not r[0].isdigit():
            continue
        try:
            inst = int(r[cols["Instructions Executed"]] or 0)
            smp = int(r[cols["# Samples"]] or 0)
        except ValueError:
            continue
        per_kernel[fn]["lines"].append((smp, inst, f, int(r[0]), r[1].strip()[:90]))
        for n, i in cols.items():
            if n.startswith("stall_") and "Not Issued" not in n:
                try:
                    per_kernel[fn]["stall"][n] += int(r[i] or 0)
                except ValueError:
                    pass
for fn, d in per_kernel.items():
    tot = sum(d["stall"].values()) or 1
    ts = sum(l[0] for l in d["lines"]) or 1
    ti = sum(l[1] for l in d["lines"]) or 1
    print(f"\n== {fn}: warp-instructions {ti}, stall samples {ts}")
    print("   stalls: " + ", ".join(f"{k[6:]} {100 * v / tot:.1f}%" for k, v in d["stall"].most_common(8)))
    byfile = collections.Counter()
    for smp, inst, f, ln, txt in d["lines"]:
        byfile[f] += inst
    print("   instructions by file: " + ", ".join(f"{f} {100 * v / ti:.1f}%" for f, v in byfile.most_common(8)))
    for smp, inst, f, ln, txt in sorted(d["lines"], reverse=True)[:12]:
        print(f"   {100 * smp / ts:5.1f}% samples {100 * inst / ti:5.1f}% inst  {f}:{ln}: {txt}")
