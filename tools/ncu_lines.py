"""Per-source-line instruction counts and lane utilisation from an .ncu-rep: python tools/ncu_lines.py <rep> [top]"""
import collections
import csv
import io
import subprocess
import sys

rep, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 60
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Line No"]
lines = []
for k, h in enumerate(hi):
    f = rows[h - 2][1].split("/")[-1]
    cols = {n: i for i, n in enumerate(rows[h])}
    end = hi[k + 1] - 2 if k + 1 < len(hi) else len(rows)
    for r in rows[h + 1:end]:
        if not r[0].isdigit():
            continue
        inst = int(r[cols["Instructions Executed"]] or 0)
        tinst = int(r[cols["Thread Instructions Executed"]] or 0)
        smp = int(r[cols["# Samples"]] or 0)
        if inst:
            lines.append((inst, tinst, smp, f, int(r[0]), r[1].strip()[:100]))
ti = sum(l[0] for l in lines)
ts = sum(l[2] for l in lines)
print(f"total warp-inst {ti}, avg lanes {sum(l[1] for l in lines) / ti:.1f}")
byf = collections.defaultdict(lambda: [0, 0, 0])
for inst, tinst, smp, f, ln, txt in lines:
    b = byf[(f, ln // 25 * 25)]
    b[0] += inst; b[1] += tinst; b[2] += smp
print("-- by 25-line block")
for (f, ln), (inst, tinst, smp) in sorted(byf.items(), key=lambda kv: -kv[1][0])[:40]:
    print(f"{100 * inst / ti:5.1f}% inst {100 * smp / ts:5.1f}% smp lanes {tinst / inst:4.1f}  {f}:{ln}")
print("-- by line")
for inst, tinst, smp, f, ln, txt in sorted(lines, reverse=True)[:top]:
    print(f"{100 * inst / ti:5.1f}% inst {100 * smp / ts:5.1f}% smp lanes {tinst / inst:4.1f}  {f}:{ln}: {txt}")
