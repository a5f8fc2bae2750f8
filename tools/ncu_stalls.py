"""Stall-reason summary of one kernel from an `ncu --set full --import-source on` report (run where ncu is installed):
    python tools/ncu_stalls.py report.ncu-rep [top_n]
Prints the share of every warp-stall reason over all samples and the instructions with the most samples."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr, data = rows[start], rows[start + 1:]
print(rows[start - 1][1][:120] if start else "")
first = hdr.index("stall_barrier")
names = hdr[first:first + 17]
isamp, iexec = hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
tot = {n: 0 for n in names}
total = 0
good = []
for r in data:
    if len(r) <= first + 16 or not r[isamp].isdigit():
        continue
    good.append(r)
    total += int(r[isamp])
    for i, n in enumerate(names):
        tot[n] += int(r[first + i])
print("samples", total, " instructions", len(good), " executed warp-instructions", sum(int(r[iexec]) for r in good))
for n, v in sorted(tot.items(), key=lambda kv: -kv[1])[:10]:
    print(f"  {n:26s}{v:8d} {v / max(total, 1):6.1%}")
pos = {id(r): i for i, r in enumerate(good)}
for r in sorted(good, key=lambda r: -int(r[isamp]))[:top_n]:
    st = sorted(((names[i], int(r[first + i])) for i in range(17) if int(r[first + i]) > 0), key=lambda kv: -kv[1])[:2]
    print(str(pos[id(r)]).rjust(5), r[isamp].rjust(6), r[iexec].rjust(9), r[1].strip()[:60].ljust(60), st)
