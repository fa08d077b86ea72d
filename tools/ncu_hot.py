#!/usr/bin/env python
"""Top stall instructions of the first kernel in an .ncu-rep (source page, SASS).
    python tools/ncu_hot.py gpurun_out/prof.ncu-rep [N]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
# first kernel only
start = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
end = next((i for i in range(start + 1, len(rows)) if rows[i] and rows[i][0] == "Kernel Name"), len(rows))
hdr = rows[start]; body = rows[start + 1:end]
ci = {h: i for i, h in enumerate(hdr)}
S = ci["Warp Stall Sampling (All Samples)"]; X = ci["Instructions Executed"]
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[S] or 0) for r in body)
print(f"kernel: {rows[start-1][1][:100]}  instructions: {len(body)}  samples: {tot}")
agg = {h: sum(int(r[ci[h]] or 0) for r in body) for h in stall_cols}
print("stall totals:", ", ".join(f"{h[6:]}={v}" for h, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v))
top = sorted(enumerate(body), key=lambda ir: -int(ir[1][S] or 0))[:N]
for i, r in sorted(top):
    why = sorted(((int(r[ci[h]] or 0), h[6:]) for h in stall_cols), reverse=True)[:2]
    print(f"{i:5d} {int(r[S]):7d} {100*int(r[S])/tot:5.1f}%  x{r[X]:>9s}  {r[1].strip()[:70]:70s} {why[0][1]}={why[0][0]} {why[1][1]}={why[1][0]}")
