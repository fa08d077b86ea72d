#!/usr/bin/env python
"""Top warp-stall locations of one kernel from an ncu report captured with --import-source on.

    ncu -i REPORT.ncu-rep --page source --csv > src.csv ; python tools/ncu_stalls.py src.csv [N]

Prints, for the N instructions with the most warp-state samples: share of all samples, SASS address, the
instruction, its two leading stall reasons, and for mbarrier try-waits the shared-memory offset (which barrier).
"""
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    print("#", rows[0][1] if len(rows[0]) > 1 else "")
    h = rows[1]
    si, src, ad = h.index("# Samples"), h.index("Source"), h.index("Address")
    stall = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    body = [r for r in rows[2:] if r[si].isdigit()]
    total = sum(int(r[si]) for r in body)
    print(f"# {total} samples over {len(body)} SASS instructions")
    order = sorted(range(len(body)), key=lambda k: -int(body[k][si]))[:top]
    for k in order:
        r = body[k]
        n = int(r[si])
        why = sorted(((int(r[i] or 0), h[i]) for i in stall), reverse=True)[:2]
        prev = body[k - 1][src] if k > 0 else ""
        note = ("   <- after " + prev.split("P0,")[-1].strip()) if "TRYWAIT" in prev else ""
        print(f"{100 * n / total:5.1f}%  {r[ad][-5:]}  {r[src][:64]:64s} {why[0][1]}={why[0][0]} {why[1][1]}={why[1][0]}{note}")


if __name__ == "__main__":
    main()
