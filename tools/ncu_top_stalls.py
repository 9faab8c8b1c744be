"""Reads `ncu -i X.ncu-rep --page source --csv` from stdin or a file and prints the instructions with the most stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
col = {n: i for i, n in enumerate(h)}
stall_cols = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
body = [r for r in rows[hi + 1:] if len(r) == len(h)]
total = sum(int(r[col["# Samples"]] or 0) for r in body)
top = sorted(body, key=lambda r: -int(r[col["# Samples"]] or 0))[: int(sys.argv[2]) if len(sys.argv) > 2 else 30]
print("total samples", total, " instructions", len(body), " executed warp-instr", sum(int(r[col["Instructions Executed"]] or 0) for r in body))
order = {id(r): i for i, r in enumerate(body)}
for r in top:
    s = int(r[col["# Samples"]] or 0)
    why = sorted(((int(r[col[c]] or 0), c) for c in stall_cols), reverse=True)[:2]
    print(f"{order[id(r)]:5d} {100.0 * s / max(total, 1):5.1f}%  {r[col['Source']].strip():60s} {why}")
