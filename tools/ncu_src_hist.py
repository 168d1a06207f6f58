"""Aggregate an `ncu --page source --csv` dump by opcode: executed warp-instructions and stall samples per opcode,
plus the hottest instructions.  Usage: ncu -i X.ncu-rep --page source --csv | python tools/ncu_src_hist.py [units]"""
import csv, sys, collections
rows = list(csv.reader(sys.stdin))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
isrc, ismp, iex = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
ops = collections.Counter(); smp = collections.Counter(); tot = 0; hot = []
for r in rows[hi + 1:]:
    if len(r) <= iex or not r[iex].isdigit():
        continue
    src = r[isrc].strip()
    toks = src.split()
    op = toks[0] if not toks[0].startswith("@") else toks[1]
    parts = op.split(".")
    op = parts[0] + ("." + parts[1] if parts[0] in ("LDS", "STS", "IMAD", "SYNCS", "BRA", "MUFU", "LDG", "STG") and len(parts) > 1 else "")
    n, s = int(r[iex]), int(r[ismp])
    ops[op] += n; smp[op] += s; tot += n
    hot.append((s, n, src))
unit = float(sys.argv[1]) if len(sys.argv) > 1 else 0.0
print(f"total warp-instructions {tot}" + (f" = {tot / unit:.1f} per unit" if unit else ""))
for op, n in ops.most_common(32):
    extra = f"  {n / unit:7.2f}/unit" if unit else ""
    print(f"{op:22s} {n:12d} {100.0 * n / tot:5.1f}%  samples {smp[op]:7d}{extra}")
print("hottest by stall samples:")
for s, n, src in sorted(hot, reverse=True)[:25]:
    print(f"{s:7d} {n:10d}  {src[:90]}")
