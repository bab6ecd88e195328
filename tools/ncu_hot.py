"""Where a kernel's warp-stall samples sit: python tools/ncu_hot.py <source-page csv> [slice=100]
Reads an exported ncu source page (tools/ncu_source.sh), prints per slice of the loop body the samples, the dominant
stall reasons and opcode mix, then the 25 hottest instructions."""
import collections
import csv
import sys

f = sys.argv[1]
step = int(sys.argv[2]) if len(sys.argv) > 2 else 100
lines = open(f).read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.reader(lines[start:]))
hdr, rows = rows[0], [r for r in rows[1:] if len(r) >= len(rows[0])]
ix = {h: i for i, h in enumerate(hdr)}
reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]


def num(v):
    try:
        return float(v.replace(",", ""))
    except ValueError:
        return 0.0


tot = sum(num(r[ix["# Samples"]]) for r in rows)
print(f"{len(rows)} instructions, {tot:.0f} samples")
rt = collections.Counter()
for r in rows:
    for h in reasons:
        rt[h] += num(r[ix[h]])
print("  ".join(f"{k[6:]} {100 * v / tot:.1f}%" for k, v in rt.most_common(9)))
for s in range(0, len(rows), step):
    sl = rows[s:s + step]
    n = sum(num(r[ix["# Samples"]]) for r in sl)
    if n < 0.002 * tot:
        continue
    c = collections.Counter()
    ops = collections.Counter()
    for r in sl:
        for h in reasons:
            c[h] += num(r[ix[h]])
        t = r[ix["Source"]].split()
        op = t[1] if t and t[0].startswith("@") and len(t) > 1 else (t[0] if t else "?")
        ops[op.split(".")[0]] += 1
    top = ", ".join(f"{k[6:]} {100 * v / max(n, 1):.0f}%" for k, v in c.most_common(4))
    mix = ", ".join(f"{k} {v}" for k, v in ops.most_common(4))
    print(f"[{s:5d}] {100 * n / tot:5.1f}% of samples, {n / len(sl):6.0f}/instr  {top:60s} {mix}")
print("hottest instructions:")
for r in sorted(rows, key=lambda r: -num(r[ix["# Samples"]]))[:25]:
    c = {h: num(r[ix[h]]) for h in reasons}
    k = max(c, key=c.get)
    print(f"  {num(r[ix['# Samples']]):7.0f}  {k[6:]:14s} {r[ix['Source']][:90]}")
