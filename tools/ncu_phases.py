"""Per-phase view of an ncu source page (SASS + warp-stall samples) of one kernel.
    python tools/ncu_phases.py report.ncu-rep <kernel regex> [--split BAR|opcode]
The loop body is cut at every BAR.SYNC (phase boundary of the dual-frame kernels); for every segment
prints instruction count, stall samples, samples per instruction and the mix of opcodes and stall reasons."""
import collections
import csv
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
split = sys.argv[4] if len(sys.argv) > 4 and sys.argv[3] == "--split" else "BAR"
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat],
                     capture_output=True, text=True, check=True).stdout
lines = raw.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.reader(lines[start:]))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]


def num(v):
    try:
        return float(v.replace(",", ""))
    except ValueError:
        return 0.0


def opclass(sass):
    t = sass.split()
    op = t[1] if t and t[0].startswith("@") else (t[0] if t else "?")
    return op.split(".")[0]


segs = []
cur = dict(n=0, samples=0.0, execd=0.0, ops=collections.Counter(), stalls=collections.Counter(), opsamp=collections.Counter())
for r in rows[1:]:
    if len(r) < len(hdr):
        continue
    sass = r[ix["Source"]].strip()
    op = opclass(sass)
    s = num(r[ix["# Samples"]])
    cur["n"] += 1
    cur["samples"] += s
    cur["execd"] += num(r[ix["Instructions Executed"]])
    cur["ops"][op] += 1
    cur["opsamp"][op] += s
    for c in stall_cols:
        cur["stalls"][c] += num(r[ix[c]])
    if op == split or (split == "BAR" and op in ("BAR", "BRA") and op == "BAR"):
        segs.append(cur)
        cur = dict(n=0, samples=0.0, execd=0.0, ops=collections.Counter(), stalls=collections.Counter(), opsamp=collections.Counter())
segs.append(cur)
tot = sum(s["samples"] for s in segs) or 1.0
print(f"kernel {pat}: {sum(s['n'] for s in segs)} SASS instructions, {tot:.0f} samples")
for i, s in enumerate(segs):
    if s["n"] == 0:
        continue
    ops = ", ".join(f"{k}:{v}" for k, v in s["ops"].most_common(7))
    osm = ", ".join(f"{k}:{100 * v / tot:.1f}%" for k, v in s["opsamp"].most_common(5))
    st = ", ".join(f"{k[6:]}:{100 * v / tot:.1f}%" for k, v in s["stalls"].most_common(6))
    print(f"seg {i:2d}: {s['n']:5d} instr  {100 * s['samples'] / tot:5.1f}% samples  {s['samples'] / max(s['n'], 1):6.1f} samp/instr  exec {s['execd'] / max(s['n'], 1):8.0f}")
    print(f"        ops    {ops}")
    print(f"        samples by op {osm}")
    print(f"        stalls {st}")
