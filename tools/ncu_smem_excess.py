"""Shared-memory wavefronts against their conflict-free minimum, per opcode and per instruction, from an exported ncu source page
(tools/ncu_source.sh):  python tools/ncu_smem_excess.py gpurun_out/<tag>_d<d>.src.csv"""
import csv, sys, collections
f=sys.argv[1]
lines=open(f).read().splitlines()
start=next(i for i,l in enumerate(lines) if l.startswith('"Address"'))
rows=list(csv.reader(lines[start:]))
hdr,rows=rows[0],[r for r in rows[1:] if len(r)>=len(rows[0])]
ix={h:i for i,h in enumerate(hdr)}
def num(v):
    try: return float(v.replace(',',''))
    except: return 0.0
tw=sum(num(r[ix["L1 Wavefronts Shared"]]) for r in rows)
te=sum(num(r[ix["L1 Wavefronts Shared Excessive"]]) for r in rows)
ti=sum(num(r[ix["L1 Wavefronts Shared Ideal"]]) for r in rows)
print("shared wavefronts",tw,"ideal",ti,"excessive",te, "%.1f%%"%(100*te/tw))
# by opcode
by=collections.Counter(); byw=collections.Counter()
for r in rows:
    op=r[ix["Source"]].split()[0] if r[ix["Source"]] else ''
    if op.startswith('@'): op=r[ix["Source"]].split()[1]
    by[op]+=num(r[ix["L1 Wavefronts Shared Excessive"]]); byw[op]+=num(r[ix["L1 Wavefronts Shared"]])
for op,v in byw.most_common(12): print(op, "wavefronts %.0f excessive %.0f"%(v,by[op]))
ex=[(num(r[ix["L1 Wavefronts Shared Excessive"]]),i,r) for i,r in enumerate(rows)]
ex.sort(reverse=True)
print("top excessive instructions:")
for v,i,r in ex[:25]:
    print(i, r[ix["Source"]][:70], "wf",r[ix["L1 Wavefronts Shared"]],"ideal",r[ix["L1 Wavefronts Shared Ideal"]],"exec",r[ix["Instructions Executed"]])
