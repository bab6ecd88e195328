#!/usr/bin/env python3
"""Static schedule of one kernel's SASS: python tools/sass_schedule.py <lib.so> <kernel-name-substring> [--regions]

Decodes the control word of every instruction (Volta+ encoding: stall count in bits 105..108 of the 128-bit
instruction, i.e. bits 41..44 of the second 64-bit word) and prints, for the largest backward-branch loop (the
persistent main loop of the fused kernels): instruction count, opcode-class mix, and the sum of stall counts -- the
number of cycles ONE warp needs to issue the loop body once if no scoreboard ever blocks it.  With W warps per
scheduler the FMA pipe sees at most (packed FP32 instructions x 2 cycles) / (that sum / W).
A tool for the container (no GPU): it tells how much of a kernel's time is fixed-latency waiting that ptxas could not
fill, before GPU time is spent on a variant."""
import re
import subprocess
import sys
from collections import Counter

CLASSES = [
    ("fp32x2", re.compile(r"^(FFMA2|FADD2|FMUL2)")),
    ("fp32", re.compile(r"^(FFMA|FADD|FMUL|FMNMX|FSEL|FSET|FSETP)")),
    ("lds", re.compile(r"^LDS")),
    ("sts", re.compile(r"^STS")),
    ("ldg", re.compile(r"^(LDG|LD\b|LDL)")),
    ("stg", re.compile(r"^(STG|ST\b|STL)")),
    ("shfl", re.compile(r"^SHFL")),
    ("mov", re.compile(r"^(MOV|IMAD\.MOV|UMOV|PRMT|SEL|R2UR|S2R|S2UR|CS2R)")),
    ("int", re.compile(r"^(IADD|IMAD|LOP|SHF|LEA|ISETP|IABS|UIADD|ULOP|USHF|UIMAD|ULEA|UISETP|VIADD|I2F|F2I|I2FP|POPC|BREV|FLO|PLOP|UPLOP|UPRMT|USEL|VOTE|MATCH)")),
    ("sync", re.compile(r"^(BAR|SYNCS|WARPSYNC|BSYNC|BSSY|MEMBAR|FENCE|NANOSLEEP|ERRBAR|DEPBAR|UBLKCP|ATOM|RED|CCTL)")),
    ("branch", re.compile(r"^(BRA|EXIT|CALL|RET|BRX|JMP)")),
]


def classify(op):
    for name, rx in CLASSES:
        if rx.match(op):
            return name
    return "other"


def parse(lib, name):
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    insts = []
    take = False
    pending = None
    for line in out.splitlines():
        if "Function :" in line:
            take = name in line
            if take and insts:
                break
            continue
        if not take:
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\* (0x[0-9a-f]{16}) \*/", line)
        if m:
            pending = (int(m.group(1), 16), m.group(2).strip(), int(m.group(3), 16))
            continue
        m = re.match(r"\s+/\* (0x[0-9a-f]{16}) \*/", line)
        if m and pending:
            hi = int(m.group(1), 16)
            stall = (hi >> 41) & 0xF
            yld = (hi >> 45) & 1
            text = pending[1]
            pred = ""
            if text.startswith("@"):
                pred, text = text.split(None, 1)
            insts.append(dict(addr=pending[0], text=pending[1], op=text.split()[0], stall=stall, yld=yld))
            pending = None
    return insts


def main():
    lib, name = sys.argv[1], sys.argv[2]
    insts = parse(lib, name)
    if not insts:
        sys.exit("kernel not found")
    by_addr = {i["addr"]: n for n, i in enumerate(insts)}
    loops = []
    for n, i in enumerate(insts):
        if i["op"].startswith("BRA"):
            m = re.search(r"0x([0-9a-f]+)", i["text"])
            if m:
                tgt = int(m.group(1), 16)
                if tgt in by_addr and by_addr[tgt] < n:
                    loops.append((n - by_addr[tgt], by_addr[tgt], n))
    print(f"{name}: {len(insts)} instructions, {len(loops)} backward branches")
    if not loops:
        return
    loops.sort(reverse=True)
    _, a, b = loops[0]
    body = insts[a:b + 1]
    mix = Counter(classify(i["op"]) for i in body)
    stall_by = Counter()
    for i in body:
        stall_by[classify(i["op"])] += max(i["stall"], 1)
    total = sum(max(i["stall"], 1) for i in body)
    print(f"main loop: {len(body)} instructions, static issue time of one warp {total} cycles ({total / len(body):.2f} per instruction)")
    for k, v in mix.most_common():
        print(f"  {k:8s} {v:6d}  stall cycles {stall_by[k]:6d}")
    if "--regions" in sys.argv:
        # cumulative profile in slices of 200 instructions: where the waiting sits
        for s in range(0, len(body), 200):
            sl = body[s:s + 200]
            c = Counter(classify(i["op"]) for i in sl)
            t = sum(max(i["stall"], 1) for i in sl)
            top = ", ".join(f"{k} {v}" for k, v in c.most_common(4))
            print(f"  [{s:5d}] {t:5d} cycles  {top}")


if __name__ == "__main__":
    main()
