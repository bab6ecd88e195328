"""Opcode histogram of every kernel in the shipped library (cuobjdump -sass on extio_sddc_b200/lib/libr2iq_cuda.so),
written to profiles/<tag>_sass_opcodes.txt: the static evidence that the hot kernels are packed-FP32 (FFMA2 / FADD2 /
FMUL2) and TMA (UBLKCP + SYNCS mbarrier waits) code with no local-memory spills (STL / LDL) and no conversion-unit
I2F on the sample path.  Runs without a GPU.   usage: python tools/sass_opcodes.py [tag=r02]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from extio_sddc_b200 import capi  # noqa: E402

tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
sass = subprocess.run(["cuobjdump", "-sass", capi.LIB_PATH], capture_output=True, text=True, check=True).stdout
demangle = {}
kernels = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        kernels[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_]+)*)", line)
    if m and cur:
        op, mods = m.group(1), m.group(2)
        kernels[cur][op] += 1
        if op in ("LDS", "STS", "LDG", "STG") and mods:
            width = [w for w in mods.split(".") if w in ("64", "128")]
            kernels[cur][op + "." + (width[0] if width else "32")] += 1
names = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
cols = ["FFMA2", "FADD2", "FMUL2", "FFMA", "FADD", "FMUL", "UBLKCP", "SYNCS", "LDS", "STS", "LDG", "STG", "STL", "LDL", "BAR",
        "I2F", "PRMT", "SHFL", "ATOMG", "total"]
out = [f"SASS opcode counts per kernel of {os.path.relpath(capi.LIB_PATH, ROOT)} ({os.path.getsize(capi.LIB_PATH)} bytes), "
       f"`cuobjdump -sass`, static instruction counts (whole function: prologue + loop)",
       "UBLKCP = TMA bulk copy, SYNCS = mbarrier arrive/try_wait, ATOMG = the device-wide work counter; STL/LDL = local-memory spills",
       ""]
out.append(f"{'kernel':78s} " + " ".join(f"{c:>6s}" for c in cols))
for (mangled, cnt), name in zip(kernels.items(), names):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\((?:r2iq::)?\w+(?: const\*| \*|\*)?(?:, [^)]*)?\)$", "", name).replace("r2iq::", "").replace("(int)", "").replace("(bool)", "")
    name = name.replace("(anonymous namespace)::", "")
    cnt["total"] = sum(v for k, v in cnt.items() if "." not in k and k != "total")
    out.append(f"{name[:78]:78s} " + " ".join(f"{cnt.get(c, 0):6d}" for c in cols))
path = os.path.join(ROOT, "profiles", f"{tag}_sass_opcodes.txt")
with open(path, "w") as f:
    f.write("\n".join(out) + "\n")
print("\n".join(out))
