"""Experiment builds of the CUDA library: python tools/exp_build.py NAME -DFOO [-DBAR ...]
compiles the product sources with extra defines into extio_sddc_b200/lib_exp/NAME/libr2iq_cuda.so (git-ignored, travels
to the GPU box) so that several variants can be timed in ONE gpurun call: python tools/gpu_check.py --lib <that file>."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from extio_sddc_b200 import build as b  # noqa: E402

name, defines = sys.argv[1], sys.argv[2:]
out_dir = os.path.join(ROOT, "extio_sddc_b200", "lib_exp", name)
os.makedirs(out_dir, exist_ok=True)


def one(src):
    obj = os.path.join(out_dir, os.path.splitext(os.path.basename(src))[0] + ".o")
    subprocess.run([b.nvcc(), *b.ARCH_FLAGS, *b.NVCC_FLAGS, *defines, "-c", os.path.join(b.CSRC, src), "-o", obj], check=True)
    return obj


with ThreadPoolExecutor(max_workers=os.cpu_count()) as pool:
    objs = list(pool.map(one, b.SOURCES))
lib = os.path.join(out_dir, "libr2iq_cuda.so")
subprocess.run([b.nvcc(), *b.ARCH_FLAGS, "--cudart", "static", "-shared", "-o", lib, *objs], check=True)
print(lib)
