"""In-tree build of the CUDA backend: nvcc -> extio_sddc_b200/lib/libr2iq_cuda.so (sm_100a only).

The library is a plain C-ABI shared object (include/r2iq_cuda.h) with the CUDA runtime linked
statically, so it loads next to torch's own runtime without symbol clashes.  Built files are
git-ignored but travel to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libr2iq_cuda.so")
# The shipped library: one fused kernel per (FFTN, decimation) plus channelizer, mixer pass and the C ABI.
SOURCES = ["r2iq_pair.cu", "r2iq_launch.cu", "r2iq_channels.cu", "r2iq_wide_pair.cu", "r2iq_wide_channels.cu", "r2iq_fft.cu", "r2iq_capi.cu", "fir_design.cpp"]
# R2IQ_AB_VARIANTS=1 at build time adds the older kernel generations and their run-time switches (R2IQ_KERNEL_GEN,
# R2IQ_VARIANT, R2IQ_CTAS_PER_SM) for comparison runs (tools/variant_sweep.py); `r2iq_cuda_version()` then says so.
SOURCES += [os.path.join("wide_tu", f"r2iq_wide_pair_d{d}{m}.cu") for d in range(7) for m in ("", "m")]   # one module per kernel
AB_VARIANTS = os.environ.get("R2IQ_AB_VARIANTS", "0") not in ("", "0")
if AB_VARIANTS:
    SOURCES += ["r2iq_wide.cu", "r2iq_gen1.cu", "r2iq_dual.cu", "r2iq_dual_d0.cu", "r2iq_dual_d1.cu", "r2iq_dual_dn.cu"]
HEADERS = ["r2iq_wide_pair.cu", "r2iq_kernels.cuh", "dft_regs.cuh", "dft_regs2.cuh", "r2iq_common.cuh", "work_dispenser.cuh", "dual_frame.cuh", "fir_design.h", os.path.join("..", "..", "include", "r2iq_cuda.h")]
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
              "--cudart", "static", "-Xfatbin", "-compress-all"] + (["-DR2IQ_AB_VARIANTS"] if AB_VARIANTS else [])


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; the r2iq CUDA backend cannot be built (there is no CPU fallback)")


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    if any(os.path.getmtime(d) > t for d in deps):
        return True
    flavour = os.path.join(LIB_DIR, ".flavour")          # a library of the other flavour is stale too
    return not os.path.exists(flavour) or open(flavour).read().strip() != ("ab" if AB_VARIANTS else "product")


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale():
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    from concurrent.futures import ThreadPoolExecutor
    hdr_time = max(os.path.getmtime(os.path.join(CSRC, h)) for h in HEADERS)
    hdr_time = max(hdr_time, os.path.getmtime(os.path.abspath(__file__)))

    def compile_one(src: str) -> str:
        obj = os.path.join(LIB_DIR, os.path.splitext(os.path.basename(src))[0] + ".o")
        path = os.path.join(CSRC, src)
        obj = obj[:-2] + ("_ab.o" if AB_VARIANTS else ".o")      # the two flavours never share objects
        if not force and os.path.exists(obj) and os.path.getmtime(obj) >= max(os.path.getmtime(path), hdr_time):
            return obj                    # translation unit and every header older than its object
        cmd = [nvcc(), *ARCH_FLAGS, *NVCC_FLAGS, "-c", path, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas")
            cmd.insert(2, "-v")
            print(" ".join(cmd), flush=True)
        subprocess.run(cmd, check=True)
        return obj

    with ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 1)) as pool:
        objs = list(pool.map(compile_one, SOURCES))
    cmd = [nvcc(), *ARCH_FLAGS, "--cudart", "static", "-shared", "-o", LIB_PATH, *objs]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.run(cmd, check=True)
    with open(os.path.join(LIB_DIR, ".flavour"), "w") as f:
        f.write("ab" if AB_VARIANTS else "product")
    return LIB_PATH


HOST = os.path.join(HERE, "host")
BIN_DIR = os.path.join(HERE, "bin")
REPLAY_PATH = os.path.join(BIN_DIR, "r2iq_replay")
FFT_BENCH_PATH = os.path.join(BIN_DIR, "fft_benchmark_cuda")


def build_tools(force: bool = False) -> str:
    """Host-side tools over the C ABI (g++ only): bin/r2iq_replay, the capture-file replayer (host/r2iq_replay.cpp)."""
    srcs = [os.path.join(HOST, "r2iq_replay.cpp"), os.path.join(HOST, "wav_capture.cpp")]
    deps = srcs + [os.path.join(HOST, f) for f in ("wav_capture.h", "fft_benchmark_cuda.cpp", "cuda_fft_backend.cpp", "cuda_fft_backend.h")] + [LIB_PATH]
    if not force and os.path.exists(REPLAY_PATH) and os.path.exists(FFT_BENCH_PATH) and all(os.path.getmtime(d) <= os.path.getmtime(REPLAY_PATH) for d in deps):
        return REPLAY_PATH
    os.makedirs(BIN_DIR, exist_ok=True)
    common = ["g++", "-std=c++17", "-O2", "-I", os.path.join(HERE, "..", "include"), "-I", HOST]
    link = ["-L", LIB_DIR, "-lr2iq_cuda", "-Wl,-rpath,$ORIGIN/../lib"]
    subprocess.run([*common, *srcs, *link, "-o", REPLAY_PATH], check=True)
    # bin/fft_benchmark_cuda: the reference's FFT backend benchmark table for CudaFFTBackend (host/fft_benchmark_cuda.cpp)
    subprocess.run([*common, os.path.join(HOST, "fft_benchmark_cuda.cpp"), os.path.join(HOST, "cuda_fft_backend.cpp"), *link,
                    "-o", FFT_BENCH_PATH], check=True)
    return REPLAY_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_tools(force="--force" in sys.argv))
