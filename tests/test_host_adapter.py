"""The C++ side of the boundary: the stand-alone ring buffer (CPU) and the `cuda_r2iq` adapter (GPU)."""
import os
import subprocess

import pytest

from conftest import ROOT
from extio_sddc_b200 import capi

HOST = os.path.join(ROOT, "extio_sddc_b200", "host")
OUT = os.path.join(ROOT, "gpurun_out")
INC = ["-I", os.path.join(ROOT, "include"), "-I", HOST]


def _compile(src, exe, extra=()):
    os.makedirs(OUT, exist_ok=True)
    cmd = ["g++", "-std=c++17", "-O2", "-pthread", *INC, src, *extra, "-o", exe]
    subprocess.run(cmd, check=True)
    return exe


def test_ringbuffer_mirror():
    exe = _compile(os.path.join(ROOT, "tests", "host", "test_ringbuffer.cpp"), os.path.join(OUT, "test_ringbuffer"))
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "ALL PASS" in r.stdout, r.stdout + r.stderr


@pytest.mark.skipif(not os.path.isdir("/root/reference/Core"), reason="reference tree not present on this box")
def test_adapter_compiles_against_reference_headers():
    # drop-in check: the adapter derives from the reference's own r2iqControlClass / ringbuffer<T>
    subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-DR2IQ_USE_REFERENCE_HEADERS", *INC, "-I", "/root/reference/Core",
                    os.path.join(HOST, "cuda_r2iq.cpp")], check=True)


@pytest.mark.gpu
def test_cuda_r2iq_adapter_streams_through_rings():
    capi.load()
    lib_dir = os.path.dirname(capi.LIB_PATH)
    exe = _compile(os.path.join(ROOT, "tests", "host", "test_cuda_r2iq.cpp"), os.path.join(OUT, "test_cuda_r2iq"),
                   extra=[os.path.join(HOST, "cuda_r2iq.cpp"), "-L", lib_dir, "-lr2iq_cuda", f"-Wl,-rpath,{lib_dir}"])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "ALL PASS" in r.stdout, r.stdout + r.stderr
