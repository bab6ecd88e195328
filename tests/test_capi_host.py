"""CPU suite, part 2: the drop-in boundary and the product's host-side logic (no compute calls).

 * the C-ABI library builds/loads and exports every symbol include/r2iq_cuda.h declares;
 * the device-free host logic (filter design, tune-bin quantisation) matches the golden vectors
   generated from the reference;
 * without a GPU the backend refuses to work instead of falling back to a CPU path.
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT
from extio_sddc_b200 import capi, r2iq


def test_library_exports_every_declared_symbol():
    lib = capi.load()
    declared = capi.declared_symbols()
    assert len(declared) >= 28
    assert set(declared) == set(capi._SIGNATURES), "capi.py binding table out of sync with include/r2iq_cuda.h"
    out = subprocess.run(["nm", "-D", "--defined-only", capi.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    missing = [s for s in declared if s not in exported]
    assert not missing, missing
    for s in declared:
        assert getattr(lib, s) is not None
    assert b"sm_100a" in lib.r2iq_cuda_version()


def test_library_contains_sm100a_sass_and_no_cufft():
    out = subprocess.run(["cuobjdump", "-lelf", capi.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump not available")
    assert "sm_100a" in out.stdout
    needed = subprocess.run(["readelf", "-d", capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "cufft" not in needed.lower()
    assert "libcudart" not in needed.lower()   # static runtime: loads beside torch's own


def test_header_is_plain_c():
    # the boundary must compile as C (no C++/torch types in the signatures)
    src = '#include "r2iq_cuda.h"\nint main(void){ r2iq_cuda_config c = {0, 1.0f, 0}; (void)c; return R2IQ_OK; }\n'
    p = os.path.join(ROOT, "gpurun_out", "_hdr_check.c")
    os.makedirs(os.path.dirname(p), exist_ok=True)
    with open(p, "w") as f:
        f.write(src)
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), p], check=True)


def test_host_filter_design_matches_reference_golden():
    z = np.load(os.path.join(GOLDEN, "filters.npz"))
    gain = 1.08e-8
    for d in range(7):
        bw = np.float32(64.0) / np.float32(1 << d)
        taps = r2iq.kaiser_taps(257, 120.0, np.float32(0.85) * bw / np.float32(128), np.float32(1.1) * bw / np.float32(128))
        assert np.array_equal(taps, z[f"taps_d{d}"]), f"Kaiser taps differ from Core/fir.cpp at d={d}"
        hw = r2iq.design_filter(gain, d)
        ref = z[f"hw_d{d}"].astype(np.complex128)
        err = np.linalg.norm(hw.astype(np.complex128) - ref) / np.linalg.norm(ref)
        assert err < 1e-6, (d, err)


def test_freq_to_tunebin_matches_reference_golden():
    rows = np.load(os.path.join(GOLDEN, "freq_offset.npy"))
    for off, d, tb, res in rows:
        got_tb, got_res = r2iq.freq_to_tunebin(off, int(d))
        assert got_tb == int(tb), (off, d)
        assert got_res == pytest.approx(res, abs=1e-9)


def test_argument_validation_without_device():
    lib = capi.load()
    assert lib.r2iq_cuda_design_filter(1.0, 7, None, 1024) == capi.E_INVALID
    assert b"design_filter" in lib.r2iq_cuda_last_error()
    assert lib.r2iq_cuda_kaiser(0, 120.0, 0.1, 0.2, None) == capi.E_INVALID
    assert lib.r2iq_cuda_create(None, None) == capi.E_INVALID
    assert lib.r2iq_cuda_destroy(None) == capi.OK
    assert lib.r2iq_cuda_set_tunebin(None, 4) == capi.E_INVALID


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_have_gpu(), reason="this box has a GPU")
def test_no_cpu_fallback_without_gpu():
    op = r2iq.CudaR2IQ()
    with pytest.raises(capi.R2iqError) as ei:
        op.Init(1.0)
    assert ei.value.code == capi.E_NODEVICE
    assert "no CPU path" in str(ei.value)


def test_product_never_imports_oracle():
    # only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may touch oracle/
    pkg = os.path.join(ROOT, "extio_sddc_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                with open(os.path.join(dirpath, fn), errors="replace") as f:
                    text = f.read()
                assert "oracle" not in text.replace("r2iq_oracle", "oracle") or "import oracle" not in text, fn
                assert "from oracle" not in text and "import oracle" not in text and "oracle/" not in text, fn


def test_numa_helpers_degrade_without_device(monkeypatch):
    """r2iq_cuda_device_numa_node / r2iq_cuda_host_alloc_on never guess: no device (or R2IQ_NUMA=0) -> node -1, and a
    failed page-locked allocation is an error with a message, not a pageable buffer."""
    import torch
    lib = capi.load()
    monkeypatch.setenv("R2IQ_NUMA", "0")
    assert lib.r2iq_cuda_device_numa_node(0) == -1
    monkeypatch.delenv("R2IQ_NUMA")
    if not torch.cuda.is_available():
        assert lib.r2iq_cuda_device_numa_node(0) == -1
        assert not lib.r2iq_cuda_host_alloc_on(4096, 0)
        assert "cudaMallocHost" in capi.last_error()


def test_fft_plan_argument_checks_need_no_device():
    """r2iq_cuda_fft_plan (Core/fft_backend.h:30-31 plan_r2c / plan_c2c) refuses unsupported lengths, kinds and batches
    with R2IQ_E_INVALID before it looks for a device; a supported request on a box without a GPU says so."""
    import ctypes

    import torch
    lib = capi.load()
    h = ctypes.c_void_p()
    for kind, n, batch in ((capi.FFT_C2C_FORWARD, 1000, 1), (capi.FFT_C2C_FORWARD, 8192, 1), (capi.FFT_R2C, 64, 1),
                           (capi.FFT_R2C, 16384, 1), (7, 1024, 1), (capi.FFT_C2C_BACKWARD, 1024, 0)):
        assert lib.r2iq_cuda_fft_plan(-1, kind, n, batch, ctypes.byref(h)) == capi.E_INVALID, (kind, n, batch)
        assert not h.value and capi.last_error()
    if not torch.cuda.is_available():
        assert lib.r2iq_cuda_fft_plan(-1, capi.FFT_R2C, 2048, 1, ctypes.byref(h)) == capi.E_NODEVICE
        assert "no CPU path" in capi.last_error()
    assert lib.r2iq_cuda_fft_destroy(None) == capi.OK
