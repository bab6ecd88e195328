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


@pytest.mark.gpu
def test_radiohandler_drop_in_matches_reference_operator():
    """the reference's UNMODIFIED RadioHandlerClass (Init/UptRand/UpdatemodeRF/TuneLO/Start/Stop, its
    radio models and its CPU fine-tune mixer, compiled in place into oracle/_ref/ref_radio_dropin) is run with its own
    fft_mt_r2iq, then with cuda_r2iq passed as the third Init argument (Core/RadioHandler.h:34), on the same mock-USB
    capture; the IQ reaching the application callback must agree within rel-L2 1e-5 / 1e-4 of full scale, and the
    device-side mixer must equal the CPU mixer bit for bit (checked inside the binary, FFTN 8192 as shipped)."""
    from oracle import ref_lib
    if not os.path.exists(ref_lib.RADIO_DROPIN):
        pytest.skip("oracle/_ref/ref_radio_dropin not built (needs /root/reference at build time)")
    capi.load()
    r = subprocess.run([ref_lib.RADIO_DROPIN], capture_output=True, text=True, timeout=600)
    lines = [l for l in r.stdout.splitlines() if l.startswith(("PASS", "FAIL", "ALL PASS"))]
    print("\n".join(lines))
    assert r.returncode == 0 and "ALL PASS" in r.stdout, "\n".join(lines) + r.stderr[-2000:]
    assert sum(l.startswith("PASS") for l in lines) == 4


# ---- capture file format: libsddc/wavehdr.h + wavewrite.c ---------------------------------------
GOLDEN = os.path.join(ROOT, "tests", "golden")
WAV_SRC = os.path.join(HOST, "wav_capture.cpp")


def test_wav_capture_matches_reference_writer():
    """Reader/writer of the reference's RIFF/auxi capture format against files written by libsddc/wavewrite.c itself
    (tests/golden/ref_capture*.wav): fields recovered, re-packed header byte-identical, round trip, unfinalised file."""
    exe = _compile(os.path.join(ROOT, "tests", "host", "test_wav_capture.cpp"), os.path.join(OUT, "test_wav_capture"), extra=[WAV_SRC])
    r = subprocess.run([exe, GOLDEN, OUT], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "ALL PASS" in r.stdout, r.stdout + r.stderr


def _build_replay():
    capi.load()
    lib_dir = os.path.dirname(capi.LIB_PATH)
    return _compile(os.path.join(HOST, "r2iq_replay.cpp"), os.path.join(OUT, "r2iq_replay"),
                    extra=[WAV_SRC, "-L", lib_dir, "-lr2iq_cuda", f"-Wl,-rpath,{lib_dir}"])


def _write_capture(path, samples):
    """A capture as the reference records it: the 120-byte header the reference's writer produced (golden file) with
    the two size fields set for this payload, then raw little-endian int16."""
    import struct
    hdr = bytearray(open(os.path.join(GOLDEN, "ref_capture.wav"), "rb").read(120))
    nbytes = samples.size * 2
    hdr[4:8] = struct.pack("<I", 112 + nbytes)
    hdr[116:120] = struct.pack("<I", nbytes)
    with open(path, "wb") as f:
        f.write(hdr)
        f.write(samples.astype("<i2").tobytes())


def test_replay_tool_has_no_cpu_path():
    import numpy as np
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by test_replay_capture_matches_oracle")
    exe = _build_replay()
    cap = os.path.join(OUT, "replay_nogpu.wav")
    _write_capture(cap, np.zeros(65536, dtype=np.int16))
    r = subprocess.run([exe, cap, os.path.join(OUT, "replay_nogpu_out.wav"), "--decimate", "0"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1 and "no CPU path" in r.stderr, r.stdout + r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("d,tune_hz,tunebin,rand", [(0, None, 256, False), (3, 9.0e6, 144, True), (5, 9.0e6, 144, True)])
def test_replay_capture_matches_oracle(d, tune_hz, tunebin, rand):
    """A recorded 128 Msps capture (reference header, 20 transfers + a partial one that must be dropped) replayed in
    batches through r2iq_replay equals one oracle pass over the same transfers; the IQ file carries the rate and
    tuned frequency in the reference's own container."""
    import struct
    import numpy as np
    from oracle import r2iq_oracle as orc
    exe = _build_replay()
    gain = orc.RX888MK2_GAINFACTOR
    nb = 20
    x = orc.rx888_blocks(nb, d, tunebin, randomized=rand, seed=77 + d).reshape(-1)
    cap = os.path.join(OUT, f"replay_d{d}.wav")
    out = os.path.join(OUT, f"replay_d{d}_iq.wav")
    _write_capture(cap, np.concatenate([x, np.arange(1000, dtype=np.int16)]))
    cmd = [exe, cap, out, "--decimate", str(d), "--batch-blocks", "8"] + (["--rand"] if rand else [])
    if tune_hz is not None:
        cmd += ["--tune-hz", repr(tune_hz)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = open(out, "rb").read()
    assert raw[:4] == b"RIFF" and raw[8:12] == b"WAVE" and raw[36:40] == b"auxi" and raw[112:116] == b"data"
    tag, ch, rate = struct.unpack("<HHI", raw[20:28])
    bits, = struct.unpack("<H", raw[34:36])
    center, adrate = struct.unpack("<II", raw[76:84])
    nbytes, = struct.unpack("<I", raw[116:120])
    assert (tag, ch, bits) == (3, 2, 32) and rate == 128000000 // (2 << d) and adrate == 128000000
    assert center == (int(tune_hz) if tune_hz else 0)
    got = np.frombuffer(raw[120:120 + nbytes], dtype=np.complex64)
    ref = orc.r2iq(x.reshape(nb, 65536), d, tunebin, gain=gain, rand=rand)
    assert got.shape == ref.shape == (nb * (32768 >> d),)
    err = orc.rel_l2(got, ref)
    mx = float(np.abs(got - ref).max()) / orc.full_scale(gain)
    assert err <= 1e-5 and mx <= 1e-4, (err, mx)   # north-star tolerance


# ---- FFT backend abstraction: Core/fft_backend.h + Core/fft_benchmark.cpp ------------------------
@pytest.mark.gpu
def test_cuda_fft_backend_through_reference_interface():
    """CudaFFTBackend driven through the FFTBackend interface alone (alloc / plan / execute / destroy / free) at the
    benchmark's six sizes, r2c + c2c both ways, against a float64 DFT: rel-L2 <= 1e-6 (tests/host/test_fft_backend.cpp)."""
    capi.load()
    lib_dir = os.path.dirname(capi.LIB_PATH)
    exe = _compile(os.path.join(ROOT, "tests", "host", "test_fft_backend.cpp"), os.path.join(OUT, "test_fft_backend"),
                   extra=[os.path.join(HOST, "cuda_fft_backend.cpp"), "-L", lib_dir, "-lr2iq_cuda", f"-Wl,-rpath,{lib_dir}"])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "ALL PASS" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_reference_fft_benchmark_runs_on_cuda_backend():
    """The reference's UNMODIFIED Core/fft_benchmark.cpp, linked with CudaFFTBackend as getFFTBackend()
    (oracle/_ref/fft_benchmark_dropin, built here where the reference tree is), prints its table for all six sizes."""
    exe = os.path.join(ROOT, "oracle", "_ref", "fft_benchmark_dropin")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/fft_benchmark_dropin not built (needs /root/reference at build time)")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Backend: CUDA sm_100a register FFT" in r.stdout
    rows = [l for l in r.stdout.splitlines() if l.strip()[:4].strip().isdigit() and "|" in l]
    assert [int(l.split("|")[0]) for l in rows] == [4096, 2048, 1024, 512, 256, 128], r.stdout
    for l in rows:                       # every timing is a positive number of microseconds
        assert all(float(c) > 0 for c in l.split("|")[1:]), l


@pytest.mark.skipif(not os.path.isdir("/root/reference/Core"), reason="reference tree not present on this box")
def test_fft_backend_compiles_against_reference_header():
    subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-DR2IQ_USE_REFERENCE_HEADERS", "-DR2IQ_FFT_BACKEND_FACTORY", *INC,
                    "-I", "/root/reference/Core", os.path.join(HOST, "cuda_fft_backend.cpp")], check=True)


# ---- consumers above RadioHandlerClass ----------------------------------------------------------
REF_BIN = os.path.join(ROOT, "oracle", "_ref")


@pytest.mark.gpu
@pytest.mark.parametrize("rate,blocks,rand", [(32000000, 48, 0), (8000000, 96, 1), (2000000, 128, 0)])
def test_libsddc_sync_ring_delivers_reference_iq(rate, blocks, rand):
    """The reference's UNMODIFIED libsddc.cpp (mock USB layer, -DLIBSDDC_TESTING) streams a capture through
    sddc_start_streaming / sddc_read_sync (callback -> 1 MiB byte ring, libsddc/libsddc.cpp:79-126,659-695) twice: linked
    with the reference's fft_mt_r2iq, and with host/fft_mt_r2iq_cuda.cpp in its place (the GPU operator behind the same
    `new fft_mt_r2iq()` of RadioHandler.cpp:94-95).  The bytes read back must agree at the north-star bar."""
    import numpy as np
    cpu, gpu = os.path.join(REF_BIN, "ref_libsddc_cpu"), os.path.join(REF_BIN, "ref_libsddc_cuda")
    if not (os.path.exists(cpu) and os.path.exists(gpu)):
        pytest.skip("oracle/_ref/ref_libsddc_{cpu,cuda} not built (needs /root/reference at build time)")
    os.makedirs(OUT, exist_ok=True)
    outs = []
    for exe, tag in ((cpu, "cpu"), (gpu, "cuda")):
        path = os.path.join(OUT, f"libsddc_{tag}_{rate}.bin")
        r = subprocess.run([exe, str(rate), str(blocks), path, str(rand)], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0 and "DONE" in r.stdout, r.stdout + r.stderr[-2000:]
        outs.append(np.fromfile(path, dtype=np.float32))
        os.remove(path)
    ref, got = outs
    d = {32000000: 0, 8000000: 2, 2000000: 4}[rate]
    assert ref.size == got.size == (blocks >> d) * 32768 * 2
    skip = 2 * (32768 >> d)                     # IQ of input block 0: the reference reads uninitialised history there
    err = np.linalg.norm(got[skip:].astype(np.float64) - ref[skip:]) / np.linalg.norm(ref[skip:].astype(np.float64))
    peak = float(np.abs(ref[skip:]).max())
    assert peak > 0
    assert err <= 1e-5, err
    assert float(np.abs(got[skip:] - ref[skip:]).max()) <= 1e-4 * peak * 4     # peak >= full scale / 4 for this capture


@pytest.mark.gpu
@pytest.mark.parametrize("rate,blocks", [(32000000, 64), (4000000, 128)])
def test_soapysddc_stream_buffers_deliver_reference_iq(rate, blocks):
    """The reference's UNMODIFIED SoapySDDC module (16 buffers of 32768 CF32 samples, overflow event: Settings.cpp:14-35,
    Streaming.cpp:110-182), built in place against declaration-only SoapySDR headers, read through readStream twice:
    with the reference's fft_mt_r2iq and with host/fft_mt_r2iq_cuda.cpp in its place.  Same IQ at the north-star bar;
    with a free-running producer the overflow event fires and streaming continues (checked inside the binary)."""
    import numpy as np
    cpu, gpu = os.path.join(REF_BIN, "ref_soapy_cpu"), os.path.join(REF_BIN, "ref_soapy_cuda")
    if not (os.path.exists(cpu) and os.path.exists(gpu)):
        pytest.skip("oracle/_ref/ref_soapy_{cpu,cuda} not built (needs /root/reference at build time)")
    os.makedirs(OUT, exist_ok=True)
    outs = []
    for exe, tag in ((cpu, "cpu"), (gpu, "cuda")):
        path = os.path.join(OUT, f"soapy_{tag}_{rate}.bin")
        r = subprocess.run([exe, str(rate), str(blocks), path], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0 and "DONE" in r.stdout and " OK" in r.stdout, r.stdout + r.stderr[-2000:]
        outs.append(np.fromfile(path, dtype=np.float32))
        os.remove(path)
    ref, got = outs
    d = {32000000: 0, 4000000: 3}[rate]
    n = (blocks >> d) * 32768 * 2
    assert ref.size >= n and got.size >= n
    ref, got = ref[:n], got[:n]
    skip = 2 * (32768 >> d)
    err = np.linalg.norm(got[skip:].astype(np.float64) - ref[skip:]) / np.linalg.norm(ref[skip:].astype(np.float64))
    peak = float(np.abs(ref[skip:]).max())
    assert peak > 0 and err <= 1e-5, err
    assert float(np.abs(got[skip:] - ref[skip:]).max()) <= 4e-4 * peak


@pytest.mark.gpu
def test_radiohandler_stability_and_rate_switch_scenarios():
    """unittest/stability_test.cpp:255-365 (20 start/stop cycles over all rates, continuity) and
    unittest/benchmark_test.cpp:601-671 (rate switching) restated at RadioHandlerClass level with cuda_r2iq as the operator
    (oracle/ref_scenarios_driver.cpp; the reference's RadioHandler / mixer / radio models compiled in place)."""
    exe = os.path.join(REF_BIN, "ref_radio_scenarios")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/ref_radio_scenarios not built (needs /root/reference at build time)")
    r = subprocess.run([exe, "15"], capture_output=True, text=True, timeout=900)
    lines = [l for l in r.stdout.splitlines() if l.startswith(("PASS", "FAIL", "ALL PASS", "  "))]
    print("\n".join(lines))
    assert r.returncode == 0 and "ALL PASS" in r.stdout, "\n".join(lines) + r.stderr[-2000:]
    assert sum(l.startswith("PASS") for l in lines) == 3


def test_cmake_switch_swaps_the_operator_sources(tmp_path):
    """extio_sddc_b200/host/r2iq_cuda.cmake, the build switch INTEGRATION.md describes: with -DSDDC_R2IQ_CUDA=ON the
    reference's operator sources leave the list and the CUDA ones enter it, the target gains the define, include paths
    and library; with the switch off nothing changes."""
    import shutil
    if not shutil.which("cmake"):
        pytest.skip("cmake not installed")
    capi.load()   # the switch refuses to configure without the built library
    src = os.path.join(ROOT, "tests", "host", "cmake_switch")

    def configure(*flags):
        build = tmp_path / ("b" + str(len(flags)))
        r = subprocess.run(["cmake", "-S", src, "-B", str(build), f"-DR2IQ_CUDA_ROOT={ROOT}", *flags],
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        return {k: v for k, v in (l[3:].split("=", 1) for l in r.stdout.splitlines() if l.startswith("-- ") and "_AFTER=" in l)}

    off = configure()
    assert "fft_mt_r2iq.cpp" in off["SRC_AFTER"] and "fft_mt_r2iq_cuda.cpp" not in off["SRC_AFTER"]
    assert "r2iq_cuda" not in off["LIBS_AFTER"]
    on = configure("-DSDDC_R2IQ_CUDA=ON", "-DSDDC_FFT_BACKEND_CUDA=ON")
    names = [os.path.basename(p) for p in on["SRC_AFTER"].split(";")]
    assert not [n for n in names if n.startswith("fft_mt_r2iq") and n != "fft_mt_r2iq_cuda.cpp"]
    assert {"fft_mt_r2iq_cuda.cpp", "cuda_r2iq.cpp", "cuda_fft_backend.cpp", "RadioHandler.cpp", "pf_mixer.cpp"} <= set(names)
    assert "fft_backend_fftw.cpp" not in names and "fft_benchmark.cpp" in names
    assert "R2IQ_USE_REFERENCE_HEADERS" in on["DEFS_AFTER"] and "r2iq_cuda" in on["LIBS_AFTER"]
