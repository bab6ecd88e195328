"""Generate tests/golden/ref_capture.wav with the REFERENCE's own writer (libsddc/wavewrite.c compiled in place
into oracle/_ref/libref_wave.so): a 16-bit one-channel ADC capture exactly as libsddc/sddc_stream_test.c:185-188
records it (waveWriteHeader -> waveWriteSamples -> waveFinalizeHeader), plus a two-channel one.
Run in the build container (needs /root/reference):  python tests/golden/make_golden_wav.py"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
so = os.path.join(ROOT, "oracle", "_ref", "libref_wave.so")
os.makedirs(os.path.dirname(so), exist_ok=True)
subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-I", os.path.join(REF, "libsddc"), "-o", so,
                os.path.join(REF, "libsddc", "wavewrite.c")], check=True)
lib = ctypes.CDLL(so)
libc = ctypes.CDLL(None)
libc.fopen.restype = ctypes.c_void_p
libc.fopen.argtypes = [ctypes.c_char_p, ctypes.c_char_p]
libc.fclose.argtypes = [ctypes.c_void_p]
lib.waveWriteHeader.argtypes = [ctypes.c_uint, ctypes.c_uint, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
lib.waveWriteSamples.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
lib.waveWriteFrames.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
lib.waveFinalizeHeader.argtypes = [ctypes.c_void_p]

rng = np.random.default_rng(2048)
for name, rate, freq, ch in (("ref_capture.wav", 128000000, 0, 1), ("ref_capture_iq.wav", 2000000, 7100000, 2)):
    x = rng.integers(-32768, 32767, 256 * ch, dtype=np.int16)
    path = os.path.join(HERE, name)
    f = libc.fopen(path.encode(), b"wb")
    lib.waveWriteHeader(rate, freq, 16, ch, f)
    if ch == 1:
        assert lib.waveWriteSamples(f, x.ctypes.data, x.size, 0) == 0
    else:
        assert lib.waveWriteFrames(f, x.ctypes.data, x.size // ch, 0) == 0
    assert lib.waveFinalizeHeader(f) == 0
    libc.fclose(f)
    np.save(os.path.join(HERE, name.replace(".wav", "_samples.npy")), x)
    print(path, os.path.getsize(path), "bytes")
