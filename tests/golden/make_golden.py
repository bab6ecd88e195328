"""Generate the golden fixtures in tests/golden/ from the UNMODIFIED reference operator.

Run in the build container (needs /root/reference):   python tests/golden/make_golden.py
The reference (Core/fft_mt_r2iq*.cpp + Core/fir.cpp, compiled in place by oracle/Makefile into
oracle/_ref/libref_r2iq_2048.so) is driven through oracle/ref_driver.cpp with the shim's float64
FFT backend, so the vectors do not depend on which FFT library a box happens to have.  The
reference's own tests hold no golden vectors for this path, which is why they
are produced from its code here.  The fixtures are small (about 1.5 MB) and are committed.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import r2iq_oracle as orc  # noqa: E402
from oracle import ref_lib  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))

# (name, d, tunebin, rand, sideband, gain, n_blocks, signal)
CASES = [
    ("hf64_sine5m_d0_default", 0, 256, False, False, orc.BBRF103_GAINFACTOR, 1, "sine5m"),
    ("hf64_sine5m_d0_tuned", 0, 160, False, False, orc.BBRF103_GAINFACTOR, 1, "sine5m"),
    ("hf64_tone_d0_default", 0, 256, False, False, orc.BBRF103_GAINFACTOR, 1, "tone64"),
    ("rx888_d1", 1, 256, True, False, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("rx888_d2", 2, 256, True, False, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("rx888_d3", 3, 256, True, False, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("rx888_d4", 4, 256, True, False, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("rx888_d5", 5, 256, True, False, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("rx888_d6_dc_edge", 6, 0, True, False, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("vhf_if_d2", 2, 144, False, False, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("vhf_if_d2_lsb", 2, 144, False, True, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("nyquist_edge_d3_lsb", 3, 1020, True, True, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("dc_edge_d5", 5, 4, True, False, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("low_tune_d1", 1, 80, False, False, orc.HF103_GAINFACTOR, 1, "comb"),
]


def comb_signal(n_blocks):
    """One shared wideband test input: 24 tones spread over the band (so every tune bin and every
    decimation sees in-band energy) plus uniform noise; stored once, clean (not randomized)."""
    rng = np.random.default_rng(888)
    n = np.arange(n_blocks * orc.TRANSFER_SAMPLES, dtype=np.float64)
    x = rng.uniform(-512.0, 512.0, n.shape[0])
    for i in range(24):
        f = (i + 0.37) / 48.5          # cycles/sample in (0, 0.5)
        x += 1200.0 * np.sin(2 * np.pi * f * n + 0.9 * i)
    return np.rint(x).astype(np.int16)


def make_signal(kind, n_blocks):
    if kind == "sine5m":
        return orc.sine_blocks(n_blocks, 5e6, 64e6)
    if kind == "tone64":
        return orc.tone_blocks(n_blocks)
    return comb_signal(n_blocks)


# FFTN_R_ADC = 8192, the value the reference tree ships with (Core/config.h:49): halfFft 4096, tune bins in [0, 4096)
CASES_8192 = [
    ("w_hf64_sine5m_d0", 0, 1024, False, False, orc.BBRF103_GAINFACTOR, 1, "sine5m"),
    ("w_rx888_d1", 1, 1024, True, False, orc.RX888MK2_GAINFACTOR, 1, "comb"),
    ("w_rx888_d2_lsb", 2, 576, True, True, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("w_rx888_d3", 3, 2000, False, False, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("w_rx888_d4", 4, 1024, True, False, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("w_rx888_d5_dc_edge", 5, 8, True, False, orc.RX888MK2_GAINFACTOR, 2, "comb"),
    ("w_rx888_d6_nyquist_edge_lsb", 6, 4092, True, True, orc.RX888MK2_GAINFACTOR, 2, "comb"),
]


def main_8192(inputs):
    """tests/golden/r2iq_cases_8192.npz: the same reference sources compiled with FFTN_R_ADC 8192
    (oracle/_ref/libref_r2iq_8192.so), on the inputs already stored in r2iq_cases.npz."""
    R = ref_lib.get(8192)
    assert R.set_fft_backend(ref_lib.BACKEND_DOUBLE) == 0
    hist = np.random.default_rng(8).integers(-2000, 2000, 4096).astype(np.int16)
    bundle = {"history": hist}
    names = []
    for (name, d, tb, rnd, sb, gain, nb, sig) in CASES_8192:
        x = inputs[sig][: nb * orc.TRANSFER_SAMPLES]
        if rnd:
            x = orc.randomize(x)
        use_hist = sig == "comb"
        h = orc.randomize(hist) if rnd else hist
        y = R.run(x, d, tb, gain=gain, rand=rnd, sideband=sb, history=h if use_hist else None)
        names.append(f"{name}:{sig}")
        bundle[f"{name}__out"] = y
        bundle[f"{name}__meta"] = np.array([d, tb, int(rnd), int(sb), nb, int(use_hist)], dtype=np.int64)
        bundle[f"{name}__gain"] = np.array([gain], dtype=np.float64)
        print(name, x.shape, y.shape, float(np.sqrt(np.mean(np.abs(y) ** 2))))
    bundle["cases"] = np.array(names)
    np.savez_compressed(os.path.join(OUT, "r2iq_cases_8192.npz"), **bundle)


def main():
    ref_lib.build()
    R = ref_lib.get(2048)
    assert R.set_fft_backend(ref_lib.BACKEND_DOUBLE) == 0
    hist = np.random.default_rng(7).integers(-2000, 2000, 1024).astype(np.int16)
    bundle = {"history": hist}
    inputs = {"sine5m": make_signal("sine5m", 1), "tone64": make_signal("tone64", 1), "comb": make_signal("comb", 2)}
    for k, v in inputs.items():
        bundle[f"input__{k}"] = v          # stored CLEAN; randomized cases apply orc.randomize() to it
    names = []
    for (name, d, tb, rnd, sb, gain, nb, sig) in CASES:
        x = inputs[sig][: nb * orc.TRANSFER_SAMPLES]
        if rnd:
            x = orc.randomize(x)
        use_hist = sig == "comb"
        h = orc.randomize(hist) if rnd else hist
        y = R.run(x, d, tb, gain=gain, rand=rnd, sideband=sb, history=h if use_hist else None)
        names.append(f"{name}:{sig}")
        bundle[f"{name}__out"] = y
        bundle[f"{name}__meta"] = np.array([d, tb, int(rnd), int(sb), nb, int(use_hist)], dtype=np.int64)
        bundle[f"{name}__gain"] = np.array([gain], dtype=np.float64)
        print(name, x.shape, y.shape, float(np.sqrt(np.mean(np.abs(y) ** 2))))
    bundle["cases"] = np.array(names)
    np.savez_compressed(os.path.join(OUT, "r2iq_cases.npz"), **bundle)
    main_8192(inputs)

    # filter design: taps and spectra for every decimation (Core/fir.cpp, fft_mt_r2iq.cpp:181-205)
    filt = {}
    for d in range(7):
        bw = np.float32(64.0) / np.float32(1 << d)
        filt[f"taps_d{d}"] = R.kaiser(257, 120.0, np.float32(0.85) * bw / np.float32(128), np.float32(1.1) * bw / np.float32(128))
        filt[f"hw_d{d}"] = R.filter_hw(orc.RX888MK2_GAINFACTOR, d)
    np.savez_compressed(os.path.join(OUT, "filters.npz"), **filt)

    # setFreqOffset (Core/fft_mt_r2iq.cpp:101-109): the TuneTest grid of unittest/core_test.cpp:186
    # (1 kHz .. 15 MHz step 377 kHz at 64 Msps) plus edge points
    offs = [f / 32e6 for f in np.arange(1e3, 15e6, 377e3)] + [0.0, 0.004, 0.1428125, 0.25, 0.5, 0.97, 0.999]
    rows = []
    for d in (0, 2, 5):
        for o in offs:
            tb, res = R.set_freq_offset(o, d)
            rows.append((o, d, tb, res))
    np.save(os.path.join(OUT, "freq_offset.npy"), np.array(rows, dtype=np.float64))

    # convert_float<rand>: the i*13-512 ramp of unittest/neon_test.cpp:234-261 plus extremes
    x = np.concatenate([(np.arange(8192) * 13 - 512).astype(np.int16),
                        np.array([-32768, -32767, -2, -1, 0, 1, 2, 32766, 32767], dtype=np.int16),
                        np.zeros(7, dtype=np.int16)])
    np.savez_compressed(os.path.join(OUT, "convert.npz"), x=x, plain=R.convert(x, False), rand=R.convert(x, True))
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
