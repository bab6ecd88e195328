"""CPU suite, part 1: the oracle itself.

Pins the NumPy restatement (oracle/r2iq_oracle.py) against (a) the golden vectors generated from
the unmodified reference operator (tests/golden/make_golden.py) and (b), where the compiled
reference library is present, the reference run live; then restates the properties the
reference's own tests assert for this path.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN, case_input
from oracle import r2iq_oracle as orc
from oracle import ref_lib

GAIN = orc.RX888MK2_GAINFACTOR


def test_golden_cases_match_restatement(golden_cases):
    for c in golden_cases["cases"]:
        x, h = case_input(c, golden_cases["history"])
        got = orc.r2iq(x, c["d"], c["tunebin"], gain=c["gain"], rand=c["rand"], sideband=c["sideband"], history=h)
        assert got.shape == c["out"].shape, c["name"]
        err = orc.rel_l2(got, c["out"])
        assert err < 5e-7, (c["name"], err)
        assert np.abs(got - c["out"]).max() <= 1e-6 * orc.full_scale(c["gain"]), c["name"]


def test_golden_cases_8192_match_restatement(golden_cases_8192):
    """The restatement at FFTN_R_ADC 8192 (what the reference tree ships, Core/config.h:49) against vectors produced by
    the reference compiled at that size."""
    for c in golden_cases_8192["cases"]:
        x, h = case_input(c, golden_cases_8192["history"])
        got = orc.r2iq(x, c["d"], c["tunebin"], gain=c["gain"], rand=c["rand"], sideband=c["sideband"], history=h, fftn=8192)
        assert got.shape == c["out"].shape, c["name"]
        err = orc.rel_l2(got, c["out"])
        assert err < 5e-7, (c["name"], err)
        assert np.abs(got - c["out"]).max() <= 1e-6 * orc.full_scale(c["gain"]), c["name"]


def test_filter_design_matches_reference_golden():
    z = np.load(os.path.join(GOLDEN, "filters.npz"))
    for d in range(7):
        bw = np.float32(64.0) / np.float32(1 << d)
        taps = orc.kaiser_window(257, 120.0, np.float32(0.85) * bw / np.float32(128), np.float32(1.1) * bw / np.float32(128))
        # float32 libm sinf vs float64 sin rounded: at most 1 ulp apart on a few taps
        assert np.allclose(taps, z[f"taps_d{d}"], rtol=0, atol=2e-8), d
        assert orc.rel_l2(orc.design_filter(GAIN, d), z[f"hw_d{d}"]) < 2e-7, d


def test_filter_shape():
    # pass-band gain == gain*2048/FFTN * sum(taps) ~ gain; stop-band >= 100 dB down (Astop 120 dB design).
    # (d = 0 has no stop band: its cut-off 0.4875 of the Fs/2 complex rate is almost Nyquist.)
    for d in (1, 3):   # 257 taps cannot resolve the d >= 5 bands (fft_mt_r2iq.cpp:187 "@todo")
        H = np.abs(orc.design_filter(1.0, d).astype(np.complex128))
        mfft = 1024 >> d
        assert abs(H[0] - 1.0) < 1e-3
        # 257 taps at beta = 12.26 give a transition about 31 bins wide around the 0.4875*mfft cut-off
        edge = int(0.4875 * mfft) + 20
        stop = H[edge: 1024 - edge]
        assert stop.max() < 1e-5, (d, stop.max())


def test_set_freq_offset_golden():
    rows = np.load(os.path.join(GOLDEN, "freq_offset.npy"))
    for off, d, tb, res in rows:
        got_tb, got_res = orc.set_freq_offset(off, int(d))
        assert got_tb == int(tb)
        assert got_res == pytest.approx(res, abs=1e-9)
        assert got_tb % 4 == 0


def test_convert_golden_bit_exact():
    z = np.load(os.path.join(GOLDEN, "convert.npz"))
    assert np.array_equal(orc.convert_float(z["x"], False), z["plain"])
    assert np.array_equal(orc.convert_float(z["x"], True), z["rand"])
    # the randomizer is an involution
    assert np.array_equal(orc.derandomize(orc.randomize(z["x"]), True), z["x"])


@pytest.mark.skipif(not ref_lib.available(2048), reason="oracle/_ref not built")
@pytest.mark.parametrize("fftn", [2048, 8192])
def test_restatement_matches_live_reference(fftn):
    if not ref_lib.available(fftn):
        pytest.skip("library for this FFTN not built")
    R = ref_lib.get(fftn)
    R.set_fft_backend(ref_lib.BACKEND_DOUBLE)
    hf = fftn // 2
    hist = np.random.default_rng(3).integers(-3000, 3000, hf).astype(np.int16)
    for (d, tb, rnd, sb) in [(0, hf // 4, False, False), (1, 144 * fftn // 2048, True, False), (3, 80, False, True),
                             (5, hf - 52, True, True), (6, 0, True, False), (2, hf - 4, False, False)]:
        x = orc.rx888_blocks(3, d, tb, fftn, randomized=rnd, seed=11 + d)
        a = R.run(x, d, tb, gain=GAIN, rand=rnd, sideband=sb, history=hist)
        b = orc.r2iq(x, d, tb, gain=GAIN, rand=rnd, sideband=sb, fftn=fftn, history=hist)
        assert orc.rel_l2(a, b) < 5e-7, (fftn, d, tb)


@pytest.mark.skipif(not ref_lib.available(2048), reason="oracle/_ref not built")
def test_reference_isa_variants_agree():
    # dispatcher's pick, scalar, and the hand-written AVX2 worker produce the same stream
    R = ref_lib.get(2048)
    R.set_fft_backend(ref_lib.BACKEND_DOUBLE)
    x = orc.rx888_blocks(2, 1, 256, randomized=True)
    base = R.run(x, 1, 256, gain=GAIN, rand=True, variant="def")
    for v in ("dispatch", "avx2"):
        assert orc.rel_l2(R.run(x, 1, 256, gain=GAIN, rand=True, variant=v), base) < 1e-6, v


# ---- properties the reference's tests pin (unittest/core_test.cpp, signal_integrity_test.cpp) ----

@pytest.mark.parametrize("d", range(7))
def test_output_count_per_block(d):
    # core_test.cpp:165-167 -- 32768 complex per callback, i.e. 32768>>d per input block
    y = orc.r2iq(np.zeros(2 * 65536, dtype=np.int16), d, 256)
    assert y.shape[0] == 2 * (32768 >> d)


def test_zero_in_zero_out():
    # signal_integrity_test.cpp:591-595
    y = orc.r2iq(np.zeros(65536, dtype=np.int16), 0, 256, gain=GAIN)
    assert np.all(y == 0)


def test_tone_lands_where_tuned():
    # signal_integrity_test.cpp:307-359 -- 5 MHz sine at 64 Msps; tuned to 5 MHz it sits at DC of the
    # IQ stream with amplitude A*gain*1024
    tb, _ = orc.set_freq_offset(5e6 / 32e6, 0)
    x = orc.sine_blocks(2, 5e6, 64e6, 10000.0)
    y = orc.r2iq(x, 0, tb, gain=orc.BBRF103_GAINFACTOR).astype(np.complex128)[4096:]
    spec = np.abs(np.fft.fft(y[:16384]))
    assert spec.argmax() == 0
    # (int16 truncation of the generator costs about 5e-5 of amplitude)
    assert np.abs(y).mean() == pytest.approx(10000.0 * orc.BBRF103_GAINFACTOR * 1024.0, rel=2e-3)


def test_linearity_in_amplitude():
    # signal_integrity_test.cpp:688-751 asserts 1 < ratio < 4; the path is exactly linear
    x = (orc.rx888_blocks(2, 2, 256, randomized=False) // 4).astype(np.int16)
    a = orc.r2iq(x, 2, 256, gain=GAIN, dtype=np.complex128)
    b = orc.r2iq((2 * x).astype(np.int16), 2, 256, gain=GAIN, dtype=np.complex128)
    assert orc.rel_l2(b, 2 * a) < 1e-12


def test_block_boundary_continuity():
    # a stream processed in one call equals the same stream processed block by block with the
    # history carried (impl.hpp:32,42)
    x = orc.rx888_blocks(3, 1, 256, randomized=False)
    whole = orc.r2iq(x, 1, 256, gain=GAIN)
    parts = []
    hist = None
    for b in range(3):
        blk = x[b * 65536:(b + 1) * 65536]
        parts.append(orc.r2iq(blk, 1, 256, gain=GAIN, history=hist))
        hist = blk[-1024:]
    assert np.array_equal(np.concatenate(parts), whole)


def test_sideband_is_conjugate():
    x = orc.rx888_blocks(1, 3, 512, randomized=False)
    a = orc.r2iq(x, 3, 512, gain=GAIN)
    b = orc.r2iq(x, 3, 512, gain=GAIN, sideband=True)
    assert np.array_equal(b, np.conj(a))
