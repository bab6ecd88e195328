"""GPU suite (-m gpu): the CUDA path, called through the C ABI, against the oracle.

Bars (BASELINE.json north_star): bit-exact through de-randomise/convert; output IQ within
relative L2 <= 1e-5 and max-abs <= 1e-4 of full scale (32768*gain*1024) of the reference.
"""
import ctypes
import os

import numpy as np
import pytest

from conftest import GOLDEN, case_input
from extio_sddc_b200 import CudaR2IQ, PinnedBuffer, capi
from oracle import r2iq_oracle as orc

pytestmark = pytest.mark.gpu

GAIN = orc.RX888MK2_GAINFACTOR
REL_L2_TOL = 1e-5          # north_star: relative L2 error <= 1e-5
MAX_ABS_TOL = 1e-4         # north_star: max abs <= 1e-4 of full scale


def check(got, ref, gain, what):
    e = orc.rel_l2(got, ref)
    m = float(np.abs(got.astype(np.complex128) - ref.astype(np.complex128)).max()) / orc.full_scale(gain)
    assert e <= REL_L2_TOL, (what, "rel_l2", e)
    assert m <= MAX_ABS_TOL, (what, "max_abs/full_scale", m)


@pytest.fixture(scope="module")
def op():
    r = CudaR2IQ()
    r.Init(GAIN)
    yield r
    r.close()


def run(op, x, d, tb, rand=False, lsb=False, history=None):
    op.setDecimate(d)
    op.updateRand(rand)
    op.setSideband(lsb)
    op.tunebin = tb
    op.TurnOn()
    y = op.process(x, history=history)
    op.TurnOff()
    return y


def test_convert_bit_exact(op):
    z = np.load(os.path.join(GOLDEN, "convert.npz"))
    x = z["x"]
    if x.shape[0] % 2:
        x = x[:-1]
    for rnd, key in ((False, "plain"), (True, "rand")):
        got = op.convert_float(x, rnd)
        assert np.array_equal(got, z[key][: x.shape[0]]), key
    # every int16 value, both modes
    allv = np.arange(-32768, 32768, dtype=np.int32).astype(np.int16)
    for rnd in (False, True):
        assert np.array_equal(op.convert_float(allv, rnd), orc.convert_float(allv, rnd))


def test_golden_cases(golden_cases):
    for c in golden_cases["cases"]:
        r = CudaR2IQ()
        r.Init(c["gain"])
        x, h = case_input(c, golden_cases["history"])
        got = run(r, x, c["d"], c["tunebin"], c["rand"], c["sideband"], history=h)
        r.close()
        assert got.shape == c["out"].shape
        check(got, c["out"], c["gain"], c["name"])


def test_golden_cases_fftn8192(golden_cases_8192):
    """The FFTN 8192 kernels against vectors from the reference compiled at 8192 (tests/golden/r2iq_cases_8192.npz)."""
    for c in golden_cases_8192["cases"]:
        r = CudaR2IQ(fftn=8192)
        r.Init(c["gain"])
        x, h = case_input(c, golden_cases_8192["history"])
        got = run(r, x, c["d"], c["tunebin"], c["rand"], c["sideband"], history=h)
        r.close()
        assert got.shape == c["out"].shape
        check(got, c["out"], c["gain"], c["name"])


@pytest.mark.parametrize("d", range(7))
def test_parity_matrix(op, d):
    # config 2 signal (rand on) at the default tune bin, plus edge tune bins with zero-filled bands
    for tb, rnd, lsb in ((256, True, False), (0, False, False), (1020, True, True), (512, False, True),
                         (4, True, False), (1024 - (1024 >> d) // 2, False, False), ((1024 >> d) // 2, True, False)):
        tb = max(0, min(1020, tb)) & ~3
        x = orc.rx888_blocks(3, d, tb, randomized=rnd, seed=1000 + 7 * d + tb)
        hist = np.random.default_rng(d).integers(-3000, 3000, 1024).astype(np.int16)
        got = run(op, x, d, tb, rnd, lsb, history=hist)
        ref = orc.r2iq(x, d, tb, gain=GAIN, rand=rnd, sideband=lsb, history=hist)
        check(got, ref, GAIN, (d, tb, rnd, lsb))


@pytest.mark.parametrize("d", [4, 5, 6])
def test_every_tune_bin_at_small_mfft(op, d):
    """d = 4, 5 run their own code paths (r2iq_pair.cu: the backward transform across the lanes by shuffle butterflies; at
    d = 5 band and mirror bins evaluated directly from the first forward pass with per-tune-bin coefficient tables) whose
    lane arithmetic depends on the tune bin modulo 32 and on where the band wraps: every tune bin setFreqOffset can produce
    (multiples of 4, Core/fft_mt_r2iq.cpp:101-109), one block with history, randomizer and sideband alternating."""
    rng = np.random.default_rng(50 + d)
    x = rng.integers(-20000, 20000, 2 * orc.TRANSFER_SAMPLES).astype(np.int16).reshape(2, -1)
    hist = rng.integers(-3000, 3000, 1024).astype(np.int16)
    for tb in range(0, 1024, 4):
        rand, lsb = bool((tb >> 2) & 1), bool((tb >> 3) & 1)
        got = run(op, x, d, tb, rand=rand, lsb=lsb, history=hist)
        ref = orc.r2iq(x, d, tb, gain=GAIN, rand=rand, sideband=lsb, history=hist)
        check(got, ref, GAIN, f"d={d} tunebin={tb}")


def test_config1_hf_64msps(op):
    # BASELINE config 1: synthetic tone 64 Msps, decimation 2 (d=0), default tune and tuned to 5 MHz
    r = CudaR2IQ()
    r.Init(orc.BBRF103_GAINFACTOR)
    x = orc.tone_blocks(4)
    check(run(r, x, 0, 256), orc.r2iq(x, 0, 256, gain=orc.BBRF103_GAINFACTOR), orc.BBRF103_GAINFACTOR, "tone")
    x = orc.sine_blocks(4, 5e6, 64e6)
    r.setDecimate(0)
    res = r.setFreqOffset(5e6 / 32e6)
    tb, want_res = orc.set_freq_offset(5e6 / 32e6, 0)
    assert r.tunebin == tb and res == pytest.approx(want_res, abs=1e-9)
    r.TurnOn()
    y = r.process(x)
    r.TurnOff()
    check(y, orc.r2iq(x, 0, tb, gain=orc.BBRF103_GAINFACTOR), orc.BBRF103_GAINFACTOR, "sine5m")
    r.close()


def test_config3_vhf_tune_sweep(op):
    # BASELINE config 3: R828D IF point 4.57 MHz (no inversion), then a retune sweep between batches
    # over the TuneTest grid of unittest/core_test.cpp:186 plus band-edge points, d = 2 and d = 5
    offsets = [f / 32e6 for f in np.arange(1e3, 15e6, 1508e3)] + [4.57e6 / 32e6, 0.0, 0.004, 0.97, 0.999]
    for d in (2, 5):
        op.setDecimate(d)
        op.updateRand(False)
        op.setSideband(False)
        op.TurnOn()
        hist = None
        for i, off in enumerate(offsets):
            res = op.setFreqOffset(off)            # retune while "streaming"
            tb, want = orc.set_freq_offset(off, d)
            assert op.tunebin == tb
            assert res == pytest.approx(want, abs=1e-9)
            x = orc.rx888_blocks(1, d, tb, randomized=False, seed=50 + i)
            got = op.process(x, history=hist if hist is not None else np.zeros(1024, np.int16))
            ref = orc.r2iq(x, d, tb, gain=GAIN, history=hist)
            check(got, ref, GAIN, ("sweep", d, off))
            hist = x[-1024:]
        op.TurnOff()
    # one sideband = true case to cover copy<true>
    x = orc.rx888_blocks(2, 2, 144, randomized=False)
    check(run(op, x, 2, 144, lsb=True), orc.r2iq(x, 2, 144, gain=GAIN, sideband=True), GAIN, "lsb")


def test_streaming_carries_history(op):
    # one call over N blocks == block-by-block calls with the state carried by the handle
    x = orc.rx888_blocks(5, 1, 256, randomized=True)
    whole = run(op, x, 1, 256, rand=True)
    op.setDecimate(1); op.updateRand(True); op.setSideband(False); op.tunebin = 256
    op.TurnOn()
    parts = [op.process(x[b * 65536:(b + 1) * 65536]) for b in range(5)]
    op.TurnOff()
    assert np.array_equal(np.concatenate(parts), whole)
    # TurnOn resets the history to zeros
    again = run(op, x, 1, 256, rand=True)
    assert np.array_equal(again, whole)


def test_chunked_pipeline_and_ragged_sizes():
    # chunk size that does not divide the block count; pinned and pageable host buffers
    r = CudaR2IQ(chunk_blocks=4)
    r.Init(GAIN)
    for nb in (1, 3, 4, 5, 11):
        for d in (0, 3, 6):
            x = orc.rx888_blocks(nb, d, 256, randomized=True, seed=nb * 10 + d)
            pin_in = PinnedBuffer(x.shape[0], np.int16)
            pin_in.array[:] = x
            pin_out = PinnedBuffer(nb * (32768 >> d), np.complex64)
            got = run(r, pin_in.array, d, 256, rand=True)
            r.setDecimate(d); r.TurnOn()
            got2 = r.process(pin_in.array, out=pin_out.array).copy()
            r.TurnOff()
            ref = orc.r2iq(x, d, 256, gain=GAIN, rand=True)
            check(got, ref, GAIN, ("chunk", nb, d))
            assert np.array_equal(got, got2)
            pin_in.free(); pin_out.free()
    st = r.stats()
    assert st.kernel_launches > 0 and st.h2d_bytes > 0 and st.d2h_bytes > 0 and st.cc_major == 10
    r.close()


def test_empty_and_error_behaviour():
    r = CudaR2IQ()
    r.Init(GAIN)
    with pytest.raises(capi.R2iqError) as ei:          # process before TurnOn
        r.process(np.zeros(65536, np.int16))
    assert ei.value.code == capi.E_STATE
    r.TurnOn()
    assert r.process(np.zeros(0, np.int16)).shape == (0,)
    with pytest.raises(ValueError):
        r.process(np.zeros(1000, np.int16))
    with pytest.raises(capi.R2iqError):
        r.tunebin = 255                                  # not a multiple of 4
    with pytest.raises(capi.R2iqError):
        r.setDecimate(7)
    assert r.IsOn()
    r.TurnOff()
    assert not r.IsOn()
    assert r.getRatio() == 1
    r.setDecimate(3)
    assert r.getRatio() == 8
    # decimation is latched at TurnOn like the reference worker (impl.hpp:3-6)
    r.TurnOn(); r.setDecimate(1)
    assert r.out_per_block() == 32768 >> 3
    # the kernels index frames with 32 bits: a launch of 2^31 frames or more (6.5 TB of samples) is refused before any
    # memory is touched (the pointers below are never dereferenced)
    with pytest.raises(capi.R2iqError) as ei:
        r.process_device(0x1000, 50_000_000, 0x1000, 0)
    assert ei.value.code == capi.E_INVALID
    r.TurnOff()
    r.close()


def test_zero_and_dc_input(op):
    y = run(op, np.zeros(2 * 65536, np.int16), 0, 256)
    assert np.all(y == 0)
    # DC input tuned away from DC: nothing in band (signal_integrity_test.cpp:367-407 analogue)
    x = np.full(2 * 65536, 1000, np.int16)
    y = run(op, x, 2, 512)
    assert np.abs(y[4096:]).max() < 1e-5 * 1000 * GAIN * 1024


def test_custom_filter_upload(op):
    # host-designed filter uploaded through set_filter (north star: "uploaded once per setDecimate/tune")
    r = CudaR2IQ()
    r.Init(GAIN)
    H = orc.design_filter(GAIN, 2) * np.complex64(0.5)
    r.set_filter_hw(2, H)
    assert np.array_equal(r.filter_hw(2), H)
    x = orc.rx888_blocks(2, 2, 256, randomized=False)
    got = run(r, x, 2, 256)
    ref = orc.r2iq(x, 2, 256, gain=GAIN, filter_hw=H)
    check(got, ref, GAIN * 0.5, "custom filter")
    r.close()


def test_device_resident_path_matches_host_path(op):
    import torch
    x = orc.rx888_blocks(6, 0, 256, randomized=True)
    host = run(op, x, 0, 256, rand=True)
    dev = torch.device("cuda:0")
    buf = torch.zeros(1024 + x.shape[0], dtype=torch.int16, device=dev)
    buf[1024:] = torch.from_numpy(x).to(dev)
    out = torch.empty(6 * 32768, dtype=torch.complex64, device=dev)
    op.setDecimate(0); op.updateRand(True); op.setSideband(False); op.tunebin = 256
    op.TurnOn()
    op.process_device(buf.data_ptr(), 6, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    op.TurnOff()
    assert np.array_equal(out.cpu().numpy(), host)


@pytest.mark.parametrize("d", [0, 5])
def test_full_size_properties(d):
    """BASELINE config 2 size (8192 blocks, 537 M samples) through size-independent properties:
    (a) time-slab split with halo == whole capture (what multi-GPU sharding relies on),
    (b) linearity: r2iq(2x) == 2*r2iq(x) to float rounding,
    (c) a sampled window agrees with the oracle."""
    import torch
    n_blocks = 8192
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev)
    g.manual_seed(888 + d)
    x = torch.randint(-4000, 4000, (1024 + n_blocks * 65536,), dtype=torch.int16, device=dev, generator=g)
    x[:1024] = 0
    opb = 32768 >> d
    r = CudaR2IQ()
    r.Init(GAIN)
    r.setDecimate(d); r.updateRand(False); r.tunebin = 256
    r.TurnOn()
    s = torch.cuda.current_stream().cuda_stream
    whole = torch.empty(n_blocks * opb, dtype=torch.complex64, device=dev)
    r.process_device(x.data_ptr(), n_blocks, whole.data_ptr(), s)
    # (a) two slabs; the second starts at a block boundary and reads its halo from the first
    half = n_blocks // 2
    parts = torch.empty_like(whole)
    r.process_device(x.data_ptr(), half, parts.data_ptr(), s)
    r.process_device(x.data_ptr() + 2 * half * 65536, n_blocks - half, parts.data_ptr() + 8 * half * opb, s)
    torch.cuda.synchronize()
    assert torch.equal(whole, parts)
    # (b)
    x2 = x * 2
    dbl = torch.empty_like(whole)
    r.process_device(x2.data_ptr(), n_blocks, dbl.data_ptr(), s)
    torch.cuda.synchronize()
    err = (torch.linalg.vector_norm(dbl - 2 * whole) / torch.linalg.vector_norm(dbl)).item()
    assert err < 1e-6, err
    # (c) blocks [4000, 4003) against the oracle
    b0 = 4000
    xin = x[1024 + b0 * 65536: 1024 + (b0 + 3) * 65536].cpu().numpy()
    hist = x[b0 * 65536: 1024 + b0 * 65536].cpu().numpy()
    ref = orc.r2iq(xin, d, 256, gain=GAIN, history=hist)
    check(whole[b0 * opb:(b0 + 3) * opb].cpu().numpy(), ref, GAIN, ("full-size window", d))
    r.TurnOff()
    r.close()


def test_config4_channelizer(op):
    """BASELINE config 4: many tuned DDC channels from one stream.  Every channel must equal an
    independent reference run tuned to that bin (64 channels at d = 5; subsets at d = 1 ... 4 and 6;
    edge channels whose bands leave [0, halfFft) included)."""
    from extio_sddc_b200 import shard
    x = orc.rx888_blocks(2, 5, 256, randomized=True, seed=4)
    hist = np.random.default_rng(4).integers(-2000, 2000, 1024).astype(np.int16)
    for d, bins in ((5, shard.channel_tunebins(64)), (3, [0, 4, 60, 256, 1020] + shard.channel_tunebins(64)[::8]),
                    (1, [0, 256, 768, 1020]), (6, shard.channel_tunebins(64) + [0, 1020] * 3),
                    (2, [0, 4, 124, 128, 132, 256, 900, 1020]), (4, [0, 4, 28, 32, 36, 60, 256, 1000, 1016, 1020] + shard.channel_tunebins(64)[::4])):
        op.setDecimate(d); op.updateRand(True); op.setSideband(False)
        op.TurnOn()
        got = op.process_channels(x, bins, history=hist)
        op.TurnOff()
        assert got.shape == (len(bins), 2 * (32768 >> d))
        for c, tb in enumerate(bins):
            ref = orc.r2iq(x, d, tb, gain=GAIN, rand=True, history=hist)
            check(got[c], ref, GAIN, ("channelizer", d, tb))
    # odd frame total (3 blocks = 129 frames: the last pair has one frame), lower sideband, more channels than one pass
    x3 = orc.rx888_blocks(3, 5, 256, randomized=False, seed=14)
    bins = shard.channel_tunebins(64) + [0, 4, 1016, 1020, 512, 500, 12]
    for d in (5, 6, 4, 3, 2):
        op.setDecimate(d); op.updateRand(False); op.setSideband(True)
        op.TurnOn()
        got = op.process_channels(x3, bins, history=hist)
        op.TurnOff()
        for c, tb in enumerate(bins):
            check(got[c], orc.r2iq(x3, d, tb, gain=GAIN, rand=False, sideband=True, history=hist), GAIN, ("channelizer lsb", d, tb))
    op.setSideband(False)
    # must agree with the single-channel kernels too
    op.setDecimate(5); op.tunebin = 264
    single = run(op, x, 5, 264, rand=True, history=hist)
    op.setDecimate(5); op.TurnOn()
    multi = op.process_channels(x, [264], history=hist)[0]
    op.TurnOff()
    assert orc.rel_l2(multi, single) < 1e-6
    # d = 0: every channel is a full-rate operator (one launch per channel, no shared transform)
    op.setDecimate(0); op.updateRand(True); op.TurnOn()
    got = op.process_channels(x, [256, 0, 1020, 512], history=hist)
    op.TurnOff()
    for c, tb in enumerate([256, 0, 1020, 512]):
        check(got[c], orc.r2iq(x, 0, tb, gain=GAIN, rand=True, history=hist), GAIN, ("channelizer", 0, tb))
    # a tune bin that is not a multiple of 4 is refused
    op.setDecimate(5); op.TurnOn()
    with pytest.raises(capi.R2iqError):
        op.process_channels(x, [258])
    op.TurnOff()


def test_config4_channelizer_fftn8192():
    """The channelizer on the FFTN the reference tree ships (Core/config.h:49): every channel equals an independent
    reference run at 8192 tuned to that bin (d = 5, 4, 3 and 0; edge bins included)."""
    from extio_sddc_b200 import CudaR2IQ
    op8 = CudaR2IQ(device=0, fftn=8192)
    op8.Init(GAIN)
    x = orc.rx888_blocks(2, 5, 1024, fftn=8192, randomized=True, seed=41)
    hist = np.random.default_rng(41).integers(-2000, 2000, 4096).astype(np.int16)
    for d, bins in ((5, [0, 64, 1024, 2048, 3000, 4092, 4, 60, 68, 124, 128, 4032, 4088] + list(range(100, 4096, 331 * 4))[:8]),
                    (4, [0, 4, 124, 128, 132, 252, 256, 2048, 3968, 4092]), (3, [4, 1024, 4000]), (0, [1024, 0])):
        op8.setDecimate(d); op8.updateRand(True); op8.setSideband(False)
        op8.TurnOn()
        got = op8.process_channels(x, bins, history=hist)
        op8.TurnOff()
        assert got.shape == (len(bins), 2 * (32768 >> d))
        for c, tb in enumerate(bins):
            ref = orc.r2iq(x, d, tb, gain=GAIN, rand=True, history=hist, fftn=8192)
            check(got[c], ref, GAIN, ("channelizer 8192", d, tb))
    # d = 4, 5 run the shared-forward kernel (r2iq_wide_channels.cu): odd frame total (3 blocks = 33 frames), lower sideband,
    # randomizer off, and agreement with the single-channel kernel
    x3 = orc.rx888_blocks(3, 5, 1024, fftn=8192, randomized=False, seed=43)
    bins = [0, 4, 1000, 2044, 2048, 2052, 4088, 4092]
    for d in (5, 4):
        op8.setDecimate(d); op8.updateRand(False); op8.setSideband(True)
        op8.TurnOn()
        got = op8.process_channels(x3, bins, history=hist)
        single = op8.process_channels(x3, [1000], history=hist)[0]      # one channel: the single-channel launch
        op8.TurnOff()
        for c, tb in enumerate(bins):
            check(got[c], orc.r2iq(x3, d, tb, gain=GAIN, rand=False, sideband=True, history=hist, fftn=8192), GAIN, ("channelizer 8192 lsb", d, tb))
        assert orc.rel_l2(got[2], single) < 1e-6
    op8.close()


def test_channelizer_more_than_256_channels(op):
    """Channel lists longer than the 256 tune bins the channelizer kernels keep in shared memory are read from global memory
    (the `in_smem == false` branch of r2iq_channels_lane_kernel, r2iq_channels_pair_kernel, r2iq_wide_channels_kernel): the
    result must equal, bit for bit, the same channels computed in two shorter lists, and the oracle on a few of them."""
    from extio_sddc_b200 import CudaR2IQ
    x = orc.rx888_blocks(1, 5, 256, randomized=True, seed=77)
    hist = np.random.default_rng(77).integers(-2000, 2000, 1024).astype(np.int16)
    bins = [(12 * i) % 1024 for i in range(261)]
    for d in (5, 3):
        op.setDecimate(d); op.updateRand(True); op.setSideband(False)
        op.TurnOn()
        got = op.process_channels(x, bins, history=hist)
        parts = np.concatenate([op.process_channels(x, bins[:130], history=hist), op.process_channels(x, bins[130:], history=hist)])
        op.TurnOff()
        assert got.shape == (261, 32768 >> d)                  # one block of complex64 per channel
        assert np.array_equal(got, parts)
        for c in (0, 129, 200, 260):
            check(got[c], orc.r2iq(x, d, bins[c], gain=GAIN, rand=True, history=hist), GAIN, ("long channel list", d, bins[c]))
    op8 = CudaR2IQ(device=0, fftn=8192)
    op8.Init(GAIN)
    x8 = orc.rx888_blocks(1, 5, 1024, fftn=8192, randomized=True, seed=78)
    hist8 = np.random.default_rng(78).integers(-2000, 2000, 4096).astype(np.int16)
    bins8 = [(44 * i) % 4096 for i in range(259)]
    op8.setDecimate(5); op8.updateRand(True); op8.setSideband(False)
    op8.TurnOn()
    got = op8.process_channels(x8, bins8, history=hist8)
    parts = np.concatenate([op8.process_channels(x8, bins8[:100], history=hist8), op8.process_channels(x8, bins8[100:], history=hist8)])
    few = [op8.process_channels(x8, bins8[:n], history=hist8) for n in (2, 3, 5)]     # fewer channels than warps in a team; a ragged last trip
    op8.TurnOff()
    assert np.array_equal(got, parts)
    for f in few:
        assert np.array_equal(f, got[:len(f)])
    for c in (0, 99, 258):
        check(got[c], orc.r2iq(x8, 5, bins8[c], gain=GAIN, rand=True, history=hist8, fftn=8192), GAIN, ("long channel list 8192", bins8[c]))
    op8.close()


def test_fine_tune_mixer_bit_exact(op):
    """Next row: the residual NCO mixer RadioHandlerClass applies after the operator, run on
    the device (mode R2IQ_NCO_REFERENCE, the default).  Applied to the same IQ, the device result must equal the
    reference's own mixer (Core/pffft/pf_mixer.cpp compiled in place) BIT FOR BIT, with the phase carried across calls."""
    from oracle import ref_lib
    if not ref_lib.mixer_available():
        pytest.skip("oracle/_ref/libref_mixer.so not built")
    x = orc.rx888_blocks(6, 1, 256, randomized=True, seed=77)
    for d, off in ((0, 0.2503), (1, 0.1431), (4, 0.6002)):
        op.setDecimate(d); op.updateRand(True); op.setSideband(False)
        fc = op.setFreqOffset(off)                    # residual after tune-bin quantisation
        assert fc != 0.0
        op.setFineTune(0.0)
        op.TurnOn()
        plain = op.process(x)
        op.TurnOff()
        assert op.setFineTune(fc) == pytest.approx(fc)
        op.TurnOn()
        mixed = np.concatenate([op.process(x[: 2 * 65536]), op.process(x[2 * 65536:])])   # phase carries over
        op.TurnOff()
        op.setFineTune(0.0)
        want = ref_lib.mixer_apply(plain, fc, slot=32768)
        assert np.array_equal(mixed.view(np.uint32), want.view(np.uint32)), (d, off)
        assert not np.array_equal(mixed, plain)
    op.tunebin = 256


def test_fine_tune_mixer_direct_mode(op):
    """R2IQ_NCO_DIRECT (include/r2iq_cuda.h): sample n since the mixer was set is multiplied by exp(2 pi j fc n), phase
    reduced in float64 on the device, inside the fused kernel's store epilogue.  Stated tolerance: rel-L2 <= 1e-6 and
    max-abs <= 1e-5 of full scale against the float64 ideal applied to the unmixed float32 stream -- at the start of the
    stream and after 2.7e8 output samples (phase carried over 65 calls); every decimation, sideband on and off, both
    FFTN families (both mix in the store epilogue of their fused kernels).  The reference's own mixer is further from that ideal than this one:
    its phase increment and its 128-sample step are float32 values, so its frequency is off by a few 1e-8 relative and
    the two streams drift apart linearly (measured 1.2e-4 rel-L2 over 2e5 samples at d = 0) -- checked loosely."""
    from extio_sddc_b200 import capi as _capi
    from oracle import ref_lib
    fs = orc.full_scale(GAIN)

    def ideal(plain, fc, n0):
        n = np.arange(n0, n0 + plain.size, dtype=np.float64)
        ph = np.float64(np.float32(fc)) * n
        ph -= np.floor(ph)
        return plain.astype(np.complex128) * np.exp(2j * np.pi * ph)

    x = orc.rx888_blocks(6, 1, 256, randomized=True, seed=78)
    for fftn in (2048, 8192):
        r = op if fftn == 2048 else CudaR2IQ(fftn=8192)
        if fftn == 8192:
            r.Init(GAIN)
        for d, off, sb in ((0, 0.2503, False), (1, 0.1431, True), (2, 0.3337, False), (4, 0.6002, False), (5, 0.0107, True), (6, 0.4411, False)):
            r.setDecimate(d); r.updateRand(True); r.setSideband(sb)
            fc = r.setFreqOffset(off)
            assert fc != 0.0
            r.setFineTune(0.0)
            r.TurnOn()
            plain = r.process(x)
            r.TurnOff()
            r.setFineTuneMode(_capi.NCO_DIRECT)
            assert r.setFineTune(fc) == pytest.approx(fc)
            r.TurnOn()
            mixed = np.concatenate([r.process(x[: 2 * 65536]), r.process(x[2 * 65536:])])   # phase carries over
            r.TurnOff()
            want = ideal(plain, fc, 0)
            err = np.linalg.norm(mixed - want) / np.linalg.norm(want)
            mx = np.abs(mixed - want).max() / fs
            assert err <= 1e-6 and mx <= 1e-5, (fftn, d, err, mx)
            assert not np.array_equal(mixed, plain)
            if ref_lib.mixer_available() and fftn == 2048:
                cpu = ref_lib.mixer_apply(plain, fc, slot=32768)
                e2 = np.linalg.norm(mixed - cpu) / np.linalg.norm(cpu)
                assert e2 <= 1e-3, (d, e2)          # the reference's float32 phase increment, not ours, carries this error
            r.setFineTune(0.0)
            r.setFineTuneMode(_capi.NCO_REFERENCE)
        if fftn == 8192:
            r.close()
    # long run at d = 0: 64 calls of 128 blocks (2.7e8 output samples) with the mixer on, then one checked call
    op.setDecimate(0); op.updateRand(True); op.setSideband(False)
    fc = op.setFreqOffset(0.2503)
    big = np.tile(x[: 4 * 65536], 32)
    op.setFineTune(0.0)
    op.TurnOn(); op.process(big); plain = op.process(x); op.TurnOff()   # same history in front of the checked call
    op.setFineTuneMode(_capi.NCO_DIRECT)
    op.setFineTune(fc)
    op.TurnOn()
    for _ in range(64):
        op.process(big)
    mixed = op.process(x)
    op.TurnOff()
    n0 = 64 * 128 * 32768
    want = ideal(plain, fc, n0)
    err = np.linalg.norm(mixed - want) / np.linalg.norm(want)
    assert err <= 1e-6 and np.abs(mixed - want).max() / fs <= 1e-5, err
    with pytest.raises(Exception):
        op.setFineTuneMode(_capi.NCO_REFERENCE)          # only while the mixer is off
    op.setFineTune(0.0)
    op.setFineTuneMode(_capi.NCO_REFERENCE)
    op.tunebin = 256


def test_async_submissions_equal_blocking_call():
    """r2iq_cuda_process_async / wait_input / wait / read / release: submissions in flight (more than there are slots),
    results delivered to host buffers or held on the device and fetched in pieces, all bit-identical to one blocking
    r2iq_cuda_process over the same stream; history is carried at submit time."""
    r = CudaR2IQ(chunk_blocks=4)
    r.Init(GAIN)
    r.updateRand(True); r.tunebin = 148
    x = orc.rx888_blocks(26, 2, 148, randomized=True, seed=19).reshape(-1)
    for d in (0, 3, 6):
        opb = 32768 >> d
        r.setDecimate(d); r.TurnOn()
        want = r.process(x)
        r.TurnOff()
        r.TurnOn()
        got = np.zeros_like(want)
        sizes = [4, 1, 3, 4, 2, 4, 4, 1, 3]          # 26 blocks, nine submissions, four slots
        tickets, pos = [], 0
        for i, nb in enumerate(sizes):
            blk = x[pos * 65536:(pos + nb) * 65536]
            if i % 2 == 0:                             # host output
                t = r.process_async(blk, out=got[pos * opb:(pos + nb) * opb])
                tickets.append((t, None, pos, nb))
            else:                                      # held on the device
                t = r.process_async(blk)
                r.wait_input(t)
                half = nb * opb // 2
                r.read(t, half, got[pos * opb + half:(pos + nb) * opb])     # second half first
                r.read(t, 0, got[pos * opb:pos * opb + half])
                r.release(t)
            pos += nb
        for t, _, _, _ in tickets:
            r.wait(t)
        r.TurnOff()
        assert np.array_equal(got, want), d
    r.setDecimate(1); r.TurnOn()
    with pytest.raises(Exception):
        r.process_async(x[: 5 * 65536])                # more than chunk_blocks
    r.TurnOff()
    r.close()


def test_process_rates_equals_one_operator_per_rate():
    """r2iq_cuda_process_rates: one upload, every decimation's stream bit-identical to r2iq_cuda_process at that
    decimation; ragged chunking, carried history across calls, latched decimation untouched."""
    r = CudaR2IQ(chunk_blocks=3)
    r.Init(GAIN)
    r.updateRand(True); r.tunebin = 148
    x = orc.rx888_blocks(8, 2, 148, randomized=True, seed=9)
    want = {}
    for d in range(7):
        r.setDecimate(d); r.TurnOn()
        want[d] = np.concatenate([r.process(x[: 5 * 65536]), r.process(x[5 * 65536:])])   # history carried
        r.TurnOff()
    r.setDecimate(3); r.TurnOn()
    decs = [0, 5, 2, 6, 1, 4, 3]
    a = r.process_rates(x[: 5 * 65536], decs)
    b = r.process_rates(x[5 * 65536:], decs)
    for i, d in enumerate(decs):
        assert np.array_equal(np.concatenate([a[i], b[i]]), want[d]), d
    assert r.getDecimate() == 3 if hasattr(r, "getDecimate") else True
    got3 = r.process(x, history=np.zeros(1024, np.int16))
    assert np.array_equal(got3, want[3])
    ref = orc.r2iq(x, 5, 148, gain=GAIN, rand=True)
    check(np.concatenate([a[1], b[1]]), ref, GAIN, "rates d5")
    with pytest.raises(Exception):
        r.process_rates(x, [7])
    for dup in ([0, 0, 1], [0, 0, 0], [3, 5, 3]):        # a repeated rate would overrun the per-chunk device buffer
        with pytest.raises(Exception, match="listed twice"):
            r.process_rates(x, dup)
    got3b = r.process(x, history=np.zeros(1024, np.int16))   # the handle stays usable after the rejected calls
    assert np.array_equal(got3b, want[3])
    r.TurnOff()
    r.close()


def test_chunks_in_flight_on_several_streams_do_not_share_work_counters():
    """The host-buffer path keeps three chunks in flight on three streams; the persistent kernels hand out
    their frame pairs through a device counter (work_dispenser.cuh).  Two launches sharing one counter would
    skip frames: run a many-chunk capture repeatedly and require the one-launch result bit for bit."""
    import torch
    d, nb = 5, 96
    x = orc.rx888_blocks(nb, d, 256, randomized=True, seed=77)
    r = CudaR2IQ(chunk_blocks=2)
    r.Init(GAIN)
    r.setDecimate(d); r.updateRand(True); r.tunebin = 256; r.TurnOn()
    dev = torch.device("cuda:0")
    xin = torch.from_numpy(np.concatenate([np.zeros(1024, np.int16), x])).to(dev)
    out = torch.empty(nb * (32768 >> d), dtype=torch.complex64, device=dev)
    r.process_device(xin.data_ptr(), nb, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    want = out.cpu().numpy()
    r.TurnOff()
    for _ in range(20):
        r.TurnOn()
        got = r.process(x)
        r.TurnOff()
        assert np.array_equal(got, want)
    r.close()


def _needs_ab_build():
    """The older kernel generations and their run-time switches exist only in a comparison build
    (R2IQ_AB_VARIANTS=1 python -m extio_sddc_b200.build --force); the shipped library holds one kernel per decimation."""
    from extio_sddc_b200 import capi as _capi
    if b"ab-variants" not in _capi.load().r2iq_cuda_version():
        pytest.skip("library built without -DR2IQ_AB_VARIANTS")


@pytest.mark.parametrize("variant", ["0", "1", "2", "3", "5", "7", "9", "17"])
def test_measurement_variants_meet_the_bar(variant):
    """Every work-distribution / staging variant kept behind R2IQ_VARIANT (r2iq_kernels.cuh) meets the parity bar,
    on the device path and through the pipelined host path (comparison builds, generation 5 = the round-1 kernels)."""
    import subprocess
    import sys
    _needs_ab_build()
    code = r'''
import sys, numpy as np
sys.path.insert(0, %r)
from extio_sddc_b200 import CudaR2IQ
from oracle import r2iq_oracle as orc
G = orc.RX888MK2_GAINFACTOR
r = CudaR2IQ(chunk_blocks=2); r.Init(G)
worst = 0.0
for d in range(7):
    for tb, rnd, lsb, nb in ((256, True, False, 5), (0, False, True, 1), (1020, True, False, 2)):
        x = orc.rx888_blocks(nb, d, tb, randomized=rnd, seed=31 + d)
        h = np.random.default_rng(d).integers(-3000, 3000, 1024).astype(np.int16)
        r.setDecimate(d); r.updateRand(rnd); r.setSideband(lsb); r.tunebin = tb; r.TurnOn()
        got = r.process(x, history=h); r.TurnOff()
        ref = orc.r2iq(x, d, tb, gain=G, rand=rnd, sideband=lsb, history=h)
        e = orc.rel_l2(got, ref); m = float(np.abs(got - ref).max()) / orc.full_scale(G)
        assert e <= 1e-5 and m <= 1e-4, (d, tb, e, m)
        worst = max(worst, e)
print("worst", worst)
''' % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))),)
    env = dict(os.environ, R2IQ_VARIANT=variant, R2IQ_KERNEL_GEN="5")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "worst" in r.stdout, r.stdout + r.stderr


@pytest.mark.parametrize("gen", ["1", "2", "3", "4", "5"])
def test_alternative_kernel_generations_meet_the_bar(gen):
    """The kernel families kept in-tree for A/B measurements (env R2IQ_KERNEL_GEN, r2iq_launch.cu
    launch_fused; comparison builds only) must meet the same parity bar as the shipped kernels."""
    import subprocess
    import sys
    _needs_ab_build()
    code = r'''
import sys, numpy as np
sys.path.insert(0, %r)
from extio_sddc_b200 import CudaR2IQ
from oracle import r2iq_oracle as orc
G = orc.RX888MK2_GAINFACTOR
r = CudaR2IQ(); r.Init(G)
worst = 0.0
for d in range(7):
    for tb, rnd, lsb in ((256, True, False), (0, False, True), (1020, True, False), (148, False, False)):
        x = orc.rx888_blocks(3, d, tb, randomized=rnd, seed=31 + d)
        h = np.random.default_rng(d).integers(-3000, 3000, 1024).astype(np.int16)
        r.setDecimate(d); r.updateRand(rnd); r.setSideband(lsb); r.tunebin = tb; r.TurnOn()
        got = r.process(x, history=h); r.TurnOff()
        ref = orc.r2iq(x, d, tb, gain=G, rand=rnd, sideband=lsb, history=h)
        e = orc.rel_l2(got, ref); m = float(np.abs(got - ref).max()) / orc.full_scale(G)
        assert e <= 1e-5 and m <= 1e-4, (d, tb, e, m)
        worst = max(worst, e)
print("worst", worst)
''' % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))),)
    env = dict(os.environ, R2IQ_KERNEL_GEN=gen)
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "worst" in r.stdout, r.stdout + r.stderr


@pytest.mark.parametrize("d", range(7))
def test_fftn8192_reference_as_shipped(d):
    """FFTN_R_ADC = 8192 is what the reference tree ships (Core/config.h:49): halfFft 4096, 11 frames per
    block, 1025 taps, mfft = 4096 >> d.  Same bars as the 2048 family, against the oracle at FFTN 8192."""
    r = CudaR2IQ(fftn=8192)
    r.Init(GAIN)
    assert r.filter_hw(d).shape[0] == 4096
    assert orc.rel_l2(r.filter_hw(d), orc.design_filter(GAIN, d, 8192)) < 1e-6
    for tb, rnd, lsb in ((1024, True, False), (0, False, False), (4092, True, True), (2048, False, True), (584, True, False),
                         (4096 - (4096 >> d) // 2, False, False)):
        tb = max(0, min(4092, tb)) & ~3
        x = orc.rx888_blocks(3, d, tb, fftn=8192, randomized=rnd, seed=8192 + 7 * d + tb)
        hist = np.random.default_rng(d).integers(-3000, 3000, 4096).astype(np.int16)
        r.setDecimate(d); r.updateRand(rnd); r.setSideband(lsb); r.tunebin = tb
        r.TurnOn()
        got = r.process(x, history=hist)
        r.TurnOff()
        ref = orc.r2iq(x, d, tb, gain=GAIN, rand=rnd, sideband=lsb, fftn=8192, history=hist)
        check(got, ref, GAIN, ("fftn8192", d, tb, rnd, lsb))
    # setFreqOffset quantises on halfFft = 4096 (VHF IF point: bin 584)
    r.setDecimate(2)
    res = r.setFreqOffset(4.57e6 / 32e6)
    tbw, want = orc.set_freq_offset(4.57e6 / 32e6, 2, 8192)
    assert r.tunebin == tbw == 584 and res == pytest.approx(want, abs=1e-9)
    # streaming: block-by-block with carried 4096-sample history == one call
    x = orc.rx888_blocks(4, d, 1024, fftn=8192, randomized=True)
    r.setDecimate(d); r.updateRand(True); r.setSideband(False); r.tunebin = 1024
    r.TurnOn(); whole = r.process(x); r.TurnOff()
    r.TurnOn(); parts = np.concatenate([r.process(x[b * 65536:(b + 1) * 65536]) for b in range(4)]); r.TurnOff()
    assert np.array_equal(parts, whole)
    r.close()


@pytest.mark.parametrize("d", [0, 4])
def test_fftn8192_full_size_properties(d):
    """FFTN 8192 at BASELINE config 2 size (8192 blocks): time slabs with a 4096-sample halo == the whole
    capture (bit-exact), linearity, and a sampled window against the oracle."""
    import torch
    n_blocks, half_fft, tb = 8192, 4096, 1024
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev)
    g.manual_seed(8192 + d)
    x = torch.randint(-4000, 4000, (half_fft + n_blocks * 65536,), dtype=torch.int16, device=dev, generator=g)
    opb = 32768 >> d
    r = CudaR2IQ(fftn=8192)
    r.Init(GAIN)
    r.setDecimate(d); r.updateRand(False); r.tunebin = tb
    r.TurnOn()
    s = torch.cuda.current_stream().cuda_stream
    whole = torch.empty(n_blocks * opb, dtype=torch.complex64, device=dev)
    r.process_device(x.data_ptr(), n_blocks, whole.data_ptr(), s)
    cut = 3000
    parts = torch.empty_like(whole)
    r.process_device(x.data_ptr(), cut, parts.data_ptr(), s)
    r.process_device(x.data_ptr() + 2 * cut * 65536, n_blocks - cut, parts.data_ptr() + 8 * cut * opb, s)
    torch.cuda.synchronize()
    assert torch.equal(whole, parts)
    dbl = torch.empty_like(whole)
    x2 = x * 2
    r.process_device(x2.data_ptr(), n_blocks, dbl.data_ptr(), s)
    torch.cuda.synchronize()
    err = (torch.linalg.vector_norm(dbl - 2 * whole) / torch.linalg.vector_norm(dbl)).item()
    assert err < 1e-6, err
    b0 = 5000
    xin = x[half_fft + b0 * 65536: half_fft + (b0 + 2) * 65536].cpu().numpy()
    hist = x[b0 * 65536: half_fft + b0 * 65536].cpu().numpy()
    ref = orc.r2iq(xin, d, tb, gain=GAIN, fftn=8192, history=hist)
    check(whole[b0 * opb:(b0 + 2) * opb].cpu().numpy(), ref, GAIN, ("fftn8192 full-size window", d))
    r.TurnOff()
    r.close()


@pytest.mark.parametrize("fftn", [2048, 8192])
def test_kernels_write_exactly_their_output_range(fftn):
    """compute-sanitizer is closed on this GPU pool, so the out-of-bounds-write check is our own: the IQ buffer sits
    between two guard regions filled with a NaN bit pattern, the input buffer is exactly halo + blocks long, and for
    every decimation, edge tune bins, odd block counts (odd frame totals: the second frame of the last pair is
    computed but must never be stored) every guard word must survive and every output sample must be written."""
    import torch
    dev = torch.device("cuda:0")
    half = fftn // 2
    guard = 1 << 16                                           # complex samples either side
    r = CudaR2IQ(fftn=fftn)
    r.Init(GAIN)
    pattern = torch.tensor([0x7FC0DEAD], dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    for nb in (1, 3):
        x = orc.rx888_blocks(nb, 0, half // 4, randomized=True, seed=nb)
        buf = torch.zeros(half + x.shape[0], dtype=torch.int16, device=dev)
        buf[half:] = torch.from_numpy(x).to(dev)
        for d in range(7):
            n_out = nb * (32768 >> d)
            for tb in (0, half // 4, half - 4):
                raw = pattern.repeat(2 * (n_out + 2 * guard))  # int32 view of [guard | out | guard] complex64
                out = raw[2 * guard: 2 * (guard + n_out)]
                r.setDecimate(d); r.updateRand(True); r.setSideband(False); r.tunebin = tb
                r.TurnOn()
                r.process_device(buf.data_ptr(), nb, out.data_ptr(), stream)
                torch.cuda.synchronize()
                r.TurnOff()
                assert bool((raw[: 2 * guard] == 0x7FC0DEAD).all()), (fftn, d, tb, nb, "write below the output range")
                assert bool((raw[2 * (guard + n_out):] == 0x7FC0DEAD).all()), (fftn, d, tb, nb, "write above the output range")
                assert not bool((out == 0x7FC0DEAD).any()), (fftn, d, tb, nb, "output sample never written")
    r.close()


def test_reference_signal_integrity_assertions():
    """The assertions of the reference's own unittest/signal_integrity_test.cpp, restated on the CUDA path (Dummy radio:
    64 Msps, BBRF103 gain, decimation index 0 -> 32 MS/s):
      :338-354  5 MHz sine, amplitude 10000 -> a clear dominant tone, RMS > 0.1 (checked at its true place: tuned to
                5 MHz the tone sits at DC + residual; untuned at 5 MHz - tunebin frequency)
      :591-595  zeros in -> RMS < 1 (here: exactly zero)
      :746-750  amplitude x2 -> RMS ratio in (1, 4) (here: 2 within 1e-3)
      :813-817  phase-derivative stddev < 1 across buffer boundaries (here: < 1e-3 for a pure tone)
      :557-559  run-to-run RMS within 5 % (here: bit-identical output, also with the dynamic work distribution)"""
    gain = orc.BBRF103_GAINFACTOR
    r = CudaR2IQ()
    r.Init(gain)
    r.setDecimate(0)
    r.updateRand(False)
    r.tunebin = 256                                   # constructor default mtunebin = halfFft/4 = 8 MHz at 64 Msps

    def go(x):
        r.TurnOn()
        y = r.process(x)
        r.TurnOff()
        return y

    x = orc.sine_blocks(8, 5e6, 64e6, amplitude=10000)
    y = go(x)
    assert y.shape == (8 * 32768,)                    # core_test.cpp:165-167: 32768 complex per block at every rate
    steady = y[4096:]                                 # skip the zero-history transient of the first frame
    rms = float(np.sqrt(np.mean(np.abs(steady) ** 2)))
    assert rms > 0.1
    spec = np.abs(np.fft.fft(steady[:65536] * np.hanning(65536)))
    k = int(np.argmax(spec))
    f_det = (k if k < 32768 else k - 65536) * 32e6 / 65536
    assert abs(f_det - (5e6 - 8e6)) <= 2 * 32e6 / 4096, f_det     # tone lands at f_in - f_tune, within the test's 2 bins
    assert spec[k] > 100 * np.median(spec)                         # dominant, not noise
    ph = np.unwrap(np.angle(steady[:10000].astype(np.complex128)))
    assert float(np.std(np.diff(ph))) < 1e-3                       # continuous phase across frame and block edges
    assert np.array_equal(go(x).view(np.uint32), y.view(np.uint32))   # run to run: the same bits
    y2 = go(orc.sine_blocks(8, 5e6, 64e6, amplitude=5000))            # half amplitude
    ratio = rms / float(np.sqrt(np.mean(np.abs(y2[4096:]) ** 2)))
    assert 1.0 < ratio < 4.0 and abs(ratio - 2.0) < 1e-3, ratio
    z = go(np.zeros(4 * 65536, dtype=np.int16))
    assert not z.any()
    r.close()


# ---- stand-alone batched transforms (Core/fft_backend.h:22-50, Core/fft_benchmark.cpp:12-15) -------
@pytest.mark.parametrize("n", [4096, 2048, 1024, 512, 256, 128])
def test_fft_backend_matches_float64_fft(n):
    """r2c / c2c forward / c2c backward of the CUDA FFT backend against numpy's float64 FFT, rel-L2 <= 1e-6 (the shim's
    float64 FFT is the same mathematics): one transform in page-locked buffers (transformed in place by the kernel),
    a ragged batch from pageable arrays (staged), and a large device-resident batch through the asynchronous entry."""
    import torch
    from extio_sddc_b200.fft_backend import BACKWARD, FORWARD, CudaFFTBackend

    rng = np.random.default_rng(n)
    for batch, pinned in ((1, True), (37, False)):
        be = CudaFFTBackend(device=0, batch=batch)
        get = (lambda cnt, dt: be.alloc(cnt, dt)) if pinned else (lambda cnt, dt: np.empty(cnt, dtype=dt))
        x = get(batch * n, np.float32)
        u = get(batch * n, np.complex64)
        X = get(batch * (n // 2 + 1), np.complex64)
        v = get(batch * n, np.complex64)
        w = get(batch * n, np.complex64)
        x[:] = rng.uniform(-0.5, 0.5, batch * n)
        u[:] = rng.uniform(-0.5, 0.5, batch * n) + 1j * rng.uniform(-0.5, 0.5, batch * n)
        pr, pf, pb = be.plan_r2c(n), be.plan_c2c(n, FORWARD), be.plan_c2c(n, BACKWARD)
        be.execute_r2c(pr, x, X)
        be.execute_c2c(pf, u, v)
        be.execute_c2c(pb, u, w)
        want_r = np.fft.rfft(x.astype(np.float64).reshape(batch, n), axis=1)
        want_f = np.fft.fft(u.astype(np.complex128).reshape(batch, n), axis=1)
        want_b = np.fft.ifft(u.astype(np.complex128).reshape(batch, n), axis=1) * n     # FFTW: unnormalised
        for got, want, what in ((X, want_r, "r2c"), (v, want_f, "forward"), (w, want_b, "backward")):
            err = np.linalg.norm(got.reshape(want.shape) - want) / np.linalg.norm(want)
            assert err <= 1e-6, (what, n, batch, err)
        for p in (pr, pf, pb):
            be.destroy_plan(p)
        if pinned:
            for a in (x, u, X, v, w):
                be.free(a)
    # device-resident, asynchronous, batch far beyond one wave of CTAs (and not a multiple of the teams per CTA)
    batch = 20011
    be = CudaFFTBackend(device=0, batch=1)
    gen = torch.Generator(device="cuda").manual_seed(n)
    d_in = torch.rand((batch, n, 2), device="cuda", generator=gen) - 0.5
    d_out = torch.empty_like(d_in)
    pf = be.plan_c2c(n, FORWARD)
    be.execute_device(pf, d_in.data_ptr(), d_out.data_ptr(), batch, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    want = torch.fft.fft(torch.view_as_complex(d_in.double()), dim=1)
    got = torch.view_as_complex(d_out.double())
    err = float(torch.linalg.norm(got - want) / torch.linalg.norm(want))
    assert err <= 1e-6, (n, err)
    d_r = torch.rand((batch, n), device="cuda", generator=gen) - 0.5
    d_R = torch.empty((batch, n // 2 + 1, 2), device="cuda")
    pr = be.plan_r2c(n)
    be.execute_device(pr, d_r.data_ptr(), d_R.data_ptr(), batch, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    want = torch.fft.rfft(d_r.double(), dim=1)
    err = float(torch.linalg.norm(torch.view_as_complex(d_R.double()) - want) / torch.linalg.norm(want))
    assert err <= 1e-6, (n, err)
    be.destroy_plan(pf)
    be.destroy_plan(pr)
