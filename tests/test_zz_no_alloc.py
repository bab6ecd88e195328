"""r2iq_cuda_process_device_static (include/r2iq_cuda.h): the device path with work counter and tables in the library's
file-scope static device array -- no cudaMalloc inside the call.  No counterpart in the reference,
so the check is against the existing entry point r2iq_cuda_process_device: same kernel, same
tables => the same bits (and the north-star tolerances against the oracle for one case).

Named test_zz_* so that pytest collects it LAST: a failure here hides nothing.
Runs alone in a fresh process and after the other GPU tests (it owns its handles; the static array is re-cleared and
its tables are re-validated by every call).
"""

import numpy as np
import pytest

from extio_sddc_b200 import CudaR2IQ, capi
from oracle import r2iq_oracle as orc

pytestmark = pytest.mark.gpu

GAIN = orc.RX888MK2_GAINFACTOR
REL_L2_TOL = 1e-5          # north_star: relative L2 error <= 1e-5
MAX_ABS_TOL = 1e-4         # north_star: max abs <= 1e-4 of full scale
STATIC_MAX_BLOCKS = 8192   # R2IQ_STATIC_MAX_BLOCKS


def _device_input(x, dev, hist=None):
    import torch
    buf = torch.zeros(1024 + x.shape[0], dtype=torch.int16, device=dev)
    if hist is not None:
        buf[:1024] = torch.from_numpy(hist).to(dev)
    buf[1024:] = torch.from_numpy(x).to(dev)
    return buf


def _setup(r, d, tb, rnd, lsb):
    r.TurnOff()
    r.setDecimate(d)
    r.updateRand(rnd)
    r.setSideband(lsb)
    r.tunebin = tb
    r.TurnOn()


@pytest.mark.parametrize("d", range(7))
def test_static_entry_matches_process_device_twice_in_a_row(d):
    import torch
    dev = torch.device("cuda:0")
    s = torch.cuda.current_stream().cuda_stream
    r = CudaR2IQ()
    r.Init(GAIN)
    n_blocks = 5
    opb = 32768 >> d
    try:
        for tb, rnd, lsb in ((256, True, False), (0, False, True), (1020, True, True), ((1024 >> d) // 2 & ~3, False, False)):
            x = orc.rx888_blocks(n_blocks, d, tb, randomized=rnd, seed=4242 + 11 * d + tb)
            hist = np.random.default_rng(100 + d).integers(-3000, 3000, 1024).astype(np.int16)
            buf = _device_input(x, dev, hist)
            _setup(r, d, tb, rnd, lsb)
            ref = torch.zeros(n_blocks * opb, dtype=torch.complex64, device=dev)
            r.process_device(buf.data_ptr(), n_blocks, ref.data_ptr(), s)
            # two calls in a row on the shared array: the second one finds the first one's counter and tables
            outs = [torch.zeros_like(ref) for _ in range(2)]
            for o in outs:
                r.process_device_static(buf.data_ptr(), n_blocks, o.data_ptr(), s)
            torch.cuda.synchronize()
            ref_h = ref.cpu().numpy()
            for k, o in enumerate(outs):
                got = o.cpu().numpy()
                e = orc.rel_l2(got, ref_h)
                m = float(np.abs(got.astype(np.complex128) - ref_h.astype(np.complex128)).max()) / orc.full_scale(GAIN)
                assert e <= REL_L2_TOL and m <= MAX_ABS_TOL, (d, tb, rnd, lsb, "call", k, e, m)
                assert np.array_equal(got, ref_h), (d, tb, rnd, lsb, "call", k, "same kernel and tables: bits must match")
            if tb == 256:   # and against the oracle itself
                want = orc.r2iq(x, d, tb, gain=GAIN, rand=rnd, sideband=lsb, history=hist)
                got = outs[1].cpu().numpy()
                assert orc.rel_l2(got, want) <= REL_L2_TOL
                assert float(np.abs(got.astype(np.complex128) - want.astype(np.complex128)).max()) / orc.full_scale(GAIN) <= MAX_ABS_TOL
    finally:
        r.close()


def test_static_entry_two_handles_share_the_array():
    """Handles with different filters alternate on the one array: every call must find (or rebuild) ITS tables."""
    import torch
    dev = torch.device("cuda:0")
    s = torch.cuda.current_stream().cuda_stream
    a, b = CudaR2IQ(), CudaR2IQ()
    a.Init(GAIN)
    b.Init(GAIN * 3.0)
    try:
        d, tb, n_blocks = 3, 512, 4
        x = orc.rx888_blocks(n_blocks, d, tb, randomized=False, seed=77)
        buf = _device_input(x, dev)
        refs, gots = [], []
        for r in (a, b):
            _setup(r, d, tb, False, False)
            ref = torch.zeros(n_blocks * (32768 >> d), dtype=torch.complex64, device=dev)
            r.process_device(buf.data_ptr(), n_blocks, ref.data_ptr(), s)
            refs.append(ref)
        for r in (a, b, a, b):
            o = torch.zeros_like(refs[0])
            r.process_device_static(buf.data_ptr(), n_blocks, o.data_ptr(), s)
            gots.append(o)
        torch.cuda.synchronize()
        assert not torch.equal(refs[0], refs[1])
        for k, o in enumerate(gots):
            assert torch.equal(o, refs[k % 2]), k
    finally:
        a.close()
        b.close()


def test_static_entry_direct_mixer_and_steady_state_memory():
    """Direct-phase fine-tune mixer through the static entry == through process_device (phase carried across calls), and
    steady-state calls leave the device's free memory untouched (cudaMemGetInfo before and after)."""
    import torch
    dev = torch.device("cuda:0")
    s = torch.cuda.current_stream().cuda_stream
    d, tb, n_blocks = 2, 640, 6
    x = orc.rx888_blocks(n_blocks, d, tb, randomized=True, seed=5)
    buf = _device_input(x, dev)
    res = []
    for static in (False, True):
        r = CudaR2IQ()
        r.Init(GAIN)
        try:
            r.setFineTuneMode(capi.NCO_DIRECT)
            _setup(r, d, tb, True, False)
            assert r.setFineTune(0.0123) != 0.0
            call = r.process_device_static if static else r.process_device
            outs = [torch.zeros(n_blocks * (32768 >> d), dtype=torch.complex64, device=dev) for _ in range(3)]
            call(buf.data_ptr(), n_blocks, outs[0].data_ptr(), s)
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info(dev)[0]
            for o in outs[1:]:
                call(buf.data_ptr(), n_blocks, o.data_ptr(), s)
            torch.cuda.synchronize()
            if static:
                assert torch.cuda.mem_get_info(dev)[0] == free0, "a steady-state static call changed the device's free memory"
            res.append(outs)
        finally:
            r.close()
    for k in range(3):
        assert torch.equal(res[0][k], res[1][k]), k
    assert not torch.equal(res[0][0], res[0][1])      # the mixer's phase moved on between the calls


def test_static_entry_limits_and_errors():
    import torch
    dev = torch.device("cuda:0")
    s = torch.cuda.current_stream().cuda_stream
    small = torch.zeros(1024 + 65536, dtype=torch.int16, device=dev)
    out = torch.zeros(32768, dtype=torch.complex64, device=dev)
    r = CudaR2IQ()
    r.Init(GAIN)
    try:
        with pytest.raises(capi.R2iqError):                       # operator off
            r.process_device_static(small.data_ptr(), 1, out.data_ptr(), s)
        _setup(r, 0, 256, False, False)
        r.process_device_static(small.data_ptr(), 0, out.data_ptr(), s)            # empty launch: nothing to do
        with pytest.raises(capi.R2iqError, match="at most 8192"):  # beyond the stated maximum: refused before any launch
            r.process_device_static(small.data_ptr(), STATIC_MAX_BLOCKS + 1, out.data_ptr(), s)
        with pytest.raises(capi.R2iqError, match="aligned"):
            r.process_device_static(small.data_ptr() + 2, 1, out.data_ptr(), s)
        with pytest.raises(capi.R2iqError):
            r.process_device_static(0, 1, out.data_ptr(), s)
        # the reference-bit mixer needs launch-sized scratch: refused with a message that says what to use instead
        assert r.setFineTune(0.01) != 0.0
        with pytest.raises(capi.R2iqError, match="R2IQ_NCO_DIRECT"):
            r.process_device_static(small.data_ptr(), 1, out.data_ptr(), s)
        r.setFineTune(0.0)
        r.process_device_static(small.data_ptr(), 1, out.data_ptr(), s)            # and the handle is still usable
        torch.cuda.synchronize()
    finally:
        r.close()
    w = CudaR2IQ(fftn=8192)
    w.Init(GAIN)
    try:
        _setup(w, 0, 1024, False, False)
        big = torch.zeros(4096 + 65536, dtype=torch.int16, device=dev)
        with pytest.raises(capi.R2iqError, match="FFTN 2048"):
            w.process_device_static(big.data_ptr(), 1, out.data_ptr(), s)
    finally:
        w.close()
