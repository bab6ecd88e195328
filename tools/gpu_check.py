"""Development aid: parity table + kernel timing of the fused r2iq kernels on a B200.
Run on the GPU box:  python tools/gpu_check.py [--blocks N]
Parity is checked against the NumPy restatement (oracle/r2iq_oracle.py); timing uses CUDA events
on torch's current stream, input and output resident in HBM.
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

from extio_sddc_b200 import CudaR2IQ  # noqa: E402
from oracle import r2iq_oracle as orc  # noqa: E402

GAIN = orc.RX888MK2_GAINFACTOR


DECIMS = list(range(7))


def parity():
    r = CudaR2IQ()
    r.Init(GAIN)
    worst = 0.0
    x = (np.arange(65536) * 13 - 512).astype(np.int16)
    for rnd in (False, True):
        ok = np.array_equal(r.convert_float(x, rnd), orc.convert_float(x, rnd))
        print(f"convert rand={rnd}: bit-exact={ok}")
    cases = []
    for d in DECIMS:
        cases += [(d, 256, True, False), (d, 0, False, False), (d, 1020, True, True), (d, 512, False, True), (d, 144, True, False)]
    for (d, tb, rnd, sb) in cases:
        nb = 3
        blocks = orc.rx888_blocks(nb, d, tb, randomized=rnd, seed=100 + d)
        hist = np.random.default_rng(d).integers(-3000, 3000, 1024).astype(np.int16)
        r.setDecimate(d); r.updateRand(rnd); r.setSideband(sb); r.tunebin = tb
        r.TurnOn()
        got = r.process(blocks, history=hist)
        r.TurnOff()
        ref = orc.r2iq(blocks, d, tb, gain=GAIN, rand=rnd, sideband=sb, history=hist)
        e = orc.rel_l2(got, ref)
        ma = np.abs(got - ref).max() / orc.full_scale(GAIN)
        worst = max(worst, e)
        print(f"d={d} tb={tb:4d} rand={int(rnd)} lsb={int(sb)}  rel_l2={e:.3e}  maxabs/fs={ma:.3e} {'OK' if e <= 1e-5 and ma <= 1e-4 else 'FAIL'}")
    print("worst rel_l2", worst)
    return worst


def timing(n_blocks, decims=range(7), fftn=2048):
    import torch
    r = CudaR2IQ(fftn=fftn)
    r.Init(GAIN)
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev); g.manual_seed(1)
    x = torch.randint(-8000, 8000, (fftn // 2 + n_blocks * 65536,), dtype=torch.int16, device=dev, generator=g)
    out = torch.empty(n_blocks * 32768, dtype=torch.complex64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    for d in decims:
        r.setDecimate(d); r.updateRand(True); r.setSideband(False); r.tunebin = fftn // 8
        r.TurnOn()
        for _ in range(3):
            r.process_device(x.data_ptr(), n_blocks, out.data_ptr(), stream)
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record()
        for _ in range(reps):
            r.process_device(x.data_ptr(), n_blocks, out.data_ptr(), stream)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        samples = n_blocks * 65536
        bps = 2 + 8 / 2 ** (d + 1)
        print(f"d={d}: {ms:8.3f} ms  {samples / ms / 1e6:8.1f} Gsamples/s  {samples * bps / ms / 1e6:8.1f} GB/s  frac(6551)={samples * bps / ms / 1e6 / 6551:.3f}")
        r.TurnOff()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--blocks", type=int, default=2048)
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--decims", default="0,1,2,3,4,5,6")
    ap.add_argument("--fftn", type=int, default=2048)
    ap.add_argument("--lib", default=None, help="an experiment build of libr2iq_cuda.so (tools/exp_build.py) instead of the product library")
    a = ap.parse_args()
    if a.lib:
        from extio_sddc_b200 import capi
        capi.LIB_PATH = os.path.abspath(a.lib)
        print("library:", capi.LIB_PATH)
    DECIMS[:] = [int(v) for v in a.decims.split(",")]
    if not a.no_parity:
        parity()
    timing(a.blocks, [int(v) for v in a.decims.split(",")], a.fftn)
