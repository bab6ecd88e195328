"""Development aid (needs a comparison build: R2IQ_AB_VARIANTS=1 python -m extio_sddc_b200.build --force; generation 5 =
the round-1 kernels): A/B timing of the measurement variants of the dual-frame kernels (env R2IQ_VARIANT,
R2IQ_CTAS_PER_SM) on a B200, with a parity check of every variant first.
    python tools/variant_sweep.py [--blocks N] [--variants 0,1,2,3] [--decims 0,2,5] [--ctas 1,2]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

from extio_sddc_b200 import CudaR2IQ  # noqa: E402
from oracle import r2iq_oracle as orc  # noqa: E402

GAIN = orc.RX888MK2_GAINFACTOR


def parity(decims):
    r = CudaR2IQ()
    r.Init(GAIN)
    worst = 0.0
    for d in decims:
        for (tb, rnd, sb, nb) in ((256, True, False, 5), (1020, True, True, 2), (0, False, False, 1)):
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
            if not (e <= 1e-5 and ma <= 1e-4):
                print(f"  PARITY FAIL d={d} tb={tb} rel_l2={e:.3e} maxabs={ma:.3e}")
    r.close()
    return worst


def timing(n_blocks, decims, reps=5):
    import torch
    r = CudaR2IQ()
    r.Init(GAIN)
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev); g.manual_seed(1)
    x = torch.randint(-8000, 8000, (1024 + n_blocks * 65536,), dtype=torch.int16, device=dev, generator=g)
    out = torch.empty(n_blocks * 32768, dtype=torch.complex64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    res = {}
    for d in decims:
        r.setDecimate(d); r.updateRand(True); r.setSideband(False); r.tunebin = 256
        r.TurnOn()
        for _ in range(3):
            r.process_device(x.data_ptr(), n_blocks, out.data_ptr(), stream)
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            r.process_device(x.data_ptr(), n_blocks, out.data_ptr(), stream)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        res[d] = n_blocks * 65536 / ms / 1e6
        r.TurnOff()
    r.close()
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--blocks", type=int, default=8192)
    ap.add_argument("--variants", default="0,1,2,3")
    ap.add_argument("--decims", default="0,1,2,3,4,5")
    ap.add_argument("--ctas", default="")
    ap.add_argument("--no-parity", action="store_true")
    a = ap.parse_args()
    decims = [int(v) for v in a.decims.split(",")]
    for v in a.variants.split(","):
        os.environ["R2IQ_VARIANT"] = v
        for c in (a.ctas.split(",") if a.ctas else [""]):
            if c:
                os.environ["R2IQ_CTAS_PER_SM"] = c
            else:
                os.environ.pop("R2IQ_CTAS_PER_SM", None)
            w = -1.0 if a.no_parity else parity(decims)
            res = timing(a.blocks, decims)
            print(f"variant {v} ctas/SM {c or 'max'}: worst rel_l2 {w:.2e} | " +
                  "  ".join(f"d{d}: {g:6.1f}" for d, g in res.items()), flush=True)
