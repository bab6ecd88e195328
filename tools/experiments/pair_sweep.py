"""Development aid: parity + A/B timing of the FMA-token kernels (r2iq_tok.cu, env R2IQ_PAIR="<wps><var>") against the
round-1 default kernels on a B200.
    python tools/tok_sweep.py [--blocks N] [--configs base,8,12] [--decims 0,1,2,3,4,5]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import numpy as np  # noqa: E402

from extio_sddc_b200 import CudaR2IQ  # noqa: E402
from oracle import r2iq_oracle as orc  # noqa: E402

GAIN = orc.RX888MK2_GAINFACTOR


def parity(decims):
    r = CudaR2IQ()
    r.Init(GAIN)
    worst = 0.0
    ok = True
    for d in decims:
        for (tb, rnd, sb, nb) in ((256, True, False, 5), (1020, True, True, 3), (0, False, False, 1), (144, True, False, 37)):
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
                ok = False
                print(f"  PARITY FAIL d={d} tb={tb} nb={nb} rel_l2={e:.3e} maxabs={ma:.3e}", flush=True)
    r.close()
    return worst, ok


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
        best = 1e9
        for _ in range(reps):
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            r.process_device(x.data_ptr(), n_blocks, out.data_ptr(), stream)
            e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        res[d] = n_blocks * 65536 / best / 1e6
        r.TurnOff()
    r.close()
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--blocks", type=int, default=8192)
    ap.add_argument("--configs", default="base,8,12")
    ap.add_argument("--decims", default="0,1,2,3,4,5")
    ap.add_argument("--no-parity", action="store_true")
    a = ap.parse_args()
    decims = [int(v) for v in a.decims.split(",")]
    for c in a.configs.split(","):
        if c == "base":
            os.environ.pop("R2IQ_PAIR", None)
        else:
            os.environ["R2IQ_PAIR"] = c
        w, ok = (-1.0, True) if a.no_parity else parity(decims)
        res = timing(a.blocks, decims)
        print(f"config {c:>4}: parity {'ok ' if ok else 'BAD'} worst rel_l2 {w:.2e} | " +
              "  ".join(f"d{d}: {g:6.1f}" for d, g in res.items()), flush=True)
