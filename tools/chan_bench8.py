"""FFTN 8192 channelizer timing (64 channels, d = 5 by default): python tools/chan_bench8.py [channels=64] [d=5] [blocks=8192]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from extio_sddc_b200 import CudaR2IQ  # noqa: E402

n_ch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
d = int(sys.argv[2]) if len(sys.argv) > 2 else 5
n_blocks = int(sys.argv[3]) if len(sys.argv) > 3 else 8192
torch.cuda.set_device(0)
op = CudaR2IQ(device=0, fftn=8192)
op.Init(1.0)
op.setDecimate(d)
op.TurnOn()
gen = torch.Generator(device="cuda").manual_seed(1)
x = torch.randint(-3000, 3000, (4096 + n_blocks * 65536,), dtype=torch.int16, device="cuda", generator=gen)
out = torch.empty((n_ch, n_blocks * (32768 >> d), 2), dtype=torch.float32, device="cuda")
bins = [(64 * c) % 4096 for c in range(n_ch)]
st = torch.cuda.current_stream().cuda_stream
for _ in range(2):
    op.process_channels_device(x.data_ptr(), n_blocks, bins, out.data_ptr(), st)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    op.process_channels_device(x.data_ptr(), n_blocks, bins, out.data_ptr(), st)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
alg = n_blocks * 65536 * (2 + n_ch * 8 / (2 << d))
print(f"fftn=8192 channels={n_ch} d={d} blocks={n_blocks} ms={ms:.3f} alg_GB={alg / 1e9:.3f} GB/s={alg / ms / 1e6:.0f}")
op.TurnOff()
