"""Development aid: channelizer timing for a few (blocks, channels, d)."""
import os, sys
sys.path.insert(0, os.getcwd())
import torch
from extio_sddc_b200 import CudaR2IQ, shard
dev = torch.device("cuda:0"); g = torch.Generator(device=dev); g.manual_seed(1)
r = CudaR2IQ(); r.Init(1.08e-8); r.updateRand(True)
s = torch.cuda.current_stream().cuda_stream
for n_blocks, nch, d in ((2048, 64, 5), (2048, 32, 5), (2048, 1, 5), (2048, 33, 5), (512, 64, 5), (2048, 8, 3), (2048, 64, 3), (2048, 64, 6)):
    x = torch.randint(-8000, 8000, (1024 + n_blocks * 65536,), dtype=torch.int16, device=dev, generator=g)
    bins = [(4 * (3 + 5 * c)) % 1024 for c in range(nch)]
    out = torch.empty(nch * n_blocks * (32768 >> d), dtype=torch.complex64, device=dev)
    r.setDecimate(d); r.TurnOn()
    for _ in range(2): r.process_channels_device(x.data_ptr(), n_blocks, bins, out.data_ptr(), s)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): r.process_channels_device(x.data_ptr(), n_blocks, bins, out.data_ptr(), s)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"blocks={n_blocks} ch={nch} d={d}: {ms:.3f} ms  input {n_blocks*65536/ms/1e6:.1f} Gs/s  us/frame/warp {ms*1e3/(n_blocks*43/1184):.2f}")
    del x, out
