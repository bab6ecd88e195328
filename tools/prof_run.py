"""Short profiling target: one warm-up + one launch of the fused kernel per decimation (d = 0..6) on a
2048-block RX-888 style capture resident in HBM (argv: blocks, decimations, fftn).  Used under ncu (see profiles/README.md)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from extio_sddc_b200 import CudaR2IQ  # noqa: E402

n_blocks = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
decims = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else list(range(7))
fftn = int(sys.argv[3]) if len(sys.argv) > 3 else 2048
dev = torch.device("cuda:0")
g = torch.Generator(device=dev)
g.manual_seed(1)
x = torch.randint(-8000, 8000, (fftn // 2 + n_blocks * 65536,), dtype=torch.int16, device=dev, generator=g)
out = torch.empty(n_blocks * 32768, dtype=torch.complex64, device=dev)
op = CudaR2IQ(fftn=fftn)
op.Init(1.08e-8)
op.updateRand(True)
op.tunebin = fftn // 8
s = torch.cuda.current_stream().cuda_stream
for rep in range(2):
    for d in decims:
        op.setDecimate(d)
        op.TurnOn()
        op.process_device(x.data_ptr(), n_blocks, out.data_ptr(), s)
torch.cuda.synchronize()
op.TurnOff()
print("ok", float(out[0].real))
