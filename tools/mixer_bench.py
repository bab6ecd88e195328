"""Cost of the residual fine-tune mixer on the device-resident path: no mixer, mixer folded into the
store epilogue of the default kernels (R2IQ_NCO_FUSED=1), stand-alone in-place pass (default).
    python tools/mixer_bench.py [blocks]      # one subprocess per mode: the switch is read per launch from the env
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODE = r'''
import os, sys
sys.path.insert(0, %r)
import torch
from extio_sddc_b200 import CudaR2IQ
nb = int(sys.argv[1]); mix = sys.argv[2] == "1"
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(1)
x = torch.randint(-8000, 8000, (1024 + nb * 65536,), dtype=torch.int16, device=dev, generator=g)
out = torch.empty(nb * 32768, dtype=torch.complex64, device=dev)
r = CudaR2IQ(); r.Init(1.08e-8); r.updateRand(True)
s = torch.cuda.current_stream().cuda_stream
res = []
for d in range(6):
    r.setDecimate(d)
    fc = r.setFreqOffset(0.2503)
    r.setFineTune(fc if mix else 0.0)
    r.TurnOn()
    for _ in range(3): r.process_device(x.data_ptr(), nb, out.data_ptr(), s)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): r.process_device(x.data_ptr(), nb, out.data_ptr(), s)
    e1.record(); torch.cuda.synchronize()
    res.append(nb * 65536 / (e0.elapsed_time(e1) / 5) / 1e6)
    r.TurnOff()
print("  ".join("d%%d %%6.1f" %% (d, v) for d, v in enumerate(res)))
''' % ROOT

nb = sys.argv[1] if len(sys.argv) > 1 else "8192"
for name, mix, env in (("no mixer", "0", {}), ("stand-alone mixer pass", "1", {}), ("mixer in the store epilogue", "1", {"R2IQ_NCO_FUSED": "1"})):
    r = subprocess.run([sys.executable, "-c", CODE, nb, mix], env=dict(os.environ, **env), capture_output=True, text=True)
    print(f"{name:28s} Gsamples/s: {r.stdout.strip() or r.stderr[-400:]}", flush=True)
