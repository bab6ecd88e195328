"""Development aid: time one decimation for ~0.5 s while sampling SM clock / power through NVML."""
import os, sys, threading, time
sys.path.insert(0, os.getcwd())
import torch, pynvml
from extio_sddc_b200 import CudaR2IQ
n_blocks = int(sys.argv[1]); d = int(sys.argv[2]); reps = int(sys.argv[3]) if len(sys.argv) > 3 else 300
pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
dev = torch.device("cuda:0"); g = torch.Generator(device=dev); g.manual_seed(1)
x = torch.randint(-8000, 8000, (1024 + n_blocks * 65536,), dtype=torch.int16, device=dev, generator=g)
out = torch.empty(n_blocks * 32768, dtype=torch.complex64, device=dev)
r = CudaR2IQ(); r.Init(1.08e-8); r.setDecimate(d); r.updateRand(True); r.tunebin = 256; r.TurnOn()
s = torch.cuda.current_stream().cuda_stream
for _ in range(5): r.process_device(x.data_ptr(), n_blocks, out.data_ptr(), s)
torch.cuda.synchronize()
clk, pw, stop = [], [], threading.Event()
def loop():
    while not stop.is_set():
        clk.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)); pw.append(pynvml.nvmlDeviceGetPowerUsage(h) / 1000); time.sleep(0.005)
t = threading.Thread(target=loop); t.start()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps): r.process_device(x.data_ptr(), n_blocks, out.data_ptr(), s)
e1.record(); torch.cuda.synchronize(); stop.set(); t.join()
ms = e0.elapsed_time(e1) / reps
clk.sort(); pw.sort()
print(f"d={d} blocks={n_blocks}: {ms:.3f} ms {n_blocks*65536/ms/1e6:.1f} Gs/s  clk med {clk[len(clk)//2]} min {clk[0]} max {clk[-1]} MHz  power med {pw[len(pw)//2]:.0f} W  ({len(clk)} samples)")
