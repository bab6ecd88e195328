"""Cost of the residual fine-tune mixer (Core/RadioHandler.cpp:33-37, Core/pffft/pf_mixer.cpp:750-856) on
the device-resident path, per decimation, both FFTN:
  * mixer off;
  * R2IQ_NCO_DIRECT: phases computed on the device -- inside the store epilogue of the FFTN 2048 kernels, as one in-place
    pass after the FFTN 8192 kernels;
  * R2IQ_NCO_REFERENCE: the reference's float32 recurrence, bit-identical stream (the start phasor of every 128-sample
    block comes from a sequential host recurrence, then a stand-alone in-place pass).
    python tools/mixer_bench.py [blocks=8192]       -> gpurun_out/mixer_bench.txt (committed as profiles/r02_mixer_bench.txt)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from extio_sddc_b200 import CudaR2IQ, capi  # noqa: E402

nb = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
dev = torch.device("cuda:0")
g = torch.Generator(device=dev)
g.manual_seed(1)
lines = []
for fftn in (2048, 8192):
    x = torch.randint(-8000, 8000, (fftn // 2 + nb * 65536,), dtype=torch.int16, device=dev, generator=g)
    out = torch.empty(nb * 32768, dtype=torch.complex64, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    table = {}
    for name, mode, mix in (("mixer off", capi.NCO_REFERENCE, False), ("direct-phase mixer (R2IQ_NCO_DIRECT)", capi.NCO_DIRECT, True),
                            ("reference-bit mixer (R2IQ_NCO_REFERENCE)", capi.NCO_REFERENCE, True)):
        r = CudaR2IQ(fftn=fftn)
        r.Init(1.08e-8)
        r.updateRand(True)
        r.setFineTuneMode(mode)
        res = []
        for d in range(6):
            r.setDecimate(d)
            fc = r.setFreqOffset(0.2503)
            r.setFineTune(fc if mix else 0.0)
            r.TurnOn()
            reps = 5 if not (mix and mode == capi.NCO_REFERENCE) else 2
            for _ in range(2):
                r.process_device(x.data_ptr(), nb, out.data_ptr(), s)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                r.process_device(x.data_ptr(), nb, out.data_ptr(), s)
            e1.record()
            torch.cuda.synchronize()
            res.append(nb * 65536 / (e0.elapsed_time(e1) / reps) / 1e6)
            r.TurnOff()
        r.close()
        table[name] = res
        lines.append(f"FFTN {fftn}  {name:42s} Gsamples/s: " + "  ".join(f"d{d} {v:7.1f}" for d, v in enumerate(res)))
    off = table["mixer off"]
    for name in list(table)[1:]:
        lines.append(f"FFTN {fftn}  {name:42s} / mixer off: " + "  ".join(f"d{d} {v / o:7.3f}" for d, (v, o) in enumerate(zip(table[name], off))))
    del x, out
text = "\n".join(lines) + "\n"
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", "mixer_bench.txt"), "w") as f:
    f.write(text)
print(text)
