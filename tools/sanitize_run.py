"""Small all-kernels exercise for compute-sanitizer: every decimation (default kernels and the older
generations), odd frame counts, edge tune bins, the channelizer and the convert hook."""
import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
from extio_sddc_b200 import CudaR2IQ
from oracle import r2iq_oracle as orc
r = CudaR2IQ(); r.Init(1.08e-8)
x = orc.rx888_blocks(3, 2, 256, randomized=True)
for d in range(7):
    for tb in (0, 256, 1020):
        r.setDecimate(d); r.updateRand(True); r.tunebin = tb; r.TurnOn()
        y = r.process(x[: (1 + d % 3) * 65536]); r.TurnOff()
        assert np.isfinite(y.view(np.float32)).all()
for d in (1, 3, 5, 6):
    r.setDecimate(d); r.TurnOn(); y = r.process_channels(x[:65536 * 2], [0, 4, 256, 1020, 512]); r.TurnOff()
    assert np.isfinite(y.view(np.float32)).all()
r.convert_float(x[:4096], True)
# several rates from one upload, and the FFTN 8192 family (the reference tree as shipped)
r.setDecimate(0); r.TurnOn()
for y in r.process_rates(x[:65536 * 2], [0, 2, 5]):
    assert np.isfinite(y.view(np.float32)).all()
r.TurnOff()
w = CudaR2IQ(fftn=8192); w.Init(1.08e-8)
for d in range(7):
    for tb in (0, 1024, 4092):
        w.setDecimate(d); w.updateRand(True); w.tunebin = tb; w.TurnOn()
        y = w.process(x[: (1 + d % 2) * 65536]); w.TurnOff()
        assert np.isfinite(y.view(np.float32)).all()
w.close()
print("sanitize run ok, generation", os.environ.get("R2IQ_KERNEL_GEN", "0"))
