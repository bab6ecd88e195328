"""FFT backend comparison: the reference's fft_benchmark table (Core/fft_benchmark.cpp:39-128) for
  * the reference's own backend on this box's host cores: oracle/_ref/fft_benchmark_ref, the UNMODIFIED
    Core/fft_benchmark.cpp + Core/fft_backend_fftw.cpp over the fftw3 shim with MKL DFTI underneath (one thread, one
    transform per call: what the reference publishes in README.md:175-180);
  * the reference's unmodified benchmark with CudaFFTBackend compiled in (oracle/_ref/fft_benchmark_dropin): one
    transform per blocking call on host buffers -- launch latency, printed because that is what the program measures;
  * CudaFFTBackend batched on host buffers (extio_sddc_b200/bin/fft_benchmark_cuda: copies inside every call);
  * CudaFFTBackend batched and device-resident (this script: CUDA events, input larger than L2, microseconds per
    transform and the fraction of the measured HBM copy bandwidth: 8 or 16 algorithmic bytes per point).
Writes gpurun_out/fft_benchmark.txt (copied to profiles/r02_fft_benchmark.txt).  Nothing here reads /root/reference."""
import json
import os
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from extio_sddc_b200 import capi  # noqa: E402
from extio_sddc_b200.fft_backend import BACKWARD, FORWARD, CudaFFTBackend  # noqa: E402

SIZES = [4096, 2048, 1024, 512, 256, 128]
OUT = os.path.join(ROOT, "gpurun_out")


def run(cmd, env=None):
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
        return r.stdout + (r.stderr if r.returncode else "")
    except Exception as exc:  # a missing binary is reported, not fatal
        return f"(not run: {exc})\n"


def device_table():
    capi.load()
    torch.cuda.set_device(0)
    peaks = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm = json.load(open(peaks)).get("hbm_gbs") if os.path.exists(peaks) else None
    hbm = float(hbm) if hbm else 6500.0
    be = CudaFFTBackend(device=0, batch=1)
    stream = torch.cuda.current_stream().cuda_stream
    lines = []
    total_points = 1 << 26            # 512 MiB of complex input per launch: four times the L2
    for n in SIZES:
        batch = total_points // n
        d_c = torch.rand((batch, n, 2), device="cuda") - 0.5
        d_o = torch.empty((batch, n + 1, 2), device="cuda")
        d_r = torch.rand((batch, n), device="cuda") - 0.5
        row = []
        for kind, plan, src, bytes_per in (("r2c", be.plan_r2c(n), d_r, n * 4 + (n // 2 + 1) * 8),
                                           ("fwd", be.plan_c2c(n, FORWARD), d_c, n * 16),
                                           ("bwd", be.plan_c2c(n, BACKWARD), d_c, n * 16)):
            for _ in range(3):
                be.execute_device(plan, src.data_ptr(), d_o.data_ptr(), batch, stream)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 10
            torch.cuda.synchronize()
            e0.record()
            for _ in range(reps):
                be.execute_device(plan, src.data_ptr(), d_o.data_ptr(), batch, stream)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            row.append((ms * 1e3 / batch, batch * bytes_per / (ms * 1e-3) / 1e9))
            be.destroy_plan(plan)
        lines.append((n, batch, row))
        del d_c, d_o, d_r
    out = ["=======================================================================",
           "  CudaFFTBackend, DEVICE-RESIDENT, batched (CUDA events, 10 launches each)",
           f"  per-transform microseconds | GB/s of algorithmic bytes (fraction of the measured {hbm:.0f} GB/s HBM copy rate)",
           "-----------------------------------------------------------------------",
           "  Size  |  batch  |   R2C (us)          |  C2C Fwd (us)       |  C2C Bwd (us)",
           "-----------------------------------------------------------------------"]
    for n, batch, row in lines:
        cells = " | ".join(f"{us:8.5f} {gbs:6.0f} ({gbs / hbm:4.2f})" for us, gbs in row)
        out.append(f"  {n:4d}  | {batch:6d}  | {cells}")
    out.append("=======================================================================")
    return "\n".join(out) + "\n"


def main():
    os.makedirs(OUT, exist_ok=True)
    parts = []
    env = dict(os.environ)
    mkl = os.path.join(os.path.dirname(torch.__file__), "lib", "libtorch_cpu.so")
    if os.path.exists(mkl):
        env["FFTW3_SHIM_MKL"] = mkl
    parts.append("### 1. reference backend on this box's host: unmodified Core/fft_benchmark.cpp + fft_backend_fftw.cpp, "
                 "fftw3 shim over MKL DFTI (one thread)\n")
    parts.append(run([os.path.join(ROOT, "oracle", "_ref", "fft_benchmark_ref")], env))
    parts.append("### 2. unmodified Core/fft_benchmark.cpp with CudaFFTBackend as getFFTBackend(): one transform per blocking "
                 "call, page-locked host buffers read and written in place by the kernel\n")
    parts.append(run([os.path.join(ROOT, "oracle", "_ref", "fft_benchmark_dropin")]))
    parts.append("### 3. CudaFFTBackend through the FFTBackend interface, host buffers (table 1: one per call; table 2: "
                 "4096 per call, copies inside every call, microseconds per transform)\n")
    parts.append(run([os.path.join(ROOT, "extio_sddc_b200", "bin", "fft_benchmark_cuda"), "4096", "1000"]))
    parts.append("### 4. device-resident\n")
    parts.append(device_table())
    text = "".join(parts)
    with open(os.path.join(OUT, "fft_benchmark.txt"), "w") as f:
        f.write(text)
    print(text)


if __name__ == "__main__":
    main()
