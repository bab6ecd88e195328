#!/usr/bin/env python
"""bench.py -- throughput of the r2iq hot path on B200, per the driver's contract.

    python bench.py --gpus N --steps K --warmup W            # our CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (rank 0 only)

Workload (BASELINE.json configs[1]): RX-888 Mk II full rate 128 Msps synthetic ADC capture, randomizer
on, decimation sweep 2/4/8/16/32/64 (d = 0..5), FFTN 2048.  One STEP = one pass of the fused kernel over
the whole capture at each of the six decimations (six launches).  `value` = input samples processed per
second over the whole job, inputs and outputs resident in HBM; `e2e` = the same sweep through the
host-buffer C-ABI call (r2iq_cuda_process) with pinned HOST buffers, H2D and D2H inside the timed region.
With N > 1 (torchrun, one rank per GPU) every rank processes its own time slab of the capture (slab +
halfFft halo, no data-path collective): weak scaling; the NCCL gather of IQ to rank 0 is timed
separately and reported under "gather".
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GAIN = 1.08e-8                 # RX888mk2_GAINFACTOR, Core/config.h:55
TUNEBIN = 256                  # setFreqOffset(0.25): SURVEY.md 8(d) config 2
DECIMS = (0, 1, 2, 3, 4, 5)    # decimation 2..64
SAMPLES_PER_BLOCK = 65536
METRIC = "r2iq input Gsamples/s"
UNIT = "Gsamples/s"
WORKLOAD = ("config2: RX-888 Mk II 128 Msps synthetic int16 ADC, randomizer on, decimation sweep "
            "2/4/8/16/32/64 (six fused-kernel launches per step), FFTN 2048, tune offset 0.25")


# The `config` both arms print, key for key: the workload only.  What differs per arm (blocks per GPU, instances ...)
# goes under "arm".
CONFIG = {"workload": WORKLOAD, "fftn": 2048, "adc_msps": 128, "randomizer": True, "decimation_indices": list(DECIMS),
          "tune_offset": 0.25, "samples_per_block": SAMPLES_PER_BLOCK}


def bytes_per_sample(d: int) -> float:
    """Algorithmic bytes per input sample: 2 B int16 in + 8 B complex64 out per 2^(d+1)."""
    return 2.0 + 8.0 / (1 << (d + 1))


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel: str, n_blocks: int):
    """dram__bytes_read + dram__bytes_write per launch from the committed `ncu --set full` capture of THIS kernel
    (profiles/traffic.json: {kernel name: {"blocks": B, "dram_bytes": N}}), scaled to this launch's block count; None
    when the kernel that ran has no capture (the name comes from the C ABI, not from a table here)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            t = json.load(f).get(kernel)
        return int(t["dram_bytes"] * n_blocks / t["blocks"]) if t else None
    except Exception:
        return None


def packed_fp32_per_pair(kernel: str):
    """Packed FP32 instructions (FFMA2 + FADD2 + FMUL2) one warp issues per frame pair in `kernel`, counted in the SASS of
    the library that is loaded (cuobjdump, loop body = everything after the prologue's block barrier).  None when
    cuobjdump is not on the box: the committed count in profiles/r02_sass_opcodes.txt documents the same number."""
    import re
    import shutil
    import subprocess
    from extio_sddc_b200 import capi
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe):
        return None
    m = re.match(r"r2iq_pair_kernel<(\d+), (\d+), (\d+)>", kernel)
    if not m:
        return None
    # the parameter list of the symbol depends on the decimation (FusedParams, or PairParams + PairNco: r2iq_pair.cu)
    stem = f"_ZN4r2iq16r2iq_pair_kernelILi{m.group(1)}ELi{m.group(2)}ELi{m.group(3)}EEEv"
    try:
        out = ""
        for mangled in (stem + "NS_11FusedParamsE", stem + "NS_10PairParamsENS_7PairNcoE"):
            out = subprocess.run([exe, "-sass", "-fun", mangled, capi.LIB_PATH], capture_output=True, text=True, timeout=120).stdout
            if "Function :" in out:
                break
    except Exception:
        return None
    # the persistent loop = the longest backward branch; count inside it only (the function also holds the prologue and
    # out-of-line wait loops)
    ins = [(int(a, 16), t) for a, t in re.findall(r"/\*([0-9a-f]{4,6})\*/\s+([^;]*);", out)]
    best = None
    for a, t in ins:
        m = re.search(r"\bBRA(?:\.\w+)*\s+(?:\w+,\s*)?0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a and (best is None or a - int(m.group(1), 16) > best[1] - best[0]):
            best = (int(m.group(1), 16), a)
    if best is None:
        return None
    n = sum(1 for a, t in ins if best[0] <= a <= best[1] and re.search(r"\b(?:FFMA2|FADD2|FMUL2)\b", t))
    return n or None


FP32_PEAK_THREAD_INST = 18.33e12   # packed FP32 thread-instructions/s, B200 at 1965 MHz: tools/ubench/fp32_pipes.cu, ffma2 at 32
                                   # warps/SM (profiles/r01_ubench_fp32_pipes.txt) = 64 lanes x 148 SMs x 1.935 GHz effective


# ------------------------------------------------------------------------------------------------
# clocks: NVML polled in a thread during the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._h = None

    def _loop(self):
        nv = self._nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                try:
                    bits = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
                except Exception:
                    bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                for b, name in self.REASONS.items():
                    if bits & b and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self._h is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self):
        if self._thr is not None:
            self._stop.set()
            self._thr.join()
        s = sorted(self.samples)
        med = s[len(s) // 2] if s else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


# ------------------------------------------------------------------------------------------------
# the reference's CPU path (oracle/_ref = the unmodified reference operator; test infrastructure,
# used here only as the timed baseline)
# ------------------------------------------------------------------------------------------------
def cpu_signal(n_blocks: int):
    from oracle import r2iq_oracle as orc
    return orc.rx888_blocks(n_blocks, 0, TUNEBIN, randomized=True, seed=888)


def reference_sweep(blocks_per_instance: int, instances: int, variant: str = "dispatch", warm: int = 64, fftn: int = 2048):
    """One bounded sample of the workload on the host cores: for every decimation of the sweep, each of
    `instances` independent reference operators processes `blocks_per_instance` blocks (saturated input
    ring, like unittest/benchmark_test.cpp:78-142).  Returns (Gsamples/s over the sweep, seconds, detail)."""
    from oracle import ref_lib
    R = ref_lib.get(fftn)
    tunebin = TUNEBIN * (fftn // 2048)
    backend = "MKL DFTI float32 (torch libtorch_cpu.so)" if R.set_fft_backend(ref_lib.BACKEND_MKL) == 0 else \
        "portable float64 FFT (MKL unavailable)"
    sig = cpu_signal(16)
    tot_s, tot_t, detail = 0.0, 0.0, []
    for d in DECIMS:
        sps, sec = R.bench(sig, d, tunebin, warm_blocks=warm, timed_blocks=blocks_per_instance, gain=GAIN, rand=True,
                           variant=variant, instances=instances)
        n = sps * sec
        tot_s += n
        tot_t += sec
        detail.append({"d": d, "gsamples_s": round(sps / 1e9, 4)})
    return tot_s / tot_t / 1e9, tot_t, detail, backend


def pick_instances() -> int:
    """The reference operator is single-worker (N_MAX_R2IQ_THREADS 1) and each instance also needs a producer
    and a consumer thread, so `nproc` instances oversubscribe the box.  Take whichever of nproc and nproc/2
    instances moves more samples on a short probe -- the better number for the CPU arm."""
    cores = os.cpu_count() or 1
    best, best_v = cores, -1.0
    for inst in sorted({cores, max(1, cores // 2)}):
        v, _, _, _ = reference_sweep(128, inst)
        if v > best_v:
            best, best_v = inst, v
    return best


def run_reference(args, rank: int, world: int):
    if rank != 0:
        return 0
    from oracle import ref_lib
    if not ref_lib.available(2048) and not ref_lib.build():
        _emit({"impl": "reference", "unavailable": "oracle/_ref is not built and /root/reference is absent"})
        return 0
    cores = pick_instances()
    blocks = args.ref_blocks
    for _ in range(args.warmup):
        reference_sweep(max(64, blocks // 8), cores)
    vals, secs = [], 0.0
    backend, detail = "", []
    for _ in range(args.steps):
        v, s, detail, backend = reference_sweep(blocks, cores)
        vals.append(v)
        secs += s
    samples_per_step = cores * blocks * SAMPLES_PER_BLOCK * len(DECIMS)
    value = samples_per_step * args.steps / secs / 1e9
    sample = (f"{cores} independent fft_mt_r2iq instances (reference ships N_MAX_R2IQ_THREADS=1 per operator) x "
              f"{blocks} blocks per decimation per step, saturated input ring; FFT behind the fftw3 shim: {backend}; "
              "FFTW itself is not installed on this box")
    line = {"impl": "reference", "metric": METRIC, "value": round(value, 4), "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(secs / args.steps * 1e3, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": CONFIG,
            "arm": {"sample_blocks_per_instance": blocks, "instances": cores, "host_cpus": os.cpu_count()},
            "cpu_baseline": {"value": round(value, 4), "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": round(value, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "per_decimation": detail, "gpu_launches": 0}
    _emit(line)
    return 0


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def synth_capture(torch, dev, n_blocks: int, seed: int, first_block: int = 0):
    """Synthetic RX-888 capture generated on the device: three tones (one in every decimation's pass-band
    around the tune bin, two out of band) + uniform noise, rounded to int16, then passed through the ADC
    randomizer (x ^= 0xFFFE when the LSB is set).  Layout: [halfFft halo | n_blocks blocks]; this rank's
    slab starts at block `first_block` of the global capture (tones are phase-continuous in the global
    sample index); the halo holds the 1024 samples before the slab, zeros for the slab that starts at 0."""
    n = n_blocks * SAMPLES_PER_BLOCK
    buf = torch.zeros(1024 + n, dtype=torch.int16, device=dev)
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    chunk = 1 << 24
    origin = first_block * SAMPLES_PER_BLOCK
    f_tune = TUNEBIN / 2048.0
    tones = ((8000.0, f_tune + 0.1 / 256, 0.0), (4000.0, (f_tune + 0.23) % 0.5, 0.7), (4000.0, (f_tune + 0.37) % 0.5, 1.9))
    lo = -1024 if first_block > 0 else 0
    for s in range(lo, n, chunk):
        m = min(chunk, n - s)
        t = torch.arange(origin + s, origin + s + m, device=dev, dtype=torch.float64)
        x = (torch.rand(m, device=dev, generator=g, dtype=torch.float32) * 1024.0 - 512.0).to(torch.float64)
        for a, f, ph in tones:
            x += a * torch.sin(2.0 * torch.pi * ((t * f) % 1.0) + ph)
        v = torch.round(x).to(torch.int32)
        v = torch.where((v & 1) != 0, v ^ 0xFFFE, v)          # randomizer on the wire
        v = ((v + 32768) & 0xFFFF) - 32768
        buf[1024 + s: 1024 + s + m] = v.to(torch.int16)
    return buf


def run_ours(args, rank: int, local_rank: int, world: int):
    import numpy as np
    import torch
    import torch.distributed as dist
    from extio_sddc_b200 import CudaR2IQ, PinnedBuffer, capi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the r2iq backend has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    n_blocks = args.blocks
    samples_per_launch = n_blocks * SAMPLES_PER_BLOCK
    x = synth_capture(torch, dev, n_blocks, seed=888 + rank, first_block=rank * n_blocks)
    out = torch.empty(n_blocks * 32768, dtype=torch.complex64, device=dev)
    op = CudaR2IQ(device=local_rank)
    op.Init(GAIN)
    op.updateRand(True)
    op.setSideband(False)
    op.tunebin = TUNEBIN
    stream = torch.cuda.current_stream().cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()

    def step(events=None):
        for d in DECIMS:
            op.setDecimate(d)
            op.TurnOn()                                   # latches the decimation (impl.hpp:3-6)
            if events is not None:
                e0 = torch.cuda.Event(enable_timing=True)
                e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
            op.process_device(x.data_ptr(), n_blocks, out.data_ptr(), stream)
            if events is not None:
                e1.record()
                events[d].append((e0, e1))

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    launches_before = op.stats().kernel_launches
    sampler = ClockSampler(local_rank)
    events = {d: [] for d in DECIMS}
    barrier()
    torch.cuda.synchronize()
    sampler.start()
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(args.steps):
        step(events)
    t1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    barrier()
    ms_total = t0.elapsed_time(t1)
    launches = int(op.stats().kernel_launches - launches_before)
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
        lt = torch.tensor([launches], dtype=torch.int64, device=dev)
        dist.all_reduce(lt, op=dist.ReduceOp.SUM)
        launches = int(lt.item())
    samples_per_step = samples_per_launch * len(DECIMS) * world
    value = samples_per_step * args.steps / (ms_total * 1e-3) / 1e9

    peak, peak_src = measured_peak()
    per_d = []
    for d in DECIMS:
        ms = sum(a.elapsed_time(b) for a, b in events[d]) / len(events[d])
        gbs = samples_per_launch * bytes_per_sample(d) / (ms * 1e-3) / 1e9
        per_d.append({"d": d, "decimation": 1 << (d + 1), "kernel_ms": round(ms, 4),
                      "gsamples_s": round(samples_per_launch / (ms * 1e-3) / 1e9, 2),
                      "alg_bytes_per_sample": bytes_per_sample(d), "achieved_gbs": round(gbs, 1),
                      "hbm_frac": round(gbs / peak, 4)})
    dom = max(per_d, key=lambda r: r["kernel_ms"])
    for r in per_d:
        r["kernel"] = op.kernel_name(r["d"])
    dom_kernel = dom["kernel"]
    roofline = {"bound": "hbm", "kernel": f"{dom_kernel} (decimation {dom['decimation']})",
                "achieved": dom["achieved_gbs"], "peak": peak, "unit": "GB/s", "frac": round(dom["achieved_gbs"] / peak, 4),
                "traffic": ncu_traffic(dom_kernel, n_blocks), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": samples_per_launch * bytes_per_sample(dom["d"])}
    # FP32 co-roof: this FFT path is bound by the FP32 pipe and what feeds it, not by HBM
    if rank == 0:
        fp = packed_fp32_per_pair(dom_kernel)
        if fp:
            pairs_per_s = (n_blocks * 43 / 2) / (dom["kernel_ms"] * 1e-3)
            ach = fp * 32 * pairs_per_s
            roofline["fp32"] = {"packed_fp32_inst_per_frame_pair": fp,
                                "packed_fp32_inst_per_sample": round(fp * 32 / 3072.0, 3),   # 32 lanes; a pair adds 2 x 1536 new samples
                                "achieved_thread_inst_per_s": round(ach / 1e12, 3), "peak_thread_inst_per_s": FP32_PEAK_THREAD_INST / 1e12,
                                "unit": "T packed-FP32 thread-instructions/s", "frac": round(ach / FP32_PEAK_THREAD_INST, 4),
                                "achieved_tflops": round(ach * 3.0 / 1e12, 1),   # ~1.5 flop per lane-op (FFMA2 : FADD2+FMUL2 = 1 : 1), 2 lanes
                                "peak_source": "tools/ubench/fp32_pipes.cu on B200 (profiles/r01_ubench_fp32_pipes.txt)"}

    # ---- e2e: same sweep through the host-buffer C-ABI call, pinned host memory -------------------
    e2e = None
    if not args.no_e2e:
        hin = PinnedBuffer(samples_per_launch, np.int16, device=local_rank)
        hout = PinnedBuffer(n_blocks * 32768, np.complex64, device=local_rank)
        torch.cuda.synchronize()
        hin.array[:] = x[1024:].cpu().numpy()
        eop = CudaR2IQ(device=local_rank, chunk_blocks=args.chunk_blocks)
        eop.Init(GAIN)
        eop.updateRand(True)
        eop.tunebin = TUNEBIN
        zero_hist = np.zeros(1024, dtype=np.int16)

        def e2e_step_per_operator():
            # one operator per decimation, as RadioHandlerClass drives the reference: the capture crosses PCIe
            # once per decimation
            acc = 0.0
            for d in DECIMS:
                eop.setDecimate(d)
                eop.TurnOn()
                y = eop.process(hin.array, history=zero_hist, out=hout.array[: n_blocks * (32768 >> d)])
                acc += float(y[0].real)                   # touch the result on the host
            return acc

        houts = [PinnedBuffer(n_blocks * (32768 >> d), np.complex64, device=local_rank) for d in DECIMS]

        def e2e_step_rates():
            # the same sweep through r2iq_cuda_process_rates: the capture crosses PCIe once (the reference's
            # operators read one input ring in place), every decimation's IQ comes back to the host
            eop.TurnOn()
            ys = eop.process_rates(hin.array, DECIMS, history=zero_hist, outs=[b.array for b in houts])
            return sum(float(y[0].real) for y in ys)

        def timed(fn):
            fn()
            st0 = eop.stats()
            barrier()
            torch.cuda.synchronize()
            w0 = time.perf_counter()
            for _ in range(args.e2e_steps):
                fn()
            torch.cuda.synchronize()
            w1 = time.perf_counter()
            barrier()
            st1 = eop.stats()
            sec = w1 - w0
            if world > 1:
                t = torch.tensor([sec], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                sec = float(t.item())
            return {"value": round(samples_per_step * args.e2e_steps / sec / 1e9, 3), "unit": UNIT,
                    "h2d_bytes_per_step": int((st1.h2d_bytes - st0.h2d_bytes) // args.e2e_steps),
                    "d2h_bytes_per_step": int((st1.d2h_bytes - st0.d2h_bytes) // args.e2e_steps),
                    "steps": args.e2e_steps, "ms_per_step": round(sec / args.e2e_steps * 1e3, 2)}

        eop.setDecimate(0)
        e2e = timed(e2e_step_rates)
        e2e["api"] = ("r2iq_cuda_process_rates (C ABI, pinned host buffers, chunked pipeline): one H2D of the capture, "
                      "six decimations on the device, every IQ stream copied back")
        e2e["host_numa_node"] = int(capi.load().r2iq_cuda_device_numa_node(local_rank))   # rank 0's; -1 = default placement
        e2e["per_operator"] = timed(e2e_step_per_operator)
        e2e["per_operator"]["api"] = ("r2iq_cuda_process once per decimation (one operator per rate): the capture is "
                                      "copied to the device six times")
        # the host-buffer path and the device-resident path give the same bits (same kernels, same capture)
        op.setDecimate(DECIMS[-1])
        op.TurnOn()
        op.process_device(x.data_ptr(), n_blocks, out.data_ptr(), stream)
        torch.cuda.synchronize()
        n_last = n_blocks * (32768 >> DECIMS[-1])
        dev_last = out[:n_last].cpu().numpy()
        e2e["verified"] = bool(np.array_equal(dev_last, houts[-1].array)) and bool(np.array_equal(dev_last, hout.array[:n_last]))
        for b in houts:
            b.free()
        # the device-resident and host-buffer paths must agree on the last decimation
        eop.TurnOff()
        eop.close()
        hin.free()
        hout.free()

    # ---- IQ of all ranks on rank 0 (north star: the only exchange on the path) --------------------------------
    gather = None
    config5 = None
    if world > 1:
        from extio_sddc_b200 import shard
        gather = []
        warm = torch.zeros(1024, dtype=torch.float32, device=dev)      # bring the NCCL communicator up untimed
        dist.gather(warm, [torch.empty_like(warm) for _ in range(world)] if rank == 0 else None, dst=0)
        torch.cuda.synchronize()
        check_rng = np.random.default_rng(2024)

        def verify(gathered, d, slabs, regen):
            """Rank 0 recomputes the last rank's slab and one more picked at random from the same seeded generator and
            requires the gathered ranges to hold the same bits."""
            picks = sorted({world - 1, int(check_rng.integers(0, world))})
            worst = 0.0
            ok = True
            for r in picks:
                first, count = slabs[r]
                xin = regen(r)
                ref = torch.empty(count * (32768 >> d), dtype=torch.complex64, device=dev)
                op.process_device(xin.data_ptr(), count, ref.data_ptr(), stream)
                torch.cuda.synchronize()
                got = gathered[first * (32768 >> d):(first + count) * (32768 >> d)]
                ok = ok and bool(torch.equal(torch.view_as_real(got), torch.view_as_real(ref)))
                worst = max(worst, float((torch.view_as_real(got) - torch.view_as_real(ref)).abs().max()))
                del xin, ref
            return ok, worst, picks

        weak_slabs = [(r * n_blocks, n_blocks) for r in range(world)]
        for d in (DECIMS[-1], DECIMS[0]):
            n_out = n_blocks * (32768 >> d)
            op.setDecimate(d)
            op.TurnOn()
            # (1) NCCL gather after the kernel (round 1's form, kept as the comparison)
            mine = torch.view_as_real(out[:n_out])
            dst = [torch.empty_like(mine) for _ in range(world)] if rank == 0 else None
            barrier()
            torch.cuda.synchronize()
            g0 = torch.cuda.Event(enable_timing=True)
            g1 = torch.cuda.Event(enable_timing=True)
            g0.record()
            op.process_device(x.data_ptr(), n_blocks, out.data_ptr(), stream)
            dist.gather(mine, dst, dst=0)
            g1.record()
            torch.cuda.synchronize()
            t = torch.tensor([g0.elapsed_time(g1)], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            entry = {"d": d, "gathered_bytes": int(n_out * 8 * (world - 1)),
                     "nccl_after_kernel_ms": round(float(t.item()), 3)}
            if rank == 0:
                allg = torch.view_as_complex(torch.cat(dst).view(-1, 2))
                ok, worst, picks = verify(allg, d, weak_slabs, lambda r: synth_capture(torch, dev, n_blocks, seed=888 + r, first_block=r * n_blocks))
                entry["nccl_verified"] = ok
                del allg
            del dst
            # (2) peer stores: every rank's kernel writes its IQ into rank 0's buffer over NVLink while it computes
            po = shard.PeerOutput(world * n_out, rank, world, local_rank)
            op.process_device(x.data_ptr(), n_blocks, po.ptr(rank * n_out), stream)      # warm the mapping
            po.fence()
            g0 = torch.cuda.Event(enable_timing=True)
            g1 = torch.cuda.Event(enable_timing=True)
            g0.record()
            op.process_device(x.data_ptr(), n_blocks, po.ptr(rank * n_out), stream)
            g1.record()
            po.fence()
            t = torch.tensor([g0.elapsed_time(g1)], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            entry["ms_compute_plus_gather"] = round(float(t.item()), 3)
            entry["method"] = "fused kernel stores into rank 0's buffer through a CUDA-IPC peer mapping (NVLink), no gather after the kernel"
            entry["into_rank0_gbs"] = round(entry["gathered_bytes"] / (float(t.item()) * 1e-3) / 1e9, 1)
            if rank == 0:
                ok, worst, picks = verify(po.tensor(), d, weak_slabs, lambda r: synth_capture(torch, dev, n_blocks, seed=888 + r, first_block=r * n_blocks))
                entry.update({"verified": ok, "max_abs_diff": worst, "slabs_recomputed_on_rank0": picks})
            po.close()
            gather.append(entry)

        # ---- config 5 (BASELINE.json configs[4]): ONE 19968-block capture, time-sharded with halos: strong scaling ----
        config5 = {"workload": "config5: 10.2 s capture (19968 blocks = 1.31 G samples) time-sharded with halfFft halos, IQ on rank 0",
                   "total_blocks": 19968, "scaling": "strong", "results": []}
        for d in (DECIMS[0], DECIMS[-1]):
            slabs = shard.time_slabs(19968, world, d)
            first, count = slabs[rank]
            opb = 32768 >> d
            xs = synth_capture(torch, dev, count, seed=555, first_block=first)
            local_out = torch.empty(count * opb, dtype=torch.complex64, device=dev)
            op.setDecimate(d)
            op.TurnOn()
            po = shard.PeerOutput(19968 * opb, rank, world, local_rank)
            for _ in range(2):
                op.process_device(xs.data_ptr(), count, local_out.data_ptr(), stream)
                op.process_device(xs.data_ptr(), count, po.ptr(first * opb), stream)
            po.fence()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            ev[0].record()
            op.process_device(xs.data_ptr(), count, local_out.data_ptr(), stream)
            ev[1].record()
            po.fence()
            ev[2].record()
            op.process_device(xs.data_ptr(), count, po.ptr(first * opb), stream)
            ev[3].record()
            po.fence()
            t = torch.tensor([ev[0].elapsed_time(ev[1]), ev[2].elapsed_time(ev[3])], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res = {"d": d, "blocks_per_gpu": count, "compute_ms": round(float(t[0]), 3), "compute_plus_gather_ms": round(float(t[1]), 3),
                   "compute_gsamples_s": round(19968 * 65536 / float(t[0]) / 1e6, 1),
                   "gathered_gsamples_s": round(19968 * 65536 / float(t[1]) / 1e6, 1), "iq_bytes_total": int(19968 * opb * 8)}
            if rank == 0:
                ok, worst, picks = verify(po.tensor(), d, slabs, lambda r: synth_capture(torch, dev, slabs[r][1], seed=555, first_block=slabs[r][0]))
                res.update({"verified": ok, "max_abs_diff": worst, "slabs_recomputed_on_rank0": picks})
            config5["results"].append(res)
            po.close()
            del xs, local_out

    # ---- config 4: 64-channel channelizer at d = 5 ------------------------------------------------------------
    # N > 1: the capture is cut into TIME SLABS (shard.time_slabs, the scheme of the headline path) and every rank
    # produces all 64 channels of its slab, so no forward FFT is computed twice; the channel-sharded alternative
    # (every rank transforms the whole capture for 64/N channels: shard.channel_shards) is timed beside it.
    channelizer = None
    if not args.no_channelizer:
        from extio_sddc_b200 import shard
        all_bins = shard.channel_tunebins(64)
        cd = 5
        op.setDecimate(cd)
        op.TurnOn()

        def time_channels(bins, first_block, blocks, out_t):
            src = x.data_ptr() + first_block * 65536 * 2      # [history = tail of the previous block | blocks]
            for _ in range(2):
                op.process_channels_device(src, blocks, bins, out_t.data_ptr(), stream)
            barrier()
            torch.cuda.synchronize()
            c0 = torch.cuda.Event(enable_timing=True)
            c1 = torch.cuda.Event(enable_timing=True)
            c0.record()
            reps = 5
            for _ in range(reps):
                op.process_channels_device(src, blocks, bins, out_t.data_ptr(), stream)
            c1.record()
            torch.cuda.synchronize()
            ms = c0.elapsed_time(c1) / reps
            if world > 1:
                t = torch.tensor([ms], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            return ms

        first, mine_blocks = shard.time_slabs(n_blocks, world, cd)[rank]
        per_ch = mine_blocks * (32768 >> cd)
        cout = torch.empty(64 * per_ch, dtype=torch.complex64, device=dev)
        cms = time_channels(all_bins, first, mine_blocks, cout)
        # every channel equals a separate operator tuned to its bin (Core/fft_mt_r2iq_impl.hpp:76-138 per channel): check
        # the first and the last channel of this rank's slab against r2iq_cuda_process_device on the same slab
        chan_ok = True
        one = torch.empty(per_ch, dtype=torch.complex64, device=dev)
        for ci in (0, 63):
            op.tunebin = all_bins[ci]
            op.process_device(x.data_ptr() + first * 65536 * 2, mine_blocks, one.data_ptr(), stream)
            torch.cuda.synchronize()
            a = torch.view_as_real(cout[ci * per_ch:(ci + 1) * per_ch])
            b = torch.view_as_real(one)
            rel = float(torch.linalg.vector_norm(a - b) / torch.linalg.vector_norm(b))
            chan_ok = chan_ok and rel <= 1e-5
        op.tunebin = TUNEBIN
        del one, cout
        cms_ch = None
        if world > 1:
            t = torch.tensor([1.0 if chan_ok else 0.0], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            chan_ok = bool(t.item() > 0.5)
            mine = [all_bins[c] for c in shard.channel_shards(64, world)[rank]]
            cout = torch.empty(len(mine) * n_blocks * (32768 >> cd), dtype=torch.complex64, device=dev)
            cms_ch = time_channels(mine, 0, n_blocks, cout)
            del cout
        channelizer = {"channels": 64, "d": cd, "decimation": 64, "blocks": n_blocks, "ms": round(cms, 3),
                       "sharding": "time slabs x all channels" if world > 1 else "none",
                       "verified": chan_ok,
                       "input_gsamples_s": round(samples_per_launch / (cms * 1e-3) / 1e9, 2),
                       "channel_gsamples_s": round(64 * samples_per_launch / (cms * 1e-3) / 1e9, 1),
                       "alg_bytes_per_input_sample": 10.0,
                       "hbm_frac": round(10.0 * samples_per_launch / world / (cms * 1e-3) / 1e9 / peak, 4),
                       "kernel": "r2iq_channels_lane_kernel<5>",
                       "ms_channel_sharded": None if cms_ch is None else round(cms_ch, 3),
                       "note": "one forward FFT per frame shared by all 64 channels; at N > 1 no rank repeats another's transforms"}

    # ---- FFTN 8192 (the reference tree as shipped, Core/config.h:49): the same sweep on the wide kernel family ----
    wide = None
    if not args.no_fftn8192:
        wop = CudaR2IQ(device=local_rank, fftn=8192)
        wop.Init(GAIN)
        wop.updateRand(True)
        wop.tunebin = TUNEBIN * 4
        wblocks = n_blocks - 1                       # the first 4096 samples of the capture serve as the halo
        wper = []
        wtot = 0.0
        for d in DECIMS:
            wop.setDecimate(d)
            wop.TurnOn()
            for _ in range(2):
                wop.process_device(x.data_ptr(), wblocks, out.data_ptr(), stream)
            barrier()
            torch.cuda.synchronize()
            w0 = torch.cuda.Event(enable_timing=True)
            w1 = torch.cuda.Event(enable_timing=True)
            w0.record()
            for _ in range(3):
                wop.process_device(x.data_ptr(), wblocks, out.data_ptr(), stream)
            w1.record()
            torch.cuda.synchronize()
            wms = w0.elapsed_time(w1) / 3
            if world > 1:
                t = torch.tensor([wms], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                wms = float(t.item())
            wtot += wms
            wper.append({"d": d, "ms": round(wms, 4), "gsamples_s": round(world * wblocks * 65536 / (wms * 1e-3) / 1e9, 1)})
            wop.TurnOff()
        wide = {"fftn": 8192, "value": round(world * len(DECIMS) * wblocks * 65536 / (wtot * 1e-3) / 1e9, 1), "unit": UNIT,
                "per_decimation": wper, "note": "same capture and sweep on a handle created with fftn = 8192 (halfFft 4096)"}
        if not args.no_channelizer:
            # config 4 at the shipped FFTN: 64 channels at d = 5 from one forward transform per frame pair
            # (r2iq_wide_channels.cu); one channel checked against the single-channel kernel on the same capture
            wd = 5
            wbins = [(64 * c) % 4096 for c in range(64)]
            wper_ch = wblocks * (32768 >> wd)
            wout = torch.empty(64 * wper_ch, dtype=torch.complex64, device=dev)
            wop.setDecimate(wd)
            wop.TurnOn()
            for _ in range(2):
                wop.process_channels_device(x.data_ptr(), wblocks, wbins, wout.data_ptr(), stream)
            torch.cuda.synchronize()
            w0 = torch.cuda.Event(enable_timing=True)
            w1 = torch.cuda.Event(enable_timing=True)
            w0.record()
            for _ in range(3):
                wop.process_channels_device(x.data_ptr(), wblocks, wbins, wout.data_ptr(), stream)
            w1.record()
            torch.cuda.synchronize()
            wcms = w0.elapsed_time(w1) / 3
            one = torch.empty(wper_ch, dtype=torch.complex64, device=dev)
            wop.tunebin = wbins[17]
            wop.process_device(x.data_ptr(), wblocks, one.data_ptr(), stream)
            torch.cuda.synchronize()
            a = torch.view_as_real(wout[17 * wper_ch:18 * wper_ch])
            b = torch.view_as_real(one)
            wrel = float(torch.linalg.vector_norm(a - b) / torch.linalg.vector_norm(b))
            wop.TurnOff()
            wide["channelizer"] = {"channels": 64, "d": wd, "blocks_per_gpu": wblocks, "ms": round(wcms, 3), "verified": wrel <= 1e-5,
                                   "input_gsamples_s": round(wblocks * 65536 / (wcms * 1e-3) / 1e9, 2),
                                   "note": "every rank its own capture (no sharding); one single-channel launch per channel costs 64 x the d = 5 time above"}
            del wout, one
        wop.close()

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cores = pick_instances()
            v, sec, detail, backend = reference_sweep(args.ref_blocks, cores)
            v1, sec1, _, _ = reference_sweep(max(256, args.ref_blocks // 2), 1)
            v2, _, _, _ = reference_sweep(max(256, args.ref_blocks // 2), 1, variant="avx2")
            cpu_baseline = {"value": round(v, 4), "unit": UNIT, "cores": cores, "kind": "reference",
                            "sample": (f"oracle/_ref (unmodified fft_mt_r2iq loop over the fftw3 shim; {backend}): {cores} "
                                       f"independent operator instances x {args.ref_blocks} blocks per decimation, "
                                       f"same sweep d=0..5, {sec:.1f} s of CPU time window"),
                            "single_instance_as_shipped": round(v1, 4),        # cpuid dispatcher's pick (avx512 TU on this host)
                            "single_instance_forced_avx2_intrinsics": round(v2, 4), "per_decimation": detail}
            if wide is not None:
                v8, _, _, _ = reference_sweep(max(256, args.ref_blocks // 4), cores, fftn=8192)
                wide["cpu_reference_gsamples_s"] = round(v8, 4)        # oracle/_ref built at FFTN_R_ADC 8192, same cores
        except Exception as ex:  # the baseline is reported, never required for the GPU arm
            cpu_baseline = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"unavailable: {ex}"}

    op.TurnOff()
    op.close()
    if rank == 0:
        line = {"metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": round(ms_total / args.steps, 4), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": CONFIG,
                "arm": {"blocks_per_gpu": n_blocks, "input_samples_per_step": samples_per_step,
                        "parallelism": f"time-slab x{world} (slab + halfFft halo per GPU, no data-path collective)",
                        "l2": f"input capture {samples_per_launch * 2 / 1e9:.2f} GB per GPU is larger than the 126 MB L2; "
                              "no explicit flush"},
                "roofline": roofline, "per_decimation": per_d, "clocks": clocks, "gpu_launches": launches}
        if e2e is not None:
            line["e2e"] = e2e
        if gather is not None:
            line["gather"] = gather
        if config5 is not None:
            line["config5"] = config5
        if channelizer is not None:
            line["channelizer"] = channelizer
        if wide is not None:
            line["fftn8192"] = wide
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
        _emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def _emit(line: dict) -> None:
    """Print the ONE JSON line on the real stdout (see main(): fd 1 is parked on stderr while the run is
    in progress so that library banners -- NCCL prints its version on stdout -- cannot precede it)."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--blocks", type=int, default=8192, help="65536-sample blocks per GPU (8192 = 4.2 s at 128 Msps)")
    ap.add_argument("--chunk-blocks", type=int, default=256)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--ref-blocks", type=int, default=2048, help="blocks per instance per decimation for the CPU arm")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-channelizer", action="store_true")
    ap.add_argument("--no-fftn8192", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank, world)
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("bench.py: --gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
    return run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    sys.exit(main())
