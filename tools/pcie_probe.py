"""PCIe / host-memory ceiling of the box for the e2e byte mix of bench.py, on 1..8 GPUs AT ONCE (launch with torchrun for
N > 1; one rank per GPU): plain pinned cudaMemcpyAsync, no kernels.
  * D2H alone, H2D alone, both directions at once (1 GiB per copy) -- per GPU and summed over the GPUs;
  * the e2e step's own mix: 1.07 GB up overlapped with 4.23 GB down per GPU (bench.py `e2e`, one upload, six rates).
The aggregate over N GPUs is the ceiling the e2e pipeline is measured against.  Prints one JSON line on
rank 0 and writes gpurun_out/pcie_probe_<N>gpu.json."""
import json
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)

GIB = 1 << 30
UP, DOWN = 1073743872, 4227858432          # bytes per e2e step and GPU: 8192 blocks in, six IQ streams out
h_in = torch.empty(max(GIB, UP), dtype=torch.uint8).pin_memory()
h_out = torch.empty(max(GIB, DOWN), dtype=torch.uint8).pin_memory()
h_in.fill_(1)                              # first touch by the owning rank
h_out.fill_(0)
d_a = torch.empty(max(GIB, DOWN), dtype=torch.uint8, device=dev)
d_b = torch.empty(max(GIB, UP), dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    if world > 1:
        dist.barrier()


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    sec = (time.perf_counter() - t0) / reps
    barrier()
    if world > 1:                          # the slowest rank sets the aggregate
        t = torch.tensor([sec], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sec = float(t.item())
    return sec


def d2h(n=GIB):
    with torch.cuda.stream(s1):
        h_out[:n].copy_(d_a[:n], non_blocking=True)


def h2d(n=GIB):
    with torch.cuda.stream(s2):
        d_b[:n].copy_(h_in[:n], non_blocking=True)


rows = []
for name, fn, up, down in (("D2H alone", d2h, 0, GIB), ("H2D alone", h2d, GIB, 0),
                           ("both directions, 1 GiB each", lambda: (d2h(), h2d()), GIB, GIB),
                           ("e2e mix: 1.07 GB up + 4.23 GB down", lambda: (d2h(DOWN), h2d(UP)), UP, DOWN)):
    sec = timed(fn)
    rows.append({"case": name, "ms": round(sec * 1e3, 2), "per_gpu_up_gbs": round(up / sec / 1e9, 1),
                 "per_gpu_down_gbs": round(down / sec / 1e9, 1), "box_total_gbs": round(world * (up + down) / sec / 1e9, 1)})
if rank == 0:
    mix = rows[-1]
    # what the e2e sweep could reach if the copies were all there is: 6 decimations x 8192 blocks x 65536 samples per GPU
    line = {"n_gpus": world, "host_cpus": os.cpu_count(), "cases": rows,
            "e2e_ceiling_gsamples_s": round(world * 6 * 8192 * 65536 / (mix["ms"] * 1e-3) / 1e9, 1)}
    print(json.dumps(line))
    os.makedirs("gpurun_out", exist_ok=True)
    with open(os.path.join("gpurun_out", f"pcie_probe_{world}gpu.json"), "w") as f:
        f.write(json.dumps(line) + "\n")
if world > 1:
    dist.destroy_process_group()
