"""BASELINE config 5: long capture (19968 blocks = 1.31 G samples = 10.2 s at 128 Msps) time-sharded with
halfFft halos across the ranks (STRONG scaling: the capture is fixed, each rank takes 19968/N blocks), IQ
gathered to rank 0 over NCCL.  Prints one JSON line on rank 0.  Launch with torchrun for N > 1."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from bench import synth_capture, GAIN, TUNEBIN
from extio_sddc_b200 import CudaR2IQ, shard

rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
TOTAL = 19968
out_line = {"workload": "config5: 19968-block capture, time-sharded, IQ gathered to rank 0", "n_gpus": world, "results": []}
op = CudaR2IQ(device=local); op.Init(GAIN); op.updateRand(True); op.tunebin = TUNEBIN
s = torch.cuda.current_stream().cuda_stream
for d in (0, 5):
    first, count = shard.time_slabs(TOTAL, world, d)[rank]
    x = synth_capture(torch, dev, count, seed=888, first_block=first)
    opb = 32768 >> d
    mine = torch.empty(count * opb, dtype=torch.complex64, device=dev)
    sizes = [2 * c * opb for _, c in shard.time_slabs(TOTAL, world, d)]
    op.setDecimate(d); op.TurnOn()
    for _ in range(2):
        op.process_device(x.data_ptr(), count, mine.data_ptr(), s)
    if world > 1:   # bring up the point-to-point connections of the gather before timing it
        w = torch.zeros(16, device=dev)
        shard.gather_to_rank0(w, [16] * world, rank, world)
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record()
    op.process_device(x.data_ptr(), count, mine.data_ptr(), s)
    e1.record()
    got = shard.gather_to_rank0(torch.view_as_real(mine).reshape(-1), sizes, rank, world) if world > 1 else mine
    e2.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1), e0.elapsed_time(e2)], dtype=torch.float64, device=dev)
    if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        n = TOTAL * 65536
        out_line["results"].append({"d": d, "compute_ms": round(float(t[0]), 3), "compute_plus_gather_ms": round(float(t[1]), 3),
                                    "compute_gsamples_s": round(n / float(t[0]) / 1e6, 1),
                                    "gathered_gsamples_s": round(n / float(t[1]) / 1e6, 1),
                                    "iq_bytes_total": int(TOTAL * opb * 8)})
    del x, mine, got
if rank == 0:
    print(json.dumps(out_line), file=sys.stderr)
    open(os.path.join("gpurun_out", f"long_capture_{world}gpu.json"), "w").write(json.dumps(out_line) + "\n")
if world > 1:
    dist.destroy_process_group()
