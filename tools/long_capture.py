"""BASELINE config 5: long capture (19968 blocks = 1.31 G samples = 10.2 s at 128 Msps) time-sharded with
halfFft halos across the ranks (STRONG scaling: the capture is fixed, each rank takes 19968/N blocks).  The IQ of all
ranks lands on rank 0 through the fused kernel's own stores into a peer-mapped buffer (shard.PeerOutput); rank 0 then
recomputes two slabs and requires the gathered ranges to hold the same bits.  bench.py --gpus N carries the same block
in its JSON line (key "config5"); this stand-alone form also runs on one GPU.  Launch with torchrun for N > 1."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from bench import synth_capture, GAIN, TUNEBIN
from extio_sddc_b200 import CudaR2IQ, shard

rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
TOTAL = 19968
out_line = {"workload": "config5: 19968-block capture, time-sharded, IQ on rank 0 via peer stores", "n_gpus": world, "results": []}
op = CudaR2IQ(device=local); op.Init(GAIN); op.updateRand(True); op.tunebin = TUNEBIN
s = torch.cuda.current_stream().cuda_stream
rng = np.random.default_rng(5)
for d in (0, 5):
    slabs = shard.time_slabs(TOTAL, world, d)
    first, count = slabs[rank]
    opb = 32768 >> d
    x = synth_capture(torch, dev, count, seed=555, first_block=first)
    mine = torch.empty(count * opb, dtype=torch.complex64, device=dev)
    po = shard.PeerOutput(TOTAL * opb, rank, world, local)
    op.setDecimate(d); op.TurnOn()
    for _ in range(2):
        op.process_device(x.data_ptr(), count, mine.data_ptr(), s)
        op.process_device(x.data_ptr(), count, po.ptr(first * opb), s)
    po.fence()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    ev[0].record(); op.process_device(x.data_ptr(), count, mine.data_ptr(), s); ev[1].record()
    po.fence()
    ev[2].record(); op.process_device(x.data_ptr(), count, po.ptr(first * opb), s); ev[3].record()
    po.fence()
    t = torch.tensor([ev[0].elapsed_time(ev[1]), ev[2].elapsed_time(ev[3])], dtype=torch.float64, device=dev)
    if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        gathered = po.tensor()
        ok, worst, picks = True, 0.0, sorted({world - 1, int(rng.integers(0, world))})
        for r in picks:
            f, c = slabs[r]
            xin = synth_capture(torch, dev, c, seed=555, first_block=f)
            ref = torch.empty(c * opb, dtype=torch.complex64, device=dev)
            op.process_device(xin.data_ptr(), c, ref.data_ptr(), s); torch.cuda.synchronize()
            got = gathered[f * opb:(f + c) * opb]
            ok = ok and bool(torch.equal(torch.view_as_real(got), torch.view_as_real(ref)))
            worst = max(worst, float((torch.view_as_real(got) - torch.view_as_real(ref)).abs().max()))
        n = TOTAL * 65536
        out_line["results"].append({"d": d, "compute_ms": round(float(t[0]), 3), "compute_plus_gather_ms": round(float(t[1]), 3),
                                    "compute_gsamples_s": round(n / float(t[0]) / 1e6, 1),
                                    "gathered_gsamples_s": round(n / float(t[1]) / 1e6, 1),
                                    "iq_bytes_total": int(TOTAL * opb * 8), "verified": ok, "max_abs_diff": worst,
                                    "slabs_recomputed_on_rank0": picks})
    po.close()
    del x, mine
if rank == 0:
    print(json.dumps(out_line), file=sys.stderr)
    os.makedirs("gpurun_out", exist_ok=True)
    open(os.path.join("gpurun_out", f"long_capture_{world}gpu.json"), "w").write(json.dumps(out_line) + "\n")
if world > 1:
    dist.destroy_process_group()
