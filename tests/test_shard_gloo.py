"""CPU suite, part 4: the N > 1 host logic over gloo (world_size 2), no GPU.

Time-slab sharding with a halfFft halo and channel sharding must reproduce the unsharded result
exactly once gathered on rank 0.  The per-rank compute is the oracle here (test infrastructure
standing in for the CUDA operator, which has the same [halo | blocks] -> IQ contract)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from extio_sddc_b200 import shard
from oracle import r2iq_oracle as orc

GAIN = orc.RX888MK2_GAINFACTOR


def test_time_slabs_properties():
    for n_blocks in (0, 1, 7, 64, 19968):
        for world in (1, 2, 3, 8):
            for d in (0, 3, 5):
                slabs = shard.time_slabs(n_blocks, world, d)
                assert len(slabs) == world
                assert slabs[0][0] == 0 and sum(c for _, c in slabs) == n_blocks
                for (f0, c0), (f1, _) in zip(slabs, slabs[1:]):
                    assert f1 == f0 + c0
                    assert f1 % (1 << d) == 0 or f1 == n_blocks
    # config 5: 19968 blocks over 8 GPUs is a whole number of output slots per rank at every d
    assert all(c == 2496 for _, c in shard.time_slabs(19968, 8, 5))
    assert shard.channel_tunebins(64)[0] == 8 and len(set(shard.channel_tunebins(64))) == 64
    assert [len(r) for r in shard.channel_shards(64, 8)] == [8] * 8


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        d, tb = 2, 256
        capture = orc.rx888_blocks(12, d, tb, randomized=True, seed=42)
        hist = np.random.default_rng(9).integers(-1000, 1000, 1024).astype(np.int16)
        if mode == "time":
            def process(slab, n):
                return orc.r2iq(slab[1024:], d, tb, gain=GAIN, rand=True, history=slab[:1024])
            local = shard.run_time_sharded(process, capture, d, rank, world, history=hist)
            sizes = [2 * c * (32768 >> d) for _, c in shard.time_slabs(12, world, d)]
        else:
            chans = shard.channel_tunebins(6)
            mine = shard.channel_shards(6, world)[rank]
            local = np.concatenate([orc.r2iq(capture, d, chans[c], gain=GAIN, rand=True, history=hist) for c in mine])
            sizes = [2 * len(r) * 12 * (32768 >> d) for r in shard.channel_shards(6, world)]
        t = torch.from_numpy(np.ascontiguousarray(local).view(np.float32).copy())
        got = shard.gather_to_rank0(t, sizes, rank, world)
        if rank == 0:
            q.put(got.numpy().view(np.complex64))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["time", "channel"])
def test_world2_gloo_matches_unsharded(mode):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    d, tb = 2, 256
    capture = orc.rx888_blocks(12, d, tb, randomized=True, seed=42)
    hist = np.random.default_rng(9).integers(-1000, 1000, 1024).astype(np.int16)
    if mode == "time":
        want = orc.r2iq(capture, d, tb, gain=GAIN, rand=True, history=hist)
    else:
        want = np.concatenate([orc.r2iq(capture, d, c, gain=GAIN, rand=True, history=hist) for c in shard.channel_tunebins(6)])
    assert got.shape == want.shape
    assert np.array_equal(got, want)
