"""Partitioning of r2iq work across the GPUs of one box (one process per GPU, torch.distributed).

The path shards without any exchange during compute:
  * by TIME SLAB -- every 65536-sample block depends only on itself and on the previous halfFft
    samples (Core/fft_mt_r2iq_impl.hpp:32,42), so rank r takes a contiguous run of blocks plus a
    halfFft-sample halo read from the slab before it; slab edges sit on multiples of 2^d blocks so
    each rank fills whole 32768-complex output slots (impl.hpp:100-148);
  * by CHANNEL -- independent tune bins over the same input (the channelizer workload).
The only collective is the gather of output IQ to one rank (NCCL over NVLink on GPUs; the same code
runs over gloo on CPU tensors in the tests).
"""
from __future__ import annotations

from typing import Callable, List, Sequence, Tuple

import numpy as np

HALF_FFT = 1024
BLOCK = 65536


def time_slabs(n_blocks: int, world: int, d: int) -> List[Tuple[int, int]]:
    """(first_block, count) per rank: contiguous, in rank order, edges on multiples of 2^d blocks,
    sizes differing by at most one output slot.  Ranks beyond the work get (n_blocks, 0)."""
    if n_blocks < 0 or world < 1 or not 0 <= d <= 6:
        raise ValueError("bad arguments")
    slot = 1 << d
    n_slots, tail = divmod(n_blocks, slot)
    base, extra = divmod(n_slots, world)
    out, first = [], 0
    for r in range(world):
        cnt = (base + (1 if r < extra else 0)) * slot
        if r == world - 1:
            cnt += tail                      # a ragged tail (partial output slot) stays with the last rank
        out.append((first, cnt))
        first += cnt
    return out


def slab_input(capture: np.ndarray, first_block: int, count: int, history: np.ndarray | None = None) -> np.ndarray:
    """[halo | blocks] for one rank: the halo is the halfFft samples before the slab (the capture's own
    history -- zeros if None -- for the rank that starts at block 0)."""
    capture = np.asarray(capture, dtype=np.int16).reshape(-1)
    out = np.empty(HALF_FFT + count * BLOCK, dtype=np.int16)
    s = first_block * BLOCK
    if first_block == 0:
        out[:HALF_FFT] = 0 if history is None else np.asarray(history, dtype=np.int16)
    else:
        out[:HALF_FFT] = capture[s - HALF_FFT:s]
    out[HALF_FFT:] = capture[s:s + count * BLOCK]
    return out


def channel_shards(n_channels: int, world: int) -> List[range]:
    """Channels c in [r*C/P, (r+1)*C/P) on rank r, remainder spread over the first ranks."""
    base, extra = divmod(n_channels, world)
    out, first = [], 0
    for r in range(world):
        cnt = base + (1 if r < extra else 0)
        out.append(range(first, first + cnt))
        first += cnt
    return out


def channel_tunebins(n_channels: int) -> List[int]:
    """Evenly spaced tune bins of the channelizer config: 4*round((c+0.5)*halfFft/C/4)."""
    return [int(4 * round((c + 0.5) * HALF_FFT / n_channels / 4)) for c in range(n_channels)]


def gather_to_rank0(local, sizes: Sequence[int], rank: int, world: int, group=None):
    """Gather per-rank 1-D tensors (float32 view of interleaved IQ) of possibly different lengths to
    rank 0, in rank order.  Returns the concatenated tensor on rank 0, None elsewhere."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return local
    # one grouped exchange (ncclGroupStart/End under NCCL): all peers' transfers are in flight together,
    # rank 0's NVLink ingest is the only serial resource
    if rank == 0:
        parts = [local] + [torch.empty(sizes[src], dtype=local.dtype, device=local.device) for src in range(1, world)]
        ops = [dist.P2POp(dist.irecv, parts[src], src, group) for src in range(1, world) if sizes[src]]
        for req in (dist.batch_isend_irecv(ops) if ops else []):
            req.wait()
        return torch.cat(parts)
    if sizes[rank]:
        for req in dist.batch_isend_irecv([dist.P2POp(dist.isend, local, 0, group)]):
            req.wait()
    return None


class PeerOutput:
    """The gathered IQ as ONE device buffer on rank 0 that every rank's kernel stores into directly (CUDA IPC mapping,
    peer stores over NVLink; include/r2iq_cuda.h "Output shared between the ranks").  ``n_complex`` = total complex64
    samples; ``ptr(offset_complex)`` is what a rank passes as ``d_out_ptr`` of ``process_device``.  Collective: every
    rank constructs it; the handle travels through torch.distributed (any backend).  ``fence()`` = stream sync + barrier,
    after which rank 0 may read ``tensor()`` (a torch view of the buffer)."""

    def __init__(self, n_complex: int, rank: int, world: int, device_index: int, group=None):
        import ctypes
        import torch
        import torch.distributed as dist
        from . import capi
        self._lib = capi.load()
        self.rank, self.world, self.n_complex, self.group = rank, world, int(n_complex), group
        self._owned = self._mapped = None
        hbytes = [None]
        if rank == 0:
            p = self._lib.r2iq_cuda_device_alloc(int(device_index), self.n_complex * 8)
            if not p:
                raise capi.R2iqError(capi.E_NOMEM, capi.last_error())
            self._owned = p
            buf = (ctypes.c_ubyte * 64)()
            capi.check(self._lib.r2iq_cuda_ipc_export(ctypes.c_void_p(p), buf))
            hbytes = [bytes(buf)]
        if world > 1:
            dist.broadcast_object_list(hbytes, src=0, group=group)
        if rank == 0:
            self.base = self._owned
        else:
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(hbytes[0])
            mapped = ctypes.c_void_p()
            capi.check(self._lib.r2iq_cuda_ipc_open(int(device_index), buf, ctypes.byref(mapped)))
            self._mapped = mapped.value
            self.base = self._mapped
        self._torch = torch
        self._dist = dist

    def ptr(self, offset_complex: int = 0) -> int:
        return self.base + 8 * int(offset_complex)

    def fence(self) -> None:
        self._torch.cuda.synchronize()
        if self.world > 1:
            self._dist.barrier(group=self.group)

    def tensor(self):
        """complex64 view of the whole buffer (rank 0 only)."""
        if self.rank != 0:
            raise RuntimeError("only the collecting rank reads the gathered buffer")
        torch = self._torch

        class _Arr:
            pass
        a = _Arr()
        a.__cuda_array_interface__ = {"shape": (2 * self.n_complex,), "typestr": "<f4", "data": (self.base, False), "version": 3}
        return torch.view_as_complex(torch.as_tensor(a, device=torch.device("cuda", torch.cuda.current_device())).view(-1, 2))

    def close(self) -> None:
        if self._mapped:
            self._lib.r2iq_cuda_ipc_close(self._mapped)
            self._mapped = None
        if self.world > 1:
            self._dist.barrier(group=self.group)      # nobody frees before every mapping is closed
        if self._owned:
            self._lib.r2iq_cuda_device_free(self._owned)
            self._owned = None


def run_time_sharded(process: Callable[[np.ndarray, int], "np.ndarray"], capture: np.ndarray, d: int, rank: int,
                     world: int, history: np.ndarray | None = None):
    """Rank-local part of a time-sharded run: `process(slab_with_halo, n_blocks)` must return the
    complex64 IQ of the slab (the CUDA operator on a GPU; the tests substitute the oracle on CPU)."""
    n_blocks = np.asarray(capture).reshape(-1).shape[0] // BLOCK
    first, count = time_slabs(n_blocks, world, d)[rank]
    if count == 0:
        return np.zeros(0, dtype=np.complex64)
    return process(slab_input(capture, first, count, history), count)
