"""Utterance sharding across the GPUs of one box: utterances are independent, so the path shards
with no data-path collective.  Longest-processing-time-first bin packing by code count (cost is linear in the
number of codes with banded attention); every rank computes the same deterministic assignment from the lengths
alone, so no communication is needed to agree on it."""
from __future__ import annotations

import heapq
import os

import numpy as np


def shard_utterances(lengths, world_size: int, rank: int) -> np.ndarray:
    """Indices (ascending) of the utterances rank `rank` decodes."""
    lengths = np.asarray(lengths, dtype=np.int64)
    if world_size <= 1:
        return np.arange(len(lengths))
    order = np.argsort(-lengths, kind="stable")
    heap = [(0, r) for r in range(world_size)]
    heapq.heapify(heap)
    mine = []
    for i in order:
        load, r = heapq.heappop(heap)
        if r == rank:
            mine.append(int(i))
        heapq.heappush(heap, (load + int(lengths[i]), r))
    return np.asarray(sorted(mine), dtype=np.int64)


def shard_counts(lengths, world_size: int, samples_per_code: int) -> list[int]:
    """Samples every rank produces (each rank can compute everybody's: the assignment is deterministic)."""
    lengths = np.asarray(lengths, dtype=np.int64)
    return [int(lengths[shard_utterances(lengths, world_size, r)].sum()) * samples_per_code for r in range(world_size)]


def utterance_slices(lengths, world_size: int, samples_per_code: int) -> list[tuple[int, int]]:
    """(offset, samples) of every utterance in the rank-major gathered buffer [rank 0's shard | rank 1's shard | ...]."""
    lengths = np.asarray(lengths, dtype=np.int64)
    out = [None] * len(lengths)
    off = 0
    for r in range(world_size):
        for i in shard_utterances(lengths, world_size, r):
            n = int(lengths[i]) * samples_per_code
            out[int(i)] = (off, n)
            off += n
    return out


class _DevicePointer:
    """A raw device allocation (e.g. mio_batch_wave_device / mio_batch_pcm16_device) as a __cuda_array_interface__ object,
    so torch.as_tensor wraps it WITHOUT a copy."""

    def __init__(self, ptr: int, n: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def device_tensor(ptr: int, n: int, dtype: str = "float32", device: int = 0):
    """Zero-copy torch view of n elements at device pointer ptr (dtype 'float32' or 'int16')."""
    import torch
    typestr = {"float32": "<f4", "int16": "<i2"}[dtype]
    return torch.as_tensor(_DevicePointer(ptr, n, typestr), device=torch.device("cuda", device))


def _wire(t):
    """Bytes on the wire: NCCL's torch binding has no int16, and a gather only moves bytes."""
    import torch
    return t.view(torch.uint8) if t.dtype == torch.int16 else t


def gather_to_rank0(local, counts: list[int], host_out=None, group=None):
    """The one exchange step of the sharded path: the variable-length shards of the waveform go to rank 0,
    SIZE-EXACT (every rank knows every count, so nothing is padded to the largest shard and no size exchange is needed).

    local:    this rank's shard: a torch tensor (CUDA for the nccl backend -- e.g. device_tensor(*codec.pcm16_device()) --,
              CPU for gloo) holding counts[rank] elements.
    host_out: rank 0 only, optional: a host tensor (pinned for full-speed copies) of sum(counts) elements; shard r is
              copied device->host on a second stream while shard r+1 is still arriving over NVLink.
    Returns on rank 0 the rank-major gathered tensor (host_out when given, else a tensor on `local`'s device), None elsewhere.
    """
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    assert local.numel() == counts[rank], (local.numel(), counts[rank])
    if rank != 0:
        if counts[rank] > 0:
            dist.send(_wire(local), dst=0, group=group)
        return None
    total = sum(counts)
    on_gpu = local.is_cuda
    offs = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    if host_out is None or not on_gpu:
        out = host_out if host_out is not None else torch.empty(total, dtype=local.dtype, device=local.device)
        out[: counts[0]].copy_(local)
        for r in range(1, world):
            if counts[r] > 0:
                dist.recv(_wire(out[int(offs[r]): int(offs[r + 1])]), src=r, group=group)
        return out
    # device shards -> one host buffer: receive into per-rank staging, copy out on a side stream as each one lands
    assert host_out.numel() >= total
    side = torch.cuda.Stream(device=local.device)
    main = torch.cuda.current_stream(local.device)
    ready = torch.cuda.Event()
    ready.record(main)
    with torch.cuda.stream(side):
        side.wait_event(ready)
        host_out[: counts[0]].copy_(local, non_blocking=True)
    stage = [None] * world
    for r in range(1, world):
        if counts[r] == 0:
            continue
        stage[r] = torch.empty(counts[r], dtype=local.dtype, device=local.device)
        dist.recv(_wire(stage[r]), src=r, group=group)
        ev = torch.cuda.Event()
        ev.record(main)
        with torch.cuda.stream(side):
            side.wait_event(ev)
            host_out[int(offs[r]): int(offs[r + 1])].copy_(stage[r], non_blocking=True)
    side.synchronize()
    return host_out


def gather_waves(local_wave: np.ndarray, local_idx: np.ndarray, lengths, samples_per_code: int, group=None):
    """Final gather of variable-length waveforms to rank 0 from HOST arrays (tests, small jobs).
    Returns the list of per-utterance waveforms on rank 0, None elsewhere.  Size-exact sends (gather_to_rank0); the
    device-resident form used by bench.py is gather_to_rank0 on device_tensor(...)."""
    import torch
    import torch.distributed as dist
    lengths = np.asarray(lengths, dtype=np.int64)
    world = dist.get_world_size(group)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    counts = shard_counts(lengths, world, samples_per_code)
    local = torch.from_numpy(np.ascontiguousarray(local_wave, dtype=np.float32)).to(dev)
    out = gather_to_rank0(local, counts, None, group)
    if out is None:
        return None
    flat = out.cpu().numpy()
    return [flat[o: o + n] for o, n in utterance_slices(lengths, world, samples_per_code)]


class SharedPinnedBuffer:
    """ONE host buffer for the whole job, written by every rank directly: a /dev/shm file mapped by all ranks, each rank
    page-locking (cudaHostRegister) only the slice its own GPU copies into.  The alternative to gather_to_rank0 when the
    consumer is one process per box: N PCIe links work in parallel instead of everything funnelling through rank 0's."""

    def __init__(self, name: str, total_elems: int, dtype, my_off: int, my_n: int, create: bool):
        import torch
        self.path = os.path.join("/dev/shm", name)
        self.dtype = np.dtype(dtype)
        nbytes = max(1, int(total_elems)) * self.dtype.itemsize
        if create:
            with open(self.path, "wb") as f:
                f.truncate(nbytes)
        self.map = np.memmap(self.path, dtype=self.dtype, mode="r+", shape=(max(1, int(total_elems)),))
        self.mine = self.map[my_off: my_off + my_n]
        self._reg = None
        if my_n > 0:
            ptr, size = self.mine.ctypes.data, my_n * self.dtype.itemsize
            lo = ptr & ~4095                                    # register whole pages
            hi = (ptr + size + 4095) & ~4095
            self.mine[:: max(1, 4096 // self.dtype.itemsize)] = 0   # touch: back the pages before pinning them
            rc = torch.cuda.cudart().cudaHostRegister(lo, hi - lo, 0)
            if int(rc) != 0:
                raise RuntimeError(f"cudaHostRegister failed ({rc})")
            self._reg = lo

    def close(self, unlink: bool):
        import torch
        if self._reg is not None:
            torch.cuda.cudart().cudaHostUnregister(self._reg)
            self._reg = None
        self.mine = None
        del self.map
        if unlink:
            try:
                os.unlink(self.path)
            except OSError:
                pass
