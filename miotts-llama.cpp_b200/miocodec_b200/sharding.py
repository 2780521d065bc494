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


def order_after_codec(codec, consumer_stream_ptr: int) -> bool:
    """Make work enqueued on the CUDA stream `consumer_stream_ptr` (torch's current stream when a collective is about to be
    issued) run after everything the codec has enqueued so far.  mio_batch_pcm16_device / mio_batch_pcm16_to_device only
    ENQUEUE the conversion on the context's stream, and torch.distributed orders a collective against torch's current stream
    alone: if the two differ -- a codec left on its private stream, or set to some other stream -- the collective could read or
    announce a buffer whose stores have not landed.  Same stream: stream order already does it, nothing happens.  Otherwise
    the host waits for the codec (batch_sync).  Returns True when it had to wait."""
    if getattr(codec, "stream_ptr", None) is not None and int(codec.stream_ptr) == int(consumer_stream_ptr or 0):
        return False
    codec.batch_sync()
    return True


def _wire(t):
    """Bytes on the wire: NCCL's torch binding has no int16, and a gather only moves bytes."""
    import torch
    return t.view(torch.uint8) if t.dtype == torch.int16 else t


def gather_to_rank0(local, counts: list[int], host_out=None, group=None):
    """The one exchange step of the sharded path: the variable-length shards of the waveform go to rank 0,
    SIZE-EXACT (every rank knows every count, so nothing is padded to the largest shard and no size exchange is needed).

    local:    this rank's shard: a torch tensor (CUDA for the nccl backend -- e.g. device_tensor(*codec.pcm16_device()),
              which is filled asynchronously on the codec's stream: call order_after_codec(codec,
              torch.cuda.current_stream().cuda_stream) before this function unless the codec runs on torch's stream --,
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


def split_groups(lengths, n_groups: int) -> list[np.ndarray]:
    """Cut a shard (utterance lengths in shard order) into n_groups CONTIGUOUS runs of about equal code count; returns the
    index arrays (possibly empty).  Deterministic: every rank can compute every rank's groups."""
    lengths = np.asarray(lengths, dtype=np.int64)
    n = len(lengths)
    if n == 0:
        return [np.arange(0)] * n_groups
    cum = np.cumsum(lengths)
    cuts = [0]
    for g in range(1, n_groups):
        cuts.append(max(cuts[-1], int(np.searchsorted(cum, cum[-1] * g / n_groups, side="left")) + 0))
    cuts.append(n)
    return [np.arange(cuts[g], max(cuts[g], cuts[g + 1])) for g in range(n_groups)]


def p2p_layout(lengths, world: int, samples_per_code: int, n_groups: int) -> dict:
    """Layout of the job buffer of P2PGather: group-major [group 0: rank 0 | rank 1 | ...][group 1: ...], every
    (rank, group) slice starting on a multiple of 8 samples (the conversion kernel stores 16 bytes at a time).
    groups[r][g] = utterance ids rank r decodes in group g; count / off[r][g] = samples / offset of that slice;
    group_range[g] = (lo, hi) of the whole group (one contiguous host copy); slices[i] = (offset, samples) of utterance i."""
    lengths = np.asarray(lengths, dtype=np.int64)
    shards = [shard_utterances(lengths, world, r) for r in range(world)]
    groups = [[sh[idx] for idx in split_groups(lengths[sh], n_groups)] for sh in shards]
    count = [[int(lengths[ids].sum()) * samples_per_code for ids in gr] for gr in groups]
    off = [[0] * n_groups for _ in range(world)]
    group_range, slices = [], [None] * len(lengths)
    o = 0
    for g in range(n_groups):
        start = o
        for r in range(world):
            o = (o + 7) & ~7
            off[r][g] = o
            oo = o
            for i in groups[r][g]:
                n = int(lengths[i]) * samples_per_code
                slices[int(i)] = (oo, n)
                oo += n
            o += count[r][g]
        group_range.append((start, o))
    return dict(groups=groups, count=count, off=off, group_range=group_range, slices=slices, total=o)


class P2PGather:
    """The gather of the sharded sweep WITHOUT a transfer step: rank 0 owns one device buffer for the whole job and maps it
    into every rank (CUDA IPC over the NVLink / NVSwitch fabric); each rank's PCM16 conversion kernel stores its result
    straight into its slice of that buffer (mio_batch_pcm16_to_device), so the audio crosses NVLink inside the kernel that
    produces it.  The shard is decoded in `n_groups` groups; after group g every rank passes a tiny stream-ordered
    all-reduce (no host synchronisation), then rank 0 copies group g's region -- laid out group-major, so it is one
    contiguous range -- to the pinned host buffer on a side stream while every GPU decodes group g+1.

    Layout of the buffer (device and host): [group 0: rank 0 | rank 1 | ... ][group 1: ...] ..., every slice starting on a
    multiple of 8 samples; `slices` maps every utterance of the job to (offset, samples)."""

    def __init__(self, lengths, samples_per_code: int, n_groups: int, local_rank: int, group=None, pin_host: bool = True):
        import torch
        import torch.distributed as dist
        self.dist, self.torch, self.group = dist, torch, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.local, self.G, self.spc = local_rank, n_groups, samples_per_code
        lay = p2p_layout(lengths, self.world, samples_per_code, n_groups)
        self.groups, self.count, self.off = lay["groups"], lay["count"], lay["off"]
        self.group_range, self.slices, self.total = lay["group_range"], lay["slices"], lay["total"]
        from . import binding
        self.binding = binding
        dev = torch.device("cuda", local_rank)
        box = [None]
        self.owner = self.rank == 0
        if self.owner:
            self.ptr = binding.device_alloc(local_rank, max(2, self.total * 2))
            box[0] = binding.ipc_export(self.ptr)              # 64-byte CUDA IPC handle
        dist.broadcast_object_list(box, src=0, group=group)
        if not self.owner:
            self.ptr = binding.ipc_open(box[0], local_rank)    # rank 0's memory mapped here, peer access enabled with it
        self.flag = torch.zeros(1, dtype=torch.float32, device=dev)
        self.host = None
        if self.owner:
            self.buf = device_tensor(self.ptr, max(1, self.total), "int16", local_rank)
            self.host = torch.empty(max(1, self.total), dtype=torch.int16)
            if pin_host:
                self.host = self.host.pin_memory()
            self.side = torch.cuda.Stream(device=dev)

    def close(self):
        self.torch.cuda.synchronize()
        self.dist.barrier(group=self.group)                    # nobody unmaps while a peer may still be storing ...
        if not self.owner:
            self.binding.ipc_close(self.ptr)
        self.dist.barrier(group=self.group)                    # ... and the owner frees only after every mapping is closed
        if self.owner:
            self.buf = None
            self.binding.device_free(self.ptr)
        self.ptr = 0

    def my_group(self, g: int) -> np.ndarray:
        """Utterance ids (of the whole job) this rank decodes in group g."""
        return self.groups[self.rank][g]

    def put(self, codec, g: int):
        """After codec.batch_run() of group g: store its PCM16 into the job buffer, then the stream-ordered rendezvous;
        rank 0 starts the host copy of the whole group's region behind it."""
        n = self.count[self.rank][g]
        if n > 0:
            codec.pcm16_to_device(self.ptr + 2 * self.off[self.rank][g], n)
            # the stores above are only enqueued, on the CODEC's stream; the rendezvous below is ordered against torch's
            # current stream: free when the two are one (codec.set_stream(torch.cuda.current_stream().cuda_stream)), a host
            # wait otherwise
            order_after_codec(codec, self.torch.cuda.current_stream().cuda_stream)
        self.dist.all_reduce(self.flag, group=self.group)      # every rank's stores of group g precede this on its stream
        if self.rank == 0:
            lo, hi = self.group_range[g]
            if hi > lo:
                ev = self.torch.cuda.Event()
                ev.record(self.torch.cuda.current_stream())
                with self.torch.cuda.stream(self.side):
                    self.side.wait_event(ev)
                    self.host[lo:hi].copy_(self.buf[lo:hi], non_blocking=True)

    def finish(self):
        """Rank 0: wait for the last host copy; returns the host buffer (None elsewhere)."""
        if self.rank == 0:
            self.side.synchronize()
        return self.host


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
