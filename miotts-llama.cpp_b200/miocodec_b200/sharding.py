"""Utterance sharding across the GPUs of one box: utterances are independent, so the path shards
with no data-path collective.  Longest-processing-time-first bin packing by code count (cost is linear in the
number of codes with banded attention); every rank computes the same deterministic assignment from the lengths
alone, so no communication is needed to agree on it."""
from __future__ import annotations

import heapq
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


def gather_waves(local_wave: np.ndarray, local_idx: np.ndarray, lengths, samples_per_code: int, group=None):
    """Final gather of variable-length waveforms to rank 0 (the one exchange step of the path).
    Returns the list of per-utterance waveforms on rank 0, None elsewhere.  Uses torch.distributed
    (NCCL on the GPU box, gloo in the CPU tests)."""
    import torch
    import torch.distributed as dist
    lengths = np.asarray(lengths, dtype=np.int64)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    counts = [int(lengths[shard_utterances(lengths, world, r)].sum()) * samples_per_code for r in range(world)]
    mx = max(counts) if counts else 0
    buf = torch.zeros(mx, dtype=torch.float32, device=dev)
    buf[: local_wave.size] = torch.from_numpy(np.ascontiguousarray(local_wave)).to(dev)
    out = [torch.empty(mx, dtype=torch.float32, device=dev) for _ in range(world)] if rank == 0 else None
    dist.gather(buf, out, dst=0, group=group)
    if rank != 0:
        return None
    waves = [None] * len(lengths)
    for r in range(world):
        idx = shard_utterances(lengths, world, r)
        flat = out[r][: counts[r]].cpu().numpy()
        off = 0
        for i in idx:
            n = int(lengths[i]) * samples_per_code
            waves[int(i)] = flat[off: off + n]
            off += n
    return waves
