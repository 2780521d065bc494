"""world_size-2 gloo test (CPU) of the multi-GPU host logic: deterministic utterance sharding agreed without
communication, and the one exchange step of the path — the final gather of variable-length waveforms to rank 0."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

SPC = 1764


def _fake_wave(i, n):
    # utterance i of n codes -> a deterministic "waveform" so rank 0 can verify placement
    return (np.arange(n * SPC, dtype=np.float32) % 977) + 1000.0 * i


def _worker(rank, world, port, lens, q):
    from miocodec_b200 import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    idx = sharding.shard_utterances(lens, world, rank)
    local = np.concatenate([_fake_wave(int(i), int(lens[i])) for i in idx]) if len(idx) else np.empty(0, np.float32)
    waves = sharding.gather_waves(local, idx, lens, SPC)
    ok = True
    if rank == 0:
        for i, n in enumerate(lens):
            ok = ok and np.array_equal(waves[i], _fake_wave(i, int(n)))
    else:
        ok = waves is None
    # the PCM16 form bench.py's c5 block uses: size-exact shards of int16 into one preallocated host buffer on rank 0
    counts = sharding.shard_counts(lens, world, SPC)
    pcm = torch.from_numpy((local % 30000).astype(np.int16))
    host = torch.empty(sum(counts), dtype=torch.int16) if rank == 0 else None
    got = sharding.gather_to_rank0(pcm, counts, host)
    if rank == 0:
        for i, (o, n) in enumerate(sharding.utterance_slices(lens, world, SPC)):
            ok = ok and np.array_equal(got[o: o + n].numpy(), (_fake_wave(i, int(lens[i])) % 30000).astype(np.int16))
    else:
        ok = ok and got is None
    # timing contract of bench.py: max over ranks
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ok = ok and float(t[0]) == float(world)
    q.put((rank, bool(ok), [int(i) for i in idx]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_and_gather():
    import sys
    from conftest import ROOT
    lens = np.array([5, 1, 9, 3, 7, 2, 8], dtype=np.int64)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    os.environ["PYTHONPATH"] = os.pathsep.join([os.path.join(ROOT, "miotts-llama.cpp_b200"), os.environ.get("PYTHONPATH", "")])
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lens, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    got = sorted(i for _, _, idx in res for i in idx)
    assert got == list(range(len(lens)))


def test_collectives_are_ordered_behind_the_codec_stream():
    """sharding.order_after_codec: the PCM16 conversion is only enqueued on the codec's stream, a torch.distributed
    collective is ordered against torch's current stream -- when the two differ (private codec stream, or another stream)
    the host must wait for the codec first; when they are the same stream nothing is needed."""
    from miocodec_b200 import sharding

    class FakeCodec:
        def __init__(self, stream_ptr):
            self.stream_ptr, self.syncs = stream_ptr, 0

        def batch_sync(self):
            self.syncs += 1

    private = FakeCodec(None)                         # never given a stream: runs on its own non-blocking stream
    assert sharding.order_after_codec(private, 0) and private.syncs == 1
    other = FakeCodec(0x7000)
    assert sharding.order_after_codec(other, 0x9000) and other.syncs == 1
    same = FakeCodec(0x7000)
    assert not sharding.order_after_codec(same, 0x7000) and same.syncs == 0
    default = FakeCodec(0)                            # set_stream(0): the legacy default stream, which is torch's default too
    assert not sharding.order_after_codec(default, 0) and default.syncs == 0
