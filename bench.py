#!/usr/bin/env python
"""Benchmark of the codec+ISTFT hot path (BASELINE.json metric: audio-seconds per second, RTF^-1).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c2|c5] [--impl reference]

A step is one pass of the hot path over one synthetic batch per GPU.  Default workload = BASELINE.json configs[2]
("batch of 256 synthetic ~10 s utterances, mixed voice embeddings, offline codec+ISTFT on 1 B200"), the
single-GPU throughput configuration the metric is quoted on; with N GPUs every rank decodes its own batch of the
same shape (utterance sharding, weak scaling, no data-path collective).  `--workload c2` is the single 700-code
utterance (latency), `--workload c5` the 4096-utterance 2-30 s sweep sharded by utterance.

One JSON line on stdout (rank 0).  `value` is timed on the device (CUDA events on the launching stream) with the
inputs resident in HBM; `e2e` goes through the host-buffer C-ABI call with the H2D/D2H copies in the timed region.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))

METRIC = "codec+ISTFT audio-sec/sec (RTF^-1)"
UNIT = "audio_s/s"
N_VOICES = 16
# algorithmic work per code: dense contractions + banded attention, upsampler kernels [9,9]
FLOP_PER_CODE = 250.45e6


def make_workload(name: str, rank: int, world: int):
    from miocodec_b200 import sharding, synth
    if name == "c2":
        lens = np.array([700], dtype=np.int64)
        desc = "C2: single utterance, 700-code synthetic sequence, batch 1"
        seed = 1234
    elif name == "c5":
        rng = np.random.default_rng(5)
        all_lens = rng.integers(50, 751, size=4096)
        idx = sharding.shard_utterances(all_lens, world, rank)
        lens = all_lens[idx]
        desc = "C5: 4096 synthetic utterances of 50-750 codes (2-30 s), sharded by utterance"
        seed = 5000 + rank
    else:
        lens = np.full(256, 250, dtype=np.int64)
        desc = "C3: batch of 256 synthetic 10 s utterances (250 codes each), 16 mixed voice embeddings, per GPU"
        seed = 3 + 1000 * rank
    cu = np.zeros(len(lens) + 1, dtype=np.int32)
    cu[1:] = np.cumsum(lens)
    codes = synth.synth_codes(int(cu[-1]), seed)
    vt = np.stack([synth.synth_voice(v) for v in range(N_VOICES)]).astype(np.float32)
    vids = (np.arange(len(lens)) % N_VOICES).astype(np.int32)
    return dict(desc=desc, lens=lens, cu=cu, codes=codes, voice_table=vt, voice_ids=vids)


def bench_config(wl, world: int):
    """The `config` object of the JSON line: identical in both arms (--impl b200 / --impl reference) by construction."""
    total_codes = int(wl["cu"][-1])
    return {"workload": wl["desc"], "codes_per_gpu_per_step": total_codes,
            "audio_s_per_gpu_per_step": total_codes * 1764 / 44100.0,
            "weights": "random-init miocodec.gguf layout, seed 0", "parallelism": f"utterance-sharded x{world}",
            "l2": "per-step working set (GBs of activations) far exceeds the 126 MB L2; no explicit flush"}


def mem_available_bytes() -> int:
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) * 1024
    except OSError:
        pass
    return 0


class stdout_to_stderr:
    """Route C-level stdout (fd 1) to stderr while NCCL initialises: its version banner goes to stdout and the
    contract is ONE JSON line there."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)

    def __exit__(self, *a):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return dict(sm_mhz=statistics.median(sm) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf=d["bf16_tflops"], tf_sus=d.get("bf16_tflops_sustained", d["bf16_tflops"]), src="measured")
    return dict(hbm=6650.0, tf=1590.0, tf_sus=1400.0, src="fallback")


# ------------------------------------------------------------------------------------------------ reference arm
def _ref_worker(args):
    gguf, jobs = args
    from oracle import ref
    R = ref.Reference(gguf)
    t0 = time.perf_counter()
    codes_total = 0
    for codes, emb in jobs:
        w = R.decode_wave(codes, emb)
        assert w.shape[0] == 1764 * len(codes)
        codes_total += len(codes)
    dt = time.perf_counter() - t0
    R.close()
    return codes_total, dt


def cpu_reference_run(wl, n_utts: int, procs: int):
    """Times the UNMODIFIED reference (oracle/_ref) on `n_utts` utterances of the workload, `procs` independent
    contexts in parallel (each uses the reference's fixed 4 ggml threads + 1-thread ISTFT)."""
    import multiprocessing as mp
    from miocodec_b200 import synth
    gguf, _ = synth.ensure_assets()
    cu, codes = wl["cu"], wl["codes"]
    n_utts = min(n_utts, len(wl["lens"]))
    jobs = [(codes[cu[i]:cu[i + 1]].copy(), wl["voice_table"][wl["voice_ids"][i]].copy()) for i in range(n_utts)]
    procs = max(1, min(procs, n_utts))
    parts = [(gguf, jobs[p::procs]) for p in range(procs)]
    t0 = time.perf_counter()
    if procs == 1:
        res = [_ref_worker(parts[0])]
    else:
        with mp.get_context("spawn").Pool(procs) as pool:
            res = pool.map(_ref_worker, parts)
    wall = time.perf_counter() - t0
    total_codes = sum(r[0] for r in res)
    busy = max(r[1] for r in res)            # excludes model load / process spawn, like the reference's own timers
    audio_s = total_codes * 1764 / 44100.0
    return audio_s / busy, audio_s, busy, wall


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref
    wl = make_workload(args.workload, 0, 1)
    ncpu = os.cpu_count() or 4
    procs = max(1, ncpu // 4)
    mean_len = float(np.mean(wl["lens"]))
    # bounded sample: about 3 s of reference work per context per step (the reference runs ~8x real time at T>=250)
    per_proc = 1 if mean_len >= 200 else 2
    n_utts = min(len(wl["lens"]), procs * per_proc)
    if not ref.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built and this host cannot load it"}))
        return
    vals = []
    for i in range(args.warmup + args.steps):
        v, audio_s, busy, wall = cpu_reference_run(wl, n_utts, procs)
        if i >= args.warmup:
            vals.append((v, busy))
    value = float(np.mean([v for v, _ in vals]))
    ms = float(np.mean([b for _, b in vals])) * 1e3
    sample = f"{n_utts} utterances of the workload ({audio_s:.0f} audio-s) per step, {procs} independent reference contexts x 4 ggml threads"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": bench_config(wl, args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs * 4, "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ streaming (C4)
def streaming_session(codec, S: int, N: int):
    """S concurrent streams of N codes in lock-step under the reference's streaming policy (full-prefix re-decode at the
    commit cadence, src/test-to-speech.cpp:499-571).  Codes are pushed 20 at a time (the cadence is evaluated per code inside
    the library, so the decode schedule is the one-code-per-tick schedule of the reference).  Each round is split into
    mio_streams_step_begin / _end with the consumer's reads in between (host copies under the GPU's decode).  Returns
    (emitted samples, wall seconds, StreamProfile counters of stream 0)."""
    from miocodec_b200 import synth
    streams = [codec.open_stream(synth.synth_voice(i % N_VOICES)) for i in range(S)]
    codes = [synth.synth_codes(N, 4000 + i) for i in range(S)]
    t0 = time.perf_counter()
    emitted = 0
    def drain():
        n = 0
        for st in streams:
            while True:                       # 256k-sample reads: a 16 MB buffer per read costs more than the copy
                a, _ = st.read(1 << 18)
                n += a.size
                if a.size < (1 << 18):
                    break
        return n

    for k in range(0, N, 20):
        for i, st in enumerate(streams):
            st.push(codes[i][k:k + 20])
            if k + 20 >= N:
                st.finish()
        # a throughput-oriented consumer: when a round is in flight it reads what EARLIER rounds emitted, i.e. its host
        # copies run under the GPU's decode (audio reaches it one round, ~0.3 s at this load, later than it could)
        if codec.step_streams_begin(streams):
            emitted += drain()
            codec.step_streams_end(streams)
    emitted += drain()
    dt = time.perf_counter() - t0
    stats = streams[0].stats()
    for st in streams:
        st.close()
    return emitted, dt, stats


def cpu_streaming_baseline(N: int):
    """The reference's OWN streaming session (oracle/_ref: unmodified src/test-to-speech.cpp LLM branch, scripted stub
    LLM, the reference's CPU codec on its fixed 4 ggml threads) for ONE stream of N codes: emitted audio-s per second."""
    from miocodec_b200 import synth
    from oracle import ref
    if not ref.available():
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref unavailable on this host"}
    gguf, voices = synth.ensure_assets()
    codes = synth.synth_codes(N, 4000)
    t0 = time.perf_counter()
    out, chunks, st = ref.ref_stream_session(gguf, voices[0], codes)
    dt = time.perf_counter() - t0
    return {"value": (len(out) / 44100.0) / dt, "unit": UNIT, "cores": 4, "kind": "reference", "seconds": dt,
            "decode_calls": st["decode_calls"], "decoded_codes": st["decoded_codes"],
            "sample": f"1 stream of {N} codes through TestToSpeech::synthesize_stream_profiled (stub LLM, model load included), 4 ggml threads"}


def c4_block(codec, S: int = 1024, N: int = 700, sessions: int = 1, cpu: bool = True):
    # warm-up = one FULL session: a first session on a fresh context spends ~2 of its ~7 s growing device and pinned host
    # buffers (cudaMalloc / cudaMallocHost of hundreds of MB to GBs, MIO_STREAM_TIMING=1 shows them); a serving process
    # pays that once
    _, cold_dt, _ = streaming_session(codec, S, N)
    runs = [streaming_session(codec, S, N) for _ in range(sessions)]
    emitted, _, st = runs[-1]
    dt = float(np.mean([r[1] for r in runs]))
    audio_s = emitted / codec.sample_rate
    redecode = st["decoded_codes"] / N
    peaks = measured_peaks()
    ach = FLOP_PER_CODE * st["decoded_codes"] * S / dt / 1e12
    blk = {"metric": "streaming codec+ISTFT emitted audio-sec/sec", "value": audio_s / dt, "unit": UNIT, "seconds_per_session": dt,
           "sessions": sessions, "first_session_seconds": cold_dt, "workload": f"C4: {S} concurrent streams x {N} codes, lock-step, reference streaming policy",
           "decode_calls_per_stream": st["decode_calls"], "decoded_codes_per_stream": st["decoded_codes"],
           "redecode_factor": redecode, "decoded_audio_s_per_s": audio_s * redecode / dt,
           "timing": "host wall clock around the whole session: pushes, rounds (mio_streams_step_begin / _end), the consumer's reads of every emitted sample into host arrays (done while the next round is in flight)",
           "roofline": {"bound": "tensor", "unit": "TFLOP/s", "peak": peaks["tf_sus"], "achieved": ach, "frac": ach / peaks["tf_sus"],
                        "note": "algorithmic FLOPs of the codes the policy re-decodes / session wall time"}}
    if cpu:
        try:
            blk["cpu_baseline"] = cpu_streaming_baseline(N)
        except Exception as e:  # noqa: BLE001
            blk["cpu_baseline"] = {"value": None, "sample": f"failed: {e}"}
    return blk


def run_streaming(args):
    """--workload c4: BASELINE config C4 alone (1024 concurrent streams per GPU, 700 codes each).  Metric: EMITTED
    audio-seconds per wall-second (the policy re-decodes ~10x the emitted audio)."""
    import torch
    from miocodec_b200 import binding, synth
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    gguf, _ = synth.ensure_assets()
    codec = binding.MioCodec(gguf, local)
    blk = c4_block(codec, sessions=max(1, min(args.steps, 3)), cpu=not args.no_cpu_baseline)
    line = {"metric": blk["metric"], "value": blk["value"], "unit": UNIT, "n_gpus": 1, "steps": blk["sessions"], "warmup": 1,
            "ms_per_step": blk["seconds_per_session"] * 1e3, "higher_is_better": True, "data": "synthetic",
            "config": {"workload": blk["workload"]}, **{k: v for k, v in blk.items() if k not in ("metric", "value", "unit", "workload")}}
    print(json.dumps(line))
    codec.close()


# ------------------------------------------------------------------------------------------------ latency (C2)
def c2_block(codec, stream, cpu: bool = True):
    """BASELINE config C2: ONE 700-code utterance, batch 1.  Device time per decode (CUDA events, resident input) and the
    host-to-host latency of the synchronous single-utterance call."""
    import torch
    wl = make_workload("c2", 0, 1)
    emb = wl["voice_table"][int(wl["voice_ids"][0])]
    codec.batch_upload(wl["codes"], wl["cu"], wl["voice_table"], wl["voice_ids"])
    for _ in range(5):
        codec.batch_run(False)
    codec.batch_sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    runs = 50
    e0.record(stream)
    for _ in range(runs):
        codec.batch_run(False)
    e1.record(stream)
    codec.batch_sync()
    dev_ms = e0.elapsed_time(e1) / runs
    launches = codec.kernel_launches()
    lat = []
    for i in range(103):
        t1 = time.perf_counter()
        codec.decode_wave(wl["codes"], emb)
        if i >= 3:
            lat.append((time.perf_counter() - t1) * 1e3)
    lat.sort()
    # the same call into PINNED host memory (mio_host_alloc): a pageable destination costs ~0.3 ms of driver staging for 4.9 MB
    from miocodec_b200 import binding
    h_out = binding.pinned_empty(700 * codec.samples_per_token, np.float32)
    h_in = binding.pinned_empty(700, np.int32); h_in[:] = wl["codes"]
    latp = []
    for i in range(103):
        t1 = time.perf_counter()
        codec.decode_wave_batch(h_in, wl["cu"], wl["voice_table"][:1], np.zeros(1, np.int32), out=h_out)
        if i >= 3:
            latp.append((time.perf_counter() - t1) * 1e3)
    latp.sort()
    audio_s = 700 * codec.samples_per_token / codec.sample_rate
    blk = {"workload": wl["desc"], "device_ms": dev_ms, "launches": launches, "value": audio_s / (dev_ms / 1e3), "unit": UNIT,
           "latency": {"p50_ms": lat[len(lat) // 2], "p99_ms": lat[min(len(lat) - 1, int(len(lat) * 0.99))], "min_ms": lat[0],
                       "runs": len(lat), "call": "mio_decode_wave (pageable host buffers, synchronous)"},
           "latency_pinned": {"p50_ms": latp[len(latp) // 2], "p99_ms": latp[min(len(latp) - 1, int(len(latp) * 0.99))], "min_ms": latp[0],
                              "runs": len(latp), "call": "mio_decode_wave_batch, one utterance, pinned host buffers (mio_host_alloc), synchronous"},
           "e2e": {"value": audio_s / (lat[len(lat) // 2] / 1e3), "unit": UNIT, "h2d_bytes_per_step": 700 * 4 + 128 * 4,
                   "d2h_bytes_per_step": 700 * codec.samples_per_token * 4}}
    if cpu:
        try:
            from oracle import ref
            if ref.available():
                v, a_s, busy, wall = cpu_reference_run(wl, 1, 1)
                blk["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": 4, "kind": "reference", "ms": busy * 1e3,
                                       "sample": "the same utterance, once, 1 reference context = 4 ggml threads + 1-thread ISTFT"}
        except Exception as e:  # noqa: BLE001
            blk["cpu_baseline"] = {"value": None, "sample": f"failed: {e}"}
    return blk


# ------------------------------------------------------------------------------------------------ strong scaling (C5)
def c5_block(codec, stream, rank: int, world: int, local: int, barrier, sweeps: int = 2):
    """BASELINE config C5 as stated: ONE sweep of 4096 utterances (50-750 codes, 2-30 s), sharded by utterance over the
    ranks (strong scaling).  Device-resident time, and three end-to-end deliveries of PCM16 audio to the host:
      per_rank_host   every rank copies its shard into its own pinned buffer (N consumers)
      gather_rank0    the path's one exchange step: shards go to rank 0 over NCCL/NVLink size-exact, rank 0 copies them to
                      ONE pinned buffer (copy of shard r overlapped with the arrival of shard r+1)
      shared_pinned   every rank copies its shard straight into its slice of ONE /dev/shm buffer page-locked by all ranks
    All times: max over ranks of the host wall clock between barriers (device idle on both sides)."""
    import torch
    import torch.distributed as dist
    from miocodec_b200 import binding, sharding
    wl = make_workload("c5", rank, world)
    rng = np.random.default_rng(5)
    all_lens = rng.integers(50, 751, size=4096)
    spc = codec.samples_per_token
    counts = sharding.shard_counts(all_lens, world, spc)
    total = int(sum(counts))
    mine = counts[rank]
    audio_all = total / codec.sample_rate
    blk = {"workload": wl["desc"], "utterances": 4096, "total_codes": int(all_lens.sum()), "audio_s": audio_all, "scaling": "strong",
           "sweeps": sweeps, "delivery": "PCM16 (2 bytes per sample; the float32 form is the headline block's e2e)"}

    def tmax(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    # ---- device resident
    codec.batch_upload(wl["codes"], wl["cu"], wl["voice_table"], wl["voice_ids"])
    codec.batch_run(False)
    codec.batch_sync()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(sweeps):
        codec.batch_run(False)
    e1.record(stream)
    barrier()
    ms = tmax(e0.elapsed_time(e1)) / sweeps
    blk.update(ms_per_sweep=ms, value=audio_all / (ms / 1e3), unit=UNIT, gpu_launches_per_sweep=codec.kernel_launches())

    need = (total + (total if world > 1 else 0)) * 2           # all ranks' pinned shards + rank 0's gathered copy, one box
    short = 1.0 if mem_available_bytes() < 3 * need + (8 << 30) else 0.0
    if tmax(short) > 0.0:                                       # a collective decision: every rank skips or none does
        blk["e2e"] = {"skipped": f"host has {mem_available_bytes() >> 30} GiB available, the pinned delivery buffers need {need >> 30} GiB"}
        return blk
    h_codes = binding.pinned_empty(len(wl["codes"]), np.int32); h_codes[:] = wl["codes"]
    h_pcm = binding.pinned_empty(mine, np.int16)

    def timed(fn):
        fn()                                   # warm-up (first touch of the buffers, NCCL channels)
        barrier()
        t0 = time.perf_counter()
        for _ in range(sweeps):
            fn()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        barrier()
        return tmax(dt) / sweeps

    def per_rank():
        codec.decode_pcm16_batch(h_codes, wl["cu"], wl["voice_table"], wl["voice_ids"], out=h_pcm)
    s1 = timed(per_rank)
    e2e = {"per_rank_host": {"value": audio_all / s1, "unit": UNIT, "ms_per_sweep": s1 * 1e3,
                             "h2d_bytes": int(len(wl["codes"]) * 4), "d2h_bytes": int(mine * 2)}}
    if world > 1:
        # every "can this rank go on" decision below is taken collectively (max over ranks of a failure flag): a rank that
        # dropped out alone would leave the others waiting in a send / barrier
        host_all, fail = None, 0.0
        if rank == 0:
            try:
                host_all = torch.empty(total, dtype=torch.int16).pin_memory()
            except Exception:  # noqa: BLE001
                fail = 1.0
        if tmax(fail) > 0.0:
            e2e["gather_rank0_nccl"] = {"skipped": "rank 0 could not pin the gathered buffer"}
        else:
            def gathered():
                codec.batch_upload(h_codes, wl["cu"], wl["voice_table"], wl["voice_ids"])
                codec.batch_run(False)
                ptr, n = codec.pcm16_device()
                sharding.gather_to_rank0(sharding.device_tensor(ptr, n, "int16", local), counts, host_all)
            s2 = timed(gathered)
            e2e["gather_rank0_nccl"] = {"value": audio_all / s2, "unit": UNIT, "ms_per_sweep": s2 * 1e3,
                                        "nvlink_bytes_to_rank0": int((total - counts[0]) * 2), "d2h_bytes_rank0": int(total * 2)}
            if rank == 0:      # placement check: rank 0's own shard lands at offset 0 of the rank-major buffer
                assert np.array_equal(host_all[: counts[0]].numpy()[:4096], h_pcm[:4096])
        del host_all
        # the same gather with no transfer step: conversion kernels store straight into rank 0's buffer over NVLink, the
        # shard is decoded in groups and rank 0's host copy of group g runs under everybody's decode of group g+1
        G = 8
        pg, fail = None, 0.0
        try:
            pg = sharding.P2PGather(all_lens, spc, G, local)
        except Exception as e:  # noqa: BLE001
            fail, why_p2p = 1.0, f"{type(e).__name__}: {e}"
        if tmax(fail) > 0.0:
            e2e["gather_rank0_p2p"] = {"skipped": why_p2p if fail else "another rank could not map the job buffer"}
            if pg is not None:
                pg.close()
        else:
            local_groups = sharding.split_groups(wl["lens"], G)
            cu = wl["cu"]

            def p2p():
                for g, ids in enumerate(local_groups):
                    if len(ids):
                        a, b = int(ids[0]), int(ids[-1]) + 1
                        codec.batch_upload(h_codes[cu[a]: cu[b]], (cu[a: b + 1] - cu[a]).astype(np.int32), wl["voice_table"], wl["voice_ids"][a:b])
                        codec.batch_run(False)
                    pg.put(codec, g)
                pg.finish()
            s4 = timed(p2p)
            e2e["gather_rank0_p2p"] = {"value": audio_all / s4, "unit": UNIT, "ms_per_sweep": s4 * 1e3, "groups": G,
                                       "nvlink_bytes_to_rank0": int((total - counts[0]) * 2), "d2h_bytes_rank0": int(total * 2),
                                       "how": "pcm16 kernels store into rank 0's IPC-mapped buffer; D2H of group g under the decode of group g+1"}
            # every rank checks its own slices in the job buffer; rank 0 checks everybody's bytes on the host by checksum
            goff, ok = 0, True
            for g, ids in enumerate(local_groups):
                n = pg.count[rank][g]
                if n:
                    got = sharding.device_tensor(pg.ptr + 2 * pg.off[rank][g], n, "int16", local).cpu().numpy()
                    ok = ok and np.array_equal(got, h_pcm[goff: goff + n])
                goff += n
            sums = torch.zeros(world, dtype=torch.float64, device="cuda")
            sums[rank] = float(np.asarray(h_pcm, dtype=np.int64).sum())
            dist.all_reduce(sums)
            if rank == 0:
                host = pg.host.numpy()
                for r in range(world):
                    tot = sum(int(host[pg.off[r][g]: pg.off[r][g] + pg.count[r][g]].astype(np.int64).sum()) for g in range(G))
                    ok = ok and float(tot) == float(sums[r])
            pg.close()
            assert ok, "P2P gather placed bytes wrongly"
        name = f"mio_c5_{os.environ.get('MASTER_PORT', '0')}"
        off = int(sum(counts[:rank]))
        shared, fail, why = None, 0.0, ""
        vfs = os.statvfs("/dev/shm")                         # a tmpfs that is too small would SIGBUS on first touch
        if vfs.f_bavail * vfs.f_frsize < total * 2 + (1 << 30):
            fail, why = 1.0, f"/dev/shm has {vfs.f_bavail * vfs.f_frsize >> 20} MiB free, the shared buffer needs {total * 2 >> 20} MiB"
        if tmax(fail) == 0.0:
            if rank == 0:
                try:
                    shared = sharding.SharedPinnedBuffer(name, total, np.int16, off, mine, create=True)
                except Exception as e:  # noqa: BLE001
                    fail, why = 1.0, f"{type(e).__name__}: {e}"
            barrier()
            if rank != 0 and fail == 0.0:
                try:
                    shared = sharding.SharedPinnedBuffer(name, total, np.int16, off, mine, create=False)
                except Exception as e:  # noqa: BLE001
                    fail, why = 1.0, f"{type(e).__name__}: {e}"
        if tmax(fail) > 0.0:
            e2e["shared_pinned"] = {"skipped": why or "another rank could not map / page-lock the shared buffer"}
        else:
            def direct():
                codec.decode_pcm16_batch(h_codes, wl["cu"], wl["voice_table"], wl["voice_ids"], out=shared.mine)
            s3 = timed(direct)
            e2e["shared_pinned"] = {"value": audio_all / s3, "unit": UNIT, "ms_per_sweep": s3 * 1e3, "d2h_bytes_per_rank": int(mine * 2)}
            assert np.array_equal(np.asarray(shared.mine[:4096]), h_pcm[:4096])
        barrier()
        if shared is not None:
            shared.close(unlink=(rank == 0))
        elif rank == 0:
            try:
                os.unlink(os.path.join("/dev/shm", name))
            except OSError:
                pass
    blk["e2e"] = e2e
    return blk


def bind_to_gpu_numa_node(local: int):
    """One process per GPU: pin this rank (and the pinned host buffers it is about to allocate) to the CPUs NVML reports
    as local to its GPU, so that with several ranks per box every D2H copy stays on its own socket.  Best effort:
    returns the CPU count bound to, or None when NVML / the cpuset does not allow it."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        p = torch.cuda.get_device_properties(local)
        bus = f"{p.pci_domain_id:08x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if len(cpus) < 8:                       # never squeeze the 4-thread CPU baseline or the copy threads
            return None
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------ our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c3", choices=["c2", "c3", "c4", "c5"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-latency", action="store_true",
                    help="skip C2's 100-call latency loop (use under ncu: the loop adds ~16k profiled launches)")
    ap.add_argument("--no-extra", action="store_true",
                    help="headline measurement only: skip the c5 (strong scaling + gather), c2 (latency) and c4 (streaming) blocks")
    ap.add_argument("--gemm", default=None)
    ap.add_argument("--chunk", type=int, default=None, help="max_chunk_codes of the context (workspace rows)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else max(args.warmup, 0)
    if args.impl == "reference":
        run_reference(args)
        return
    if args.workload == "c4":
        run_streaming(args)
        return

    import torch
    import torch.distributed as dist
    from miocodec_b200 import binding, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa_node(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        with stdout_to_stderr():
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()                       # first collective: communicator set-up (and NCCL's banner) happens here
            torch.cuda.synchronize()

    if rank == 0:
        synth.ensure_assets()
    if world > 1:
        dist.barrier()
    gguf, _ = synth.ensure_assets()
    codec = binding.MioCodec(gguf, local, gemm=args.gemm, max_chunk_codes=args.chunk)
    stream = torch.cuda.current_stream()
    codec.set_stream(stream.cuda_stream)
    wl = make_workload(args.workload, rank, world)
    total_codes = int(wl["cu"][-1])
    audio_s = total_codes * codec.samples_per_token / codec.sample_rate

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput (`value`)
    codec.batch_upload(wl["codes"], wl["cu"], wl["voice_table"], wl["voice_ids"])
    codec.batch_sync()
    for _ in range(args.warmup):
        codec.batch_run(False)
    codec.batch_sync()
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        codec.batch_run(False)
    ev1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    codec.batch_sync()
    dev_ms = ev0.elapsed_time(ev1)
    launches = codec.kernel_launches() * args.steps
    # ---- per-stage device times: a separate pass with CUDA events around every launch (the events cost a few
    # microseconds of serialisation per launch, so this pass is NOT the one `value` is taken from)
    prof_steps = min(args.steps, 3)
    codec.profile_enable(True)
    pe0, pe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pe0.record(stream)
    for _ in range(prof_steps):
        codec.batch_run(False)
    pe1.record(stream)
    codec.batch_sync()
    prof_ms = pe0.elapsed_time(pe1)
    prof = codec.profile_read()
    codec.profile_enable(False)

    # ---- end to end through the host-buffer C-ABI call (pinned host buffers; H2D + D2H inside the timed region)
    h_codes = binding.pinned_empty(len(wl["codes"]), np.int32); h_codes[:] = wl["codes"]
    h_wave = binding.pinned_empty(total_codes * codec.samples_per_token, np.float32)
    for _ in range(2):
        codec.decode_wave_batch(h_codes, wl["cu"], wl["voice_table"], wl["voice_ids"], out=h_wave)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        codec.decode_wave_batch(h_codes, wl["cu"], wl["voice_table"], wl["voice_ids"], out=h_wave)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    h2d = codec.h2d_bytes()
    d2h = int(h_wave.nbytes)
    # the same call delivering PCM16 (what wav_write stores; half the D2H bytes, converted on the device)
    h_pcm = binding.pinned_empty(total_codes * codec.samples_per_token, np.int16)
    for _ in range(2):
        codec.decode_pcm16_batch(h_codes, wl["cu"], wl["voice_table"], wl["voice_ids"], out=h_pcm)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        codec.decode_pcm16_batch(h_codes, wl["cu"], wl["voice_table"], wl["voice_ids"], out=h_pcm)
    torch.cuda.synchronize()
    e2e_pcm_s = time.perf_counter() - t0
    d2h_pcm = int(h_pcm.nbytes)
    del h_pcm
    # C2 is the latency configuration: p50 / p99 of 100 synchronous single-utterance calls, host to host
    latency = None
    if args.workload == "c2" and not args.no_latency:
        emb = wl["voice_table"][int(wl["voice_ids"][0])]
        lat = []
        for i in range(103):
            t1 = time.perf_counter()
            codec.decode_wave(wl["codes"], emb)
            if i >= 3:
                lat.append((time.perf_counter() - t1) * 1e3)
        lat.sort()
        latency = {"p50_ms": lat[len(lat) // 2], "p99_ms": lat[min(len(lat) - 1, int(len(lat) * 0.99))], "min_ms": lat[0],
                   "runs": len(lat), "call": "mio_decode_wave (pageable host buffers, synchronous)"}
    assert os.environ.get("MIO_EXP_SKIP_NORMS") == "1" or np.isfinite(h_wave[:: max(1, len(h_wave) // 4096)]).all()

    # ---- the other BASELINE configurations, as blocks of the same line (default run only)
    extra = {}
    if args.workload == "c3" and not args.no_extra:
        del h_wave
        extra["c5"] = c5_block(codec, stream, rank, world, local, barrier)
        if world == 1:
            extra["c2"] = c2_block(codec, stream, cpu=not args.no_cpu_baseline)
            extra["c4"] = c4_block(codec, cpu=not args.no_cpu_baseline)

    # ---- max over ranks / whole-job aggregation
    t = torch.tensor([dev_ms, e2e_s * 1e3, e2e_pcm_s * 1e3], dtype=torch.float64, device="cuda")
    a = torch.tensor([audio_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(a, op=dist.ReduceOp.SUM)
    dev_ms_max, e2e_ms_max = float(t[0]), float(t[1])
    audio_all = float(a[0])
    value = audio_all * args.steps / (dev_ms_max / 1e3)
    e2e_value = audio_all * args.steps / (e2e_ms_max / 1e3)
    e2e_pcm_ms_max = float(t[2])

    if rank == 0:
        peaks = measured_peaks()
        g = prof["gemm_split3"]
        ach_tf = g["flops"] / (g["ms"] / 1e3) / 1e12 if g["ms"] > 0 else 0.0
        stages = {}
        for k, v in prof.items():
            if v["launches"] == 0 or v["ms"] <= 0:
                continue
            s = dict(ms_per_step=v["ms"] / prof_steps, launches_per_step=v["launches"] / prof_steps, share=v["ms"] / prof_ms)
            if v["flops"] > 0:
                s.update(bound="tensor", achieved=v["flops"] / (v["ms"] / 1e3) / 1e12, unit="TFLOP/s")
                s["frac"] = s["achieved"] / peaks["tf_sus"]
            elif v["bytes"] > 0:
                s.update(bound="hbm", achieved=v["bytes"] / (v["ms"] / 1e3) / 1e9, unit="GB/s")
                s["frac"] = s["achieved"] / peaks["hbm"]
            stages[k] = s
        # DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture of this workload
        traffic, traffic_src = None, None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if args.workload == "c3" and os.path.exists(tpath):
            tj = json.load(open(tpath))
            traffic, traffic_src = tj["gemm_split3"]["dram_bytes_per_launch"], tj["source"]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True,
            "scaling": "strong" if args.workload == "c5" else "weak",   # C5 shards ONE 4096-utterance sweep over the ranks
            "vs_baseline": None,
            "dtype": "f16x2-split operands, f32 accumulate (f16 operands in conv1d like the reference)", "data": "synthetic",
            "config": bench_config(wl, world), "cpus_bound_per_rank": numa,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms_max / args.steps},
            "e2e_pcm16": {"value": audio_all * args.steps / (e2e_pcm_ms_max / 1e3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                          "d2h_bytes_per_step": d2h_pcm, "ms_per_step": e2e_pcm_ms_max / args.steps,
                          "call": "mio_decode_pcm16_batch (SURVEY 8f F1: PCM16 made on the device)"},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"kernel": "gemm_tc3_kernel<BN, 3, epilogue> (tcgen05 split-precision tap-GEMM, all linear/ConvT layers)",
                         "bound": "tensor", "achieved": ach_tf, "peak": peaks["tf_sus"], "unit": "TFLOP/s",
                         "frac": ach_tf / peaks["tf_sus"], "traffic": traffic,
                         "executed_mma_frac": 3.0 * ach_tf / peaks["tf_sus"], "traffic_source": traffic_src,
                         "note": f"algorithmic FLOPs 2*M*N*K*taps per launch over the LIVE rows (the pad rows of the 16-row utterance slots are computed but not counted; 3 MMA passes executed per algorithmic FLOP; executed_mma_frac can "
                                 f"pass 1.0 because the peak is cuBLAS's measured sustained bf16 rate, not a hardware bound); "
                                 f"peak = {peaks['src']} sustained dense bf16; avg launch {g['ms'] / max(1, g['launches']):.3f} ms "
                                 f"over {int(g['launches'])} launches (separate event-instrumented pass of {prof_steps} steps, {prof_ms / prof_steps:.1f} ms/step); "
                                 f"share of step {g['ms'] / prof_ms:.2f}"},
            "stages": stages,
            **({"latency": latency} if latency else {}),
            **extra,
            "whole_path": {"algorithmic_tflops": FLOP_PER_CODE * total_codes * world * args.steps / (dev_ms_max / 1e3) / 1e12},
        }
        if world == 1 and not args.no_cpu_baseline:
            try:
                from oracle import ref
                if ref.available():
                    n_utts = 8 if args.workload != "c2" else 1
                    v, a_s, busy, wall = cpu_reference_run(wl, n_utts, 1)
                    line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": 4, "kind": "reference",
                                            "sample": f"first {min(n_utts, len(wl['lens']))} utterances of the workload ({a_s:.0f} audio-s, {busy:.1f} s of CPU time), "
                                                      f"1 reference context = 4 ggml threads + 1-thread ISTFT (the reference's fixed default)"}
                else:
                    line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref unavailable on this host"}
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"failed: {e}"}
        # the figures a reader of a truncated log needs, as the LAST key of the line (a 1 500-byte tail of the output keeps it)
        c5, c2, c4 = line.get("c5") or {}, line.get("c2") or {}, line.get("c4") or {}
        line["summary"] = {"n_gpus": world, "value": value, "ms_per_step": dev_ms_max / args.steps, "e2e": e2e_value,
                           "e2e_pcm16": line["e2e_pcm16"]["value"], "roofline_frac": line["roofline"]["frac"],
                           "c5": c5.get("value"),
                           "c5_e2e": {k: v.get("value") for k, v in (c5.get("e2e") or {}).items() if isinstance(v, dict) and "value" in v},
                           "c2_device_ms": c2.get("device_ms"), "c4": c4.get("value")}
        print(json.dumps(line))
    codec.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
