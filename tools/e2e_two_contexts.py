"""End-to-end C3 throughput with ONE context (serial calls) and with TWO contexts driven by two host threads on the same
GPU (the upload / tail / D2H of one batch under the compute of the other)."""
import os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))
import numpy as np
from miocodec_b200 import binding, synth

gguf, _ = synth.ensure_assets()
lens = np.full(256, 250)
cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
vt = np.stack([synth.synth_voice(v) for v in range(16)]).astype(np.float32)
vids = (np.arange(256) % 16).astype(np.int32)
audio_s = int(cu[-1]) * 1764 / 44100.0


def make(seed):
    c = binding.MioCodec(gguf, 0)
    codes = binding.pinned_empty(int(cu[-1]), np.int32); codes[:] = synth.synth_codes(int(cu[-1]), seed)
    out = binding.pinned_empty(int(cu[-1]) * 1764, np.float32)
    for _ in range(3):
        c.decode_wave_batch(codes, cu, vt, vids, out=out)
    return c, codes, out


def loop(ctx, n):
    c, codes, out = ctx
    for _ in range(n):
        c.decode_wave_batch(codes, cu, vt, vids, out=out)


a, b = make(3), make(4)
STEPS = 10
for rep in range(2):
    t0 = time.perf_counter(); loop(a, STEPS); dt1 = time.perf_counter() - t0
    th = [threading.Thread(target=loop, args=(x, STEPS // 2)) for x in (a, b)]
    t0 = time.perf_counter()
    for t in th: t.start()
    for t in th: t.join()
    dt2 = time.perf_counter() - t0
    print(f"one context: {audio_s * STEPS / dt1:.0f} audio-s/s ({dt1 / STEPS * 1e3:.2f} ms/batch); two contexts: {audio_s * STEPS / dt2:.0f} audio-s/s ({dt2 / STEPS * 1e3:.2f} ms/batch)")
