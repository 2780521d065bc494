"""GPU experiment: end-to-end (host buffers) time of the C3 batch against the chunk size / tapering of the
host-delivery pipeline, next to the raw pinned D2H bandwidth.  usage: python tools/e2e_sweep.py"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))
from miocodec_b200 import binding, synth  # noqa: E402
import bench  # noqa: E402

gguf, _ = synth.ensure_assets()
codec = binding.MioCodec(gguf, 0)
wl = bench.make_workload("c3", 0, 1)
total = int(wl["cu"][-1])
h_codes = binding.pinned_empty(len(wl["codes"]), np.int32); h_codes[:] = wl["codes"]
h_wave = binding.pinned_empty(total * codec.samples_per_token, np.float32)
h_pcm = binding.pinned_empty(total * codec.samples_per_token, np.int16)

# raw D2H rate
d = torch.empty(h_wave.size, dtype=torch.float32, device="cuda")
hp = torch.from_numpy(h_wave)
for _ in range(2):
    hp.copy_(d, non_blocking=True); torch.cuda.synchronize()
t0 = time.perf_counter(); hp.copy_(d, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print(f"raw pinned D2H: {h_wave.nbytes / dt / 1e9:.1f} GB/s ({dt * 1e3:.2f} ms for {h_wave.nbytes / 1e6:.0f} MB)")
del d

def run(fn, n=4):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    return min(ts), sum(ts) / len(ts)

codec.batch_upload(wl["codes"], wl["cu"], wl["voice_table"], wl["voice_ids"]); codec.batch_sync()
print("device only (one chunk): min %.2f avg %.2f ms" % run(lambda: (codec.batch_run(False), codec.batch_sync())))
print("upload only: min %.2f avg %.2f ms" % run(lambda: (codec.batch_upload(h_codes, wl["cu"], wl["voice_table"], wl["voice_ids"]), codec.batch_sync())))
for tail in (0, 8, 12, 15, 18, 22, 30):
    codec.set_option("fetch_tail_percent", str(tail))
    a = run(lambda: codec.decode_wave_batch(h_codes, wl["cu"], wl["voice_table"], wl["voice_ids"], out=h_wave))
    b = run(lambda: codec.decode_pcm16_batch(h_codes, wl["cu"], wl["voice_table"], wl["voice_ids"], out=h_pcm))
    print(f"tail {tail:2d} %: f32 min {a[0]:.2f} avg {a[1]:.2f} ms | pcm16 min {b[0]:.2f} avg {b[1]:.2f} ms")
