"""Small end-to-end workload for compute-sanitizer (memcheck / racecheck / synccheck / initcheck): the reference's
20-code utterance, a ragged 3-utterance batch and a 3-stream streaming session.
usage: compute-sanitizer --tool memcheck python tools/sanitize_run.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))
from miocodec_b200 import binding, synth  # noqa: E402

gguf, voices = synth.ensure_assets()
c = binding.MioCodec(gguf, 0, max_chunk_codes=512)
emb = synth.synth_voice(0)
w = c.decode_wave(synth.REFERENCE_TEST_CODES, emb)
assert np.isfinite(w).all() and w.shape == (20 * 1764,)
lens = [5, 33, 17]
cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
vt = np.stack([synth.synth_voice(v) for v in range(2)])
out = c.decode_wave_batch(synth.synth_codes(int(cu[-1]), 3), cu, vt, np.array([0, 1, 0], dtype=np.int32), peak_normalize=True)
assert np.isfinite(out).all()
pcm = c.decode_pcm16_batch(synth.synth_codes(int(cu[-1]), 3), cu, vt, np.array([0, 1, 0], dtype=np.int32))
assert pcm.shape == out.shape
N = [61, 40, 23]
streams = [c.open_stream(synth.synth_voice(i)) for i in range(3)]
codes = [synth.synth_codes(n, 50 + i) for i, n in enumerate(N)]
tot = [0, 0, 0]
for tick in range(max(N)):
    for i, s in enumerate(streams):
        if tick < N[i]:
            s.push(codes[i][tick])
        if tick == N[i] - 1:
            s.finish()
    c.step_streams(streams)
    for i, s in enumerate(streams):
        a, last = s.read()
        tot[i] += len(a)
assert tot == [n * 1764 for n in N], tot
for s in streams:
    s.close()
spec = np.random.default_rng(0).standard_normal((45, 394)).astype(np.float32)
assert c.istft(spec).shape == (45 * 98,)
c.close()
print("sanitize_run: ok")
