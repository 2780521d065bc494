"""Where a C4 streaming session spends its wall time: pushes, mio_streams_step (device work + emit copy), reads."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))
import numpy as np
from miocodec_b200 import binding, synth

gguf, _ = synth.ensure_assets()
codec = binding.MioCodec(gguf, 0)
S, N = 1024, 700
SPLIT = os.environ.get('C4_SPLIT', '0') == '1'


def session(S, N):
    streams = [codec.open_stream(synth.synth_voice(i % 16)) for i in range(S)]
    codes = [synth.synth_codes(N, 4000 + i) for i in range(S)]
    t_push = t_step = t_read = 0.0
    emitted = 0
    for k in range(0, N, 20):
        t0 = time.perf_counter()
        for i, st in enumerate(streams):
            st.push(codes[i][k:k + 20])
            if k + 20 >= N:
                st.finish()
        t1 = time.perf_counter()
        if SPLIT:
            codec.step_streams_begin(streams)
        else:
            codec.step_streams(streams)
        t2 = time.perf_counter()
        for st in streams:
            while True:
                a, _ = st.read(1 << 18)
                emitted += a.size
                if a.size < (1 << 18):
                    break
        t3 = time.perf_counter()
        if SPLIT:
            codec.step_streams_end(streams)
        t4 = time.perf_counter()
        t_push += t1 - t0; t_step += (t2 - t1) + (t4 - t3); t_read += t3 - t2
    if SPLIT:
        t3 = time.perf_counter()
        for st in streams:
            while True:
                a, _ = st.read(1 << 18)
                emitted += a.size
                if a.size < (1 << 18):
                    break
        t_read += time.perf_counter() - t3
    for st in streams:
        st.close()
    return emitted, t_push, t_step, t_read


session(32, 100)
if os.environ.get('C4_PROFILE'):
    codec.profile_enable(True)
e, a, b, c = session(S, N)
print(f'first session (cold buffers): push {a:.3f} s, step {b:.3f} s, read {c:.3f} s; total {a + b + c:.3f} s')
print('---- second session', file=sys.stderr, flush=True)
e, a, b, c = session(S, N)
prof = codec.profile_read()
print('device kernel time by category (s):', {k: round(v['ms'] / 1e3, 3) for k, v in prof.items() if v['ms'] > 0}, 'sum', round(sum(v['ms'] for v in prof.values()) / 1e3, 3), 'launches', sum(v['launches'] for v in prof.values()))
print(f"emitted {e / 44100:.0f} audio-s; push {a:.3f} s, step {b:.3f} s, read {c:.3f} s; total {a + b + c:.3f} s")
