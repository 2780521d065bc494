"""Per-round wall times of a C4 session in the split (begin / read / end) and synchronous forms."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))
import numpy as np
from miocodec_b200 import binding, synth
gguf, _ = synth.ensure_assets()
codec = binding.MioCodec(gguf, 0)
S, N = 1024, 700


def drain(streams):
    n = 0
    for st in streams:
        while True:
            a, _ = st.read(1 << 18)
            n += a.size
            if a.size < (1 << 18):
                break
    return n


def session(split, verbose):
    streams = [codec.open_stream(synth.synth_voice(i % 16)) for i in range(S)]
    codes = [synth.synth_codes(N, 4000 + i) for i in range(S)]
    t00 = time.perf_counter()
    for k in range(0, N, 20):
        for i, st in enumerate(streams):
            st.push(codes[i][k:k + 20])
            if k + 20 >= N:
                st.finish()
        t0 = time.perf_counter()
        if split:
            fl = codec.step_streams_begin(streams); t1 = time.perf_counter()
            n = drain(streams) if fl else 0; t2 = time.perf_counter()
            codec.step_streams_end(streams); t3 = time.perf_counter()
        else:
            codec.step_streams(streams); t1 = t2 = time.perf_counter()
            n = drain(streams); t3 = time.perf_counter()
            t2 = t3
        if verbose and k >= 400:
            print(f"  k={k}: begin/step {1e3 * (t1 - t0):.1f} ms, read {1e3 * ((t2 - t1) if split else (t3 - t1)):.1f} ms ({n} samples), end {1e3 * (t3 - t2):.1f} ms")
    drain(streams)
    dt = time.perf_counter() - t00
    for st in streams:
        st.close()
    return dt


session(False, False)
for split in (False, True, False, True):
    print("split" if split else "sync", "%.3f s" % session(split, split is True or True))
