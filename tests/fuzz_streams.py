"""Randomised streaming check: S concurrent streams with random lengths, push bursts, read sizes and read laziness; every
stream's concatenated output must equal the streaming-policy port (oracle/port.stream_policy, itself pinned bit for bit to
the reference's orchestrator) driven by single-utterance decodes of the same decoder, bit for bit, and the StreamProfile
counters must match.  TEST TOOL (imports oracle/).  usage: python tests/fuzz_streams.py [rounds] [seed]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))
import numpy as np
from miocodec_b200 import binding, synth
from oracle import port

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 6
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 3)
gguf, _ = synth.ensure_assets()
codec = binding.MioCodec(gguf, 0)
bad = 0
for r in range(rounds):
    S = int(rng.integers(1, 13))
    lens = rng.integers(1, 420, size=S)
    embs = [synth.synth_voice(int(rng.integers(0, 8))) for _ in range(S)]
    codes = [synth.synth_codes(int(n), 7000 + 100 * r + i) for i, n in enumerate(lens)]
    lazy = rng.random(S) < 0.3                      # these read nothing until the end (pinned pool exhaustion path)
    streams = [codec.open_stream(e) for e in embs]
    pos = np.zeros(S, dtype=np.int64)
    got = [[] for _ in range(S)]
    done = [False] * S
    while not all(pos[i] >= lens[i] for i in range(S)):
        for i, s in enumerate(streams):
            if pos[i] < lens[i] and rng.random() < 0.8:
                n = int(min(lens[i] - pos[i], rng.integers(1, 38)))
                s.push(codes[i][pos[i]: pos[i] + n])
                pos[i] += n
                if pos[i] >= lens[i]:
                    s.finish()
        codec.step_streams(streams)
        for i, s in enumerate(streams):
            if not lazy[i] and not done[i] and rng.random() < 0.7:
                a, last = s.read(int(rng.integers(1, 1 << 17)))
                got[i].append(a)
                done[i] = bool(last)
    codec.step_streams(streams)
    for i, s in enumerate(streams):
        while not done[i]:
            a, last = s.read(int(rng.integers(1000, 1 << 18)))
            got[i].append(a)
            done[i] = bool(last)
            if a.size == 0 and not last:
                codec.step_streams(streams)
        out = np.concatenate(got[i]) if got[i] else np.zeros(0, np.float32)
        want, st = port.stream_policy(lambda c, e=embs[i]: codec.decode_wave(c, e), codes[i])
        mine = s.stats()
        ok = out.shape == want.shape and np.array_equal(out, want) and mine["decode_calls"] == st["decode_calls"] and \
            mine["decoded_codes"] == st["decoded_codes"] and mine["emitted_samples"] == len(want)
        if not ok:
            bad += 1
            print(f"round {r} stream {i} (T={lens[i]}, lazy={bool(lazy[i])}): MISMATCH {out.shape} vs {want.shape}, {mine} vs {st}")
        s.close()
    print(f"round {r}: {S} streams, lengths {list(map(int, lens))}: {'ok' if not bad else 'failures so far %d' % bad}", flush=True)
print("STREAM FUZZ", "FAILED" if bad else "PASSED", f"({rounds} rounds)")
sys.exit(1 if bad else 0)
