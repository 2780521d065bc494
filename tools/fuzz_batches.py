"""Randomised batch-invariance check: every utterance of random ragged batches (random lengths, voices, workspace sizes,
delivery calls) must equal its single-utterance decode bit for bit.  usage: python tools/fuzz_batches.py [rounds] [seed]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))
import numpy as np
from miocodec_b200 import binding, synth

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 24
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 7)
gguf, _ = synth.ensure_assets()
single = binding.MioCodec(gguf, 0)
cache = {}


def ref_wave(codes, vid):
    key = (codes.tobytes(), vid)
    if key not in cache:
        cache[key] = single.decode_wave(codes, synth.synth_voice(vid))
    return cache[key]


bad = 0
for r in range(rounds):
    chunk = int(rng.choice([512, 1024, 4096, 65536]))
    c = binding.MioCodec(gguf, 0, max_chunk_codes=chunk)
    n = int(rng.integers(1, 40))
    lens = rng.integers(1, min(400, chunk - 32), size=n)
    if rng.random() < 0.3:
        lens[rng.integers(0, n)] = 1
    vids = rng.integers(0, 6, size=n).astype(np.int32)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    codes = synth.synth_codes(int(cu[-1]), 1000 + r)
    vt = np.stack([synth.synth_voice(v) for v in range(6)]).astype(np.float32)
    mode = int(rng.integers(0, 3))
    if mode == 0:
        wave = c.decode_wave_batch(codes, cu, vt, vids)
    elif mode == 1:
        c.batch_upload(codes, cu, vt, vids); c.batch_run(False); c.batch_sync(); wave = c.batch_fetch()
    else:
        pcm = c.decode_pcm16_batch(codes, cu, vt, vids)
        wave = None
    for i in range(n):
        want = ref_wave(codes[cu[i]:cu[i + 1]].copy(), int(vids[i]))
        a, b = int(cu[i]) * 1764, int(cu[i + 1]) * 1764
        if wave is not None:
            ok = np.array_equal(wave[a:b], want)
        else:
            ok = np.array_equal(pcm[a:b], np.clip(want * 32767.0, -32768.0, 32767.0).astype(np.int32).astype(np.int16))
        if not ok:
            bad += 1
            print(f"round {r} (chunk {chunk}, mode {mode}): utterance {i} of {n} (T={lens[i]}) differs")
    c.close()
    print(f"round {r}: {n} utterances, {int(cu[-1])} codes, workspace {chunk}, mode {mode}: ok" if not bad else f"round {r}: failures so far {bad}", flush=True)
print("FUZZ", "FAILED" if bad else "PASSED", f"({rounds} rounds)")
sys.exit(1 if bad else 0)
