"""the batched / multi-stream caller (shim/batch-tts.h, the skip_llm and live-stream forms of
the reference's TestToSpeech, src/test-to-speech.h:26-124) — the C++ test program drives it, this file checks what
it produced against the single-utterance path and the streaming policy."""
import os
import struct
import subprocess
import tempfile

import numpy as np
import pytest

from conftest import ROOT, snr_db

PKG = os.path.join(ROOT, "miotts-llama.cpp_b200", "miocodec_b200")


def lcg_codes(n, seed):
    s, out = seed, []
    for _ in range(n):
        s = (s * 1664525 + 1013904223) & 0xFFFFFFFF
        out.append((s >> 8) % 12800)
    return np.array(out, dtype=np.int32)


def test_batch_tts_symbols_are_exported():
    """CPU check: the shim library carries the BatchTestToSpeech entry points."""
    lib = os.path.join(PKG, "libmiocodec_shim.so")
    if not os.path.exists(lib):
        import __graft_entry__
        __graft_entry__.build()
    syms = subprocess.run(["nm", "-D", "--defined-only", "-C", lib], capture_output=True, text=True, check=True).stdout
    for name in ("BatchTestToSpeech::synthesize_to_vectors", "BatchTestToSpeech::synthesize_to_files",
                 "BatchTestToSpeech::synthesize_streams", "BatchTestToSpeech::open_stream", "BatchTestToSpeech::push_code",
                 "BatchTestToSpeech::finish_stream", "BatchTestToSpeech::poll", "BatchTestToSpeech::stream_profile"):
        assert name in syms, name


@pytest.mark.gpu
def test_batch_tts_demo(assets, codec):
    from miocodec_b200 import binding
    exe = os.path.join(PKG, "batch_tts_demo")
    assert os.path.exists(exe)
    with tempfile.TemporaryDirectory() as td:
        prefix = os.path.join(td, "out")
        r = subprocess.run([exe, assets[0], assets[1][0], assets[1][1], prefix], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout + r.stderr
        # the reference's diagnostics for the two rejected utterances
        assert "no speech codes parsed from text" in r.stderr and "voice model is not ready" in r.stderr
        v = [binding.load_voice_embedding(assets[1][i]) for i in (0, 1)]
        c0, c1 = lcg_codes(20, 1), lcg_codes(33, 2)
        w0, w1 = codec.decode_wave(c0, v[0]), codec.decode_wave(c1, v[1])
        a0 = np.fromfile(prefix + ".v0.f32", dtype=np.float32)
        a1 = np.fromfile(prefix + ".v1.f32", dtype=np.float32)
        assert snr_db(a0, w0) > 55.0 and snr_db(a1, w1) > 55.0
        assert snr_db(np.fromfile(prefix + ".s0.f32", dtype=np.float32), w0) > 55.0
        # WAV files: the reference writer's header, PCM16 of the peak-normalised waveform
        for path, w in ((prefix + ".0.wav", w0), (prefix + ".1.wav", w1)):
            raw = open(path, "rb").read()
            assert raw[:44] == binding.wav_header(44100, len(w))
            pcm = np.frombuffer(raw[44:], dtype=np.int16)
            want = w * (0.95 / np.abs(w).max())
            assert len(pcm) == len(w) and abs(int(np.abs(pcm).max()) - int(0.95 * 32767)) <= 2
            assert snr_db(pcm.astype(np.float32) / 32767.0, want) > 50.0
        # live streams against the Python mirror of the same policy driven the same way (one code per tick)
        for i, (seed, vi) in enumerate(((11, 0), (12, 1), (13, 0))):
            gen = lcg_codes(100, seed)
            s = codec.open_stream(v[vi])
            chunks = []
            for t in range(100):
                s.push(gen[t])
                codec.step_streams([s])
                a, _ = s.read()
                chunks.append(a)
            s.finish()
            codec.step_streams([s])
            a, last = s.read()
            chunks.append(a)
            assert last
            want = np.concatenate(chunks)
            s.close()
            got = np.fromfile(prefix + f".l{i}.f32", dtype=np.float32)
            assert got.shape == want.shape == (100 * 1764,)
            assert snr_db(got, want) > 55.0
