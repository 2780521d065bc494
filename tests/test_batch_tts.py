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


@pytest.fixture(scope="module")
def demo(assets):
    """Runs tests/batch_tts_demo.cpp once; yields (output prefix, stderr)."""
    exe = os.path.join(PKG, "batch_tts_demo")
    assert os.path.exists(exe)
    with tempfile.TemporaryDirectory() as td:
        prefix = os.path.join(td, "out")
        r = subprocess.run([exe, assets[0], assets[1][0], assets[1][1], prefix], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout + r.stderr
        yield prefix, r.stderr


def token_text(codes):
    return "noise " + "".join(f"<|s_{int(c)}|>" for c in codes) + " tail"


@pytest.mark.gpu
def test_batch_tts_against_the_reference_test_to_speech(assets, demo):
    """F3 against the reference's OWN caller: TestToSpeech::synthesize_to_vector / synthesize_to_file / synthesize_stream
    with skip_llm (oracle/_ref: unmodified src/test-to-speech.cpp, CPU codec) on the very texts and voices the demo fed
    to BatchTestToSpeech.  Same refusals, same sample counts, same WAV header, same stream chunking; samples at the
    parity floor (the codecs differ, the caller logic must not)."""
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref is not built on this host")
    prefix, _ = demo
    gguf, voices = assets
    c0, c1 = lcg_codes(20, 1), lcg_codes(33, 2)
    # refusals: nothing parses -> the reference returns false as well
    assert ref.ref_tts_skip_llm(gguf, voices[0], "no tokens here <|s_|>") is None
    for i, (codes, vi) in enumerate(((c0, 0), (c1, 1))):
        # raw vectors (peak normalisation off)
        want = ref.ref_tts_skip_llm(gguf, voices[vi], token_text(codes), peak_normalize=False)
        got = np.fromfile(prefix + f".v{i}.f32", dtype=np.float32)
        assert got.shape == want.shape == (len(codes) * 1764,)
        s, m = snr_db(got, want), float(np.abs(got - want).max())
        print(f"utterance {i}: BatchTestToSpeech vs TestToSpeech (vector) SNR {s:.2f} dB, max-abs {m:.2e}")
        assert s > 58.0 and m < 4.5e-3
        # WAV files (peak normalisation on, the reference default): header identical, PCM within the codec distance
        ref_wav = prefix + f".ref{i}.wav"
        normed = ref.ref_tts_skip_llm(gguf, voices[vi], token_text(codes), peak_normalize=True, wav_path=ref_wav)
        a, b = open(prefix + f".{i}.wav", "rb").read(), open(ref_wav, "rb").read()
        assert len(a) == len(b) and a[:44] == b[:44]
        pa, pb = np.frombuffer(a[44:], dtype=np.int16).astype(np.int32), np.frombuffer(b[44:], dtype=np.int16).astype(np.int32)
        assert abs(int(np.abs(pa).max()) - int(np.abs(pb).max())) <= 1           # both peak at 0.95 full scale
        s_pcm = snr_db(pa, pb)
        print(f"utterance {i}: WAV PCM16 SNR {s_pcm:.2f} dB, max |diff| {int(np.abs(pa - pb).max())} LSB")
        assert s_pcm > 57.0 and np.abs(pa - pb).max() <= int(4.5e-3 * 0.95 / np.abs(want).max() * 32767) + 2
        assert abs(np.abs(normed).max() - 0.95) < 1e-6
    # skip_llm stream of utterance 0: the reference's chunk sizes (src/test-to-speech.cpp:419-428 -> emit_range)
    want, chunks, st = ref.ref_stream_session(gguf, voices[0], c0, skip_llm=True, chunk_samples=4096)
    got = np.fromfile(prefix + ".s0.f32", dtype=np.float32)
    assert got.shape == want.shape
    assert len(chunks) == (len(want) + 4095) // 4096 and st["last_flags"] == 1      # what the demo CHECKs for its own callbacks
    assert snr_db(got, want) > 58.0


@pytest.mark.gpu
def test_batch_tts_demo(assets, codec, demo):
    from miocodec_b200 import binding
    if True:
        prefix, stderr = demo
        # the reference's diagnostics for the two rejected utterances
        assert "no speech codes parsed from text" in stderr and "voice model is not ready" in stderr
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
