"""CPU tests of the oracle itself: the torch restatement (oracle/port.py) against golden vectors produced by the
unmodified reference, the compiled reference (oracle/_ref) against the same vectors when it is present, and the
known-answer structure of the path (frame / sample counts)."""
import hashlib
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_golden, snr_db


def test_weights_digest_matches_golden(weights):
    # the goldens are only meaningful for bit-identical synthetic weights
    _, W = weights
    from miocodec_b200 import synth
    _, tensors = synth.synth_tensors(0)
    h = hashlib.sha256()
    for name, arr in tensors:
        h.update(name.encode())
        h.update(np.ascontiguousarray(arr).tobytes())
    meta = dict(l.strip().split("=", 1) for l in open(os.path.join(GOLDEN, "meta.txt")))
    assert h.hexdigest() == meta["weights_sha256"]
    assert len(tensors) == 227                      # 227 tensors
    assert sum(int(np.prod(a.shape)) for _, a in tensors if a.dtype != np.int32) == 92_344_074


@pytest.mark.parametrize("case", ["ref20", "rand33", "one"])
def test_port_matches_reference_golden(weights, case):
    """The port is an independent fp32 evaluation; the reference's own FP16 operand rounding in conv1d
    (ggml.c:4416) makes its output discontinuous in fp32 round-off, which caps ANY independent evaluation at
    ~60 dB (DESIGN.md 'parity floor'); counts and indexing are exact."""
    from miocodec_b200 import synth
    from oracle import port
    hp, W = weights
    g = load_golden(case)
    codes, vid = g["codes"], int(g["voice"])
    spec = port.decode(W, hp, codes, synth.synth_voice(vid))
    assert spec.shape == g["spec"].shape == (18 * len(codes), 394)
    assert snr_db(spec, g["spec"]) > 55.0
    wave = port.istft(spec)
    assert wave.shape == g["wave"].shape == (1764 * len(codes),)
    assert snr_db(wave, g["wave"]) > 55.0
    assert np.abs(wave - g["wave"]).max() < 5e-3 * max(1.0, np.abs(g["wave"]).max())


def test_port_istft_matches_reference_golden():
    from oracle import port
    for case in ("ref20", "istft50"):
        g = load_golden(case)
        wave = port.istft(g["spec"])
        assert wave.shape == g["wave"].shape
        assert snr_db(wave, g["wave"]) > 110.0
        assert np.abs(wave - g["wave"]).max() < 2e-5 * max(1.0, np.abs(g["wave"]).max())


def test_compiled_reference_reproduces_golden(assets):
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built on this machine")
    from miocodec_b200 import synth
    R = ref.Reference(assets[0])
    g = load_golden("ref20")
    spec = R.decode_spec(g["codes"], synth.synth_voice(int(g["voice"])))
    meta = dict(l.strip().split("=", 1) for l in open(os.path.join(GOLDEN, "meta.txt")))
    if R.variant == meta["variant"]:
        assert np.array_equal(spec, g["spec"])          # bitwise deterministic
    else:
        assert snr_db(spec, g["spec"]) > 55.0           # other SIMD width => other summation order
    wave = R.istft(g["spec"])
    assert snr_db(wave, g["wave"]) > 120.0
    assert R.samples_per_token == 1764 and R.n_fft == 392 and R.hop == 98 and R.sample_rate == 44100
    R.close()


def test_stream_policy_cadence():
    """Cadence known answers: decodes fire at n = 60, 100, 140, ... and once more at the end."""
    from oracle import port
    calls = []

    def fake_decode(c):
        calls.append(len(c))
        return np.arange(len(c) * 1764, dtype=np.float32)

    codes = np.zeros(230, dtype=np.int32)
    out, st = port.stream_policy(fake_decode, codes)
    assert calls == [60, 100, 140, 180, 220, 230]
    assert st["decoded_codes"] == sum(calls) and st["emitted_samples"] == 230 * 1764 == len(out)
    # with a waveform that does not depend on the prefix, only the cross-faded heads differ from offline
    ref = np.arange(230 * 1764, dtype=np.float32)
    diff = np.nonzero(out != ref)[0]
    assert len(diff) == 0 or diff.max() < 230 * 1764
    # a multiple of 20 decodes the last prefix twice but emits nothing twice
    calls.clear()
    out, st = port.stream_policy(fake_decode, np.zeros(100, dtype=np.int32))
    assert calls == [60, 100, 100] and len(out) == 100 * 1764
    # short utterance: single final decode
    calls.clear()
    out, st = port.stream_policy(fake_decode, np.zeros(7, dtype=np.int32))
    assert calls == [7] and len(out) == 7 * 1764


# ---- streaming policy: the port against the reference's OWN orchestrator ---------------------------------------------
@pytest.mark.parametrize("N", [7, 47, 100, 230, 700])
def test_stream_policy_matches_reference_golden_counters(N):
    """Goldens were produced by the UNMODIFIED TestToSpeech::synthesize_stream_profiled (LLM branch, src/test-to-speech.cpp
    :430-625) with oracle/stub_llama.cpp emitting the codes (tests/golden/make_golden.py).  The port must make the same
    decisions: decode calls, decoded codes, emitted samples, and the size of every callback chunk."""
    from oracle import port
    g = load_golden(f"stream{N}")
    sizes = []

    def fake_decode(c):
        return np.zeros(len(c) * 1764, dtype=np.float32)

    out, st = port.stream_policy(fake_decode, g["codes"], chunk_log=sizes)
    assert st["decode_calls"] == int(g["decode_calls"]) and st["decoded_codes"] == int(g["decoded_codes"])
    assert st["emitted_samples"] == int(g["emitted_samples"]) == 1764 * N == len(out)
    assert int(g["llm_tokens"]) == N
    assert sizes == [int(v) for v in g["chunks"]]


@pytest.mark.parametrize("N", [7, 47, 100])
def test_stream_policy_is_bit_exact_against_the_reference_orchestrator(assets, N):
    """Same codec (the compiled reference) under both policies: the port's emitted audio must equal the reference
    orchestrator's bit for bit — cadence, llround indexing, chunking and the crossfade arithmetic (:388-395)."""
    from miocodec_b200 import synth
    from oracle import port, ref
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    g = load_golden(f"stream{N}")
    R = ref.Reference(assets[0])
    emb = synth.synth_voice(int(g["voice"]))
    out, st = port.stream_policy(lambda c: R.decode_wave(c, emb), g["codes"])
    meta = dict(l.strip().split("=", 1) for l in open(os.path.join(GOLDEN, "meta.txt")))
    if R.variant == meta["variant"]:
        assert np.array_equal(out, g["audio"])
        assert hashlib.sha256(out.tobytes()).digest() == bytes(g["audio_sha256"])
    else:
        assert out.shape == g["audio"].shape and snr_db(out, g["audio"]) > 57.0
    R.close()


def test_reference_orchestrator_live_session(assets):
    """The reference's own streaming code runs here (stub LLM): a live session reproduces its golden, and its skip_llm
    branch (:419-428) equals the offline decode bit for bit."""
    from miocodec_b200 import synth
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    g = load_golden("stream100")
    vid = int(g["voice"])
    audio, chunks, st = ref.ref_stream_session(assets[0], assets[1][vid], g["codes"])
    assert [int(v) for v in chunks] == [int(v) for v in g["chunks"]]
    assert (st["decode_calls"], st["decoded_codes"], st["emitted_samples"], st["llm_tokens"], st["last_flags"]) == (3, 260, 176400, 100, 1)
    R = ref.Reference(assets[0])
    a2, ch2, st2 = ref.ref_stream_session(assets[0], assets[1][vid], g["codes"][:30], skip_llm=True)
    assert np.array_equal(a2, R.decode_wave(g["codes"][:30], synth.synth_voice(vid))) and st2["decode_calls"] == 0
    R.close()


VARIANTS = {
    "kernels [7,5]": (dict(hp=dict(up_kernels=(7, 5))), 57.0),
    "kernels [15,3]": (dict(hp=dict(up_kernels=(15, 3))), 57.0),
    "factors [2,2], widths 384/192, out_proj 448": (dict(hp=dict(up_factors=(2, 2), up_kernels=(4, 10), up_channels=(384, 192), out_proj_dim=448)), 57.0),
    "factors [1,9]": (dict(hp=dict(up_factors=(1, 9), up_kernels=(3, 27))), 57.0),
    "F16 matrices + kernels": (dict(matrices="f16"), 47.0),
    "BF16 matrices": (dict(matrices="bf16"), 31.0),
}


@pytest.mark.parametrize("name", list(VARIANTS))
def test_port_follows_the_reference_on_other_geometries_and_storage_types(tmp_path, name):
    """SURVEY.md row A16 / Appendix A 'Linear': the geometry of the upsampler is read from the file, and matrices
    stored as F16 / BF16 make ggml round the activations to that type.  The port, told the storage types, must land at
    the reference's own cross-build distance for such a file (F32 ~59.5 dB, F16 ~49.5 dB, BF16 ~34.5 dB); told
    nothing, it sits measurably further away (checked for F16)."""
    from miocodec_b200 import synth
    from oracle import port, ref
    if not ref.available():
        pytest.skip("oracle/_ref is not built on this host")
    kw, bar = VARIANTS[name]
    path = str(tmp_path / "variant.gguf")
    hp = synth.write_codec_gguf(path, seed=2, **kw)
    _, tensors = synth.synth_tensors(2, kw.get("hp"))
    W, store = dict(tensors), {}
    kind = kw.get("matrices", "f32")
    if kind != "f32":
        for n, a in tensors:
            if a.dtype == np.float32 and a.ndim >= 2:
                if a.ndim == 2:
                    W[n] = a.astype(np.float16).astype(np.float32) if kind == "f16" else synth.bf16_round(a)
                    store[n] = kind
                elif kind == "f16":
                    W[n] = a.astype(np.float16).astype(np.float32)
                    store[n] = "f16"
    codes, emb = synth.synth_codes(40, 3), synth.synth_voice(2)
    R = ref.Reference(path)
    want = R.decode_spec(codes, emb)
    R.close()
    got = port.decode(W, hp, codes, emb, store=store)
    fpc = hp["samples_per_token"] // hp["hop_length"]
    assert got.shape == want.shape == (fpc * 40, 394)
    s = snr_db(got, want)
    print(f"{name}: port vs reference {s:.2f} dB")
    assert s > bar
    if kind == "f16":
        blind = snr_db(port.decode(W, hp, codes, emb), want)
        print(f"{name}: port ignoring the storage type {blind:.2f} dB")
        assert blind < s - 0.2
