"""Generates tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref, compiled from /root/reference by
oracle/build_ref.sh) on the synthetic assets of miocodec_b200.synth.  Run in the build container (where
/root/reference exists):   python tests/golden/make_golden.py

Stored per case: the input codes, the voice id, the reference spectrogram (float16-free: float32) and waveform.
The weights are NOT stored (352 MiB); they are regenerated from seed 0 and their checksum is stored instead.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))
from miocodec_b200 import synth  # noqa: E402
from oracle import ref  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def weights_digest(seed=0):
    _, tensors = synth.synth_tensors(seed)
    h = hashlib.sha256()
    for name, arr in tensors:
        h.update(name.encode())
        h.update(np.ascontiguousarray(arr).tobytes())
    return h.hexdigest()


def main():
    g, voices = synth.ensure_assets()
    R = ref.Reference(g)
    cases = {
        "ref20": (synth.REFERENCE_TEST_CODES, 0),          # the reference's own 20-code vector (tools/compare_codec.py:50-51)
        "rand33": (synth.synth_codes(33, 11), 3),           # odd length, another voice
        "one": (synth.synth_codes(1, 12), 1),               # shortest utterance
    }
    meta = dict(weights_sha256=weights_digest(0), variant=R.variant)
    for name, (codes, vid) in cases.items():
        emb = synth.synth_voice(vid)
        spec = R.decode_spec(codes, emb)
        wave = R.istft(spec)
        assert wave.shape[0] == 1764 * len(codes)
        # bitwise determinism of the reference
        assert np.array_equal(spec, R.decode_spec(codes, emb))
        np.savez_compressed(os.path.join(HERE, name + ".npz"), codes=codes.astype(np.int32), voice=np.int32(vid),
                            spec=spec.astype(np.float32), wave=wave.astype(np.float32))
        print(name, spec.shape, wave.shape, float(np.abs(wave).max()))
    # ISTFT alone on a random spectrogram (no weights involved)
    rng = np.random.default_rng(5)
    spec = rng.standard_normal((50, 394)).astype(np.float32)
    wave = ref.ref_istft_only(spec)
    np.savez_compressed(os.path.join(HERE, "istft50.npz"), spec=spec, wave=wave)
    # streaming sessions of the reference's OWN orchestrator (TestToSpeech::synthesize_stream_profiled, LLM branch, with
    # oracle/stub_llama.cpp emitting the codes): chunk sizes + StreamProfile counters for every N, the emitted audio for
    # the short ones (the long ones are compared live on the GPU box, where oracle/_ref is present)
    import hashlib as _h
    for N, vid in ((7, 0), (47, 1), (100, 2), (230, 4), (700, 5)):
        codes = synth.synth_codes(N, 70 + N)
        audio, chunks, st = ref.ref_stream_session(g, voices[vid], codes)
        assert len(audio) == 1764 * N == st["emitted_samples"]
        keep = audio if N <= 100 else np.empty(0, dtype=np.float32)
        np.savez_compressed(os.path.join(HERE, f"stream{N}.npz"), codes=codes.astype(np.int32), voice=np.int32(vid), chunks=chunks.astype(np.int64),
                            decode_calls=np.int64(st["decode_calls"]), decoded_codes=np.int64(st["decoded_codes"]),
                            emitted_samples=np.int64(st["emitted_samples"]), llm_tokens=np.int64(st["llm_tokens"]), audio=keep,
                            audio_sha256=np.frombuffer(_h.sha256(audio.tobytes()).digest(), dtype=np.uint8))
        print("stream", N, st, len(chunks))
    with open(os.path.join(HERE, "meta.txt"), "w") as f:
        for k, v in meta.items():
            f.write(f"{k}={v}\n")
    print(meta)


if __name__ == "__main__":
    main()
