"""TEST INFRASTRUCTURE ONLY — ctypes binding of oracle/_ref/libmio_ref_*.so.

That library is the UNMODIFIED reference hot path (miocodec_decode + istft, compiled from
/root/reference by oracle/build_ref.sh).  Only tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py may import this module; the product
(miocodec_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def _cpu_flags() -> set[str]:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def lib_path() -> str | None:
    flags = _cpu_flags()
    order = []
    if {"avx512f", "avx512bw", "avx512dq", "avx512vl", "avx512cd"} <= flags:
        order.append("v4")
    if {"avx2", "fma", "f16c", "bmi2"} <= flags:
        order.append("v3")
    forced = os.environ.get("MIO_REF_VARIANT")
    if forced:
        order = [forced]
    for tag in order:
        p = os.path.join(_HERE, "_ref", f"libmio_ref_{tag}.so")
        if os.path.exists(p):
            return p
    return None


def available() -> bool:
    return lib_path() is not None


class Reference:
    """One reference miocodec context (4 ggml threads, like the reference always uses)."""

    def __init__(self, gguf_path: str):
        p = lib_path()
        if p is None:
            raise RuntimeError("oracle/_ref not built (run oracle/build_ref.sh where /root/reference exists)")
        self.lib = L = C.CDLL(p)
        self.variant = os.path.basename(p)
        L.mio_ref_load.restype = C.c_void_p
        L.mio_ref_load.argtypes = [C.c_char_p]
        L.mio_ref_free.argtypes = [C.c_void_p]
        for n in ("sample_rate", "n_fft", "hop_length", "samples_per_token"):
            getattr(L, "mio_ref_" + n).argtypes = [C.c_void_p]
            getattr(L, "mio_ref_" + n).restype = C.c_int
        L.mio_ref_decode.restype = C.c_long
        L.mio_ref_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_long,
                                     C.POINTER(C.c_int), C.POINTER(C.c_double)]
        L.mio_ref_istft.restype = C.c_long
        L.mio_ref_istft.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_long,
                                    C.POINTER(C.c_double)]
        L.mio_ref_decode_wave.restype = C.c_long
        L.mio_ref_decode_wave.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_long,
                                          C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.mio_ref_load_voice.restype = C.c_int
        L.mio_ref_load_voice.argtypes = [C.c_char_p, C.c_void_p, C.c_int]
        self.ctx = L.mio_ref_load(gguf_path.encode())
        if not self.ctx:
            raise RuntimeError("reference miocodec_load failed: " + gguf_path)
        self.n_fft = L.mio_ref_n_fft(self.ctx)
        self.hop = L.mio_ref_hop_length(self.ctx)
        self.samples_per_token = L.mio_ref_samples_per_token(self.ctx)
        self.sample_rate = L.mio_ref_sample_rate(self.ctx)
        self.last_codec_sec = 0.0
        self.last_istft_sec = 0.0

    def close(self):
        if self.ctx:
            self.lib.mio_ref_free(self.ctx)
            self.ctx = None

    def decode_spec(self, codes: np.ndarray, emb: np.ndarray) -> np.ndarray:
        codes = np.ascontiguousarray(codes, dtype=np.int32)
        emb = np.ascontiguousarray(emb, dtype=np.float32)
        n_frames_guess = 18 * len(codes)
        out = np.empty(n_frames_guess * (self.n_fft + 2) + 16, dtype=np.float32)
        nf = C.c_int(0)
        sec = C.c_double(0)
        n = self.lib.mio_ref_decode(self.ctx, codes.ctypes.data, len(codes), emb.ctypes.data,
                                    out.ctypes.data, out.size, C.byref(nf), C.byref(sec))
        if n < 0:
            raise RuntimeError("reference decode failed (%d)" % n)
        self.last_codec_sec = sec.value
        return out[:n].reshape(nf.value, -1).copy()

    def istft(self, spec: np.ndarray) -> np.ndarray:
        spec = np.ascontiguousarray(spec, dtype=np.float32)
        nf = spec.shape[0]
        out = np.empty((nf + 4) * self.hop + self.n_fft, dtype=np.float32)
        sec = C.c_double(0)
        n = self.lib.mio_ref_istft(spec.ctypes.data, nf, self.n_fft, self.hop, out.ctypes.data, out.size,
                                   C.byref(sec))
        if n < 0:
            raise RuntimeError("reference istft failed")
        self.last_istft_sec = sec.value
        return out[:n].copy()

    def decode_wave(self, codes: np.ndarray, emb: np.ndarray) -> np.ndarray:
        codes = np.ascontiguousarray(codes, dtype=np.int32)
        emb = np.ascontiguousarray(emb, dtype=np.float32)
        out = np.empty(len(codes) * self.samples_per_token + 4096, dtype=np.float32)
        cs, is_ = C.c_double(0), C.c_double(0)
        n = self.lib.mio_ref_decode_wave(self.ctx, codes.ctypes.data, len(codes), emb.ctypes.data,
                                         out.ctypes.data, out.size, C.byref(cs), C.byref(is_))
        if n < 0:
            raise RuntimeError("reference decode_wave failed")
        self.last_codec_sec, self.last_istft_sec = cs.value, is_.value
        return out[:n].copy()

    def load_voice(self, path: str) -> np.ndarray:
        out = np.empty(4096, dtype=np.float32)
        n = self.lib.mio_ref_load_voice(path.encode(), out.ctypes.data, out.size)
        if n < 0:
            raise RuntimeError("reference load_voice_embedding failed")
        return out[:n].copy()


def ref_stream_session(gguf_path: str, voice_path: str, codes, skip_llm: bool = False, chunk_samples: int = 4096):
    """The reference's OWN streaming code: TestToSpeech::synthesize_stream_profiled (src/test-to-speech.cpp:348-626,
    unmodified) with oracle/stub_llama.cpp emitting `codes` one per generated token (LLM branch: maybe_emit cadence /
    holdback / llround indexing, emit_range chunking + crossfade) or, with skip_llm, the single-decode branch.
    Returns (emitted audio, chunk sizes, stats dict with the StreamProfile counters)."""
    L = C.CDLL(lib_path())
    L.mio_ref_stream_session.restype = C.c_long
    L.mio_ref_stream_session.argtypes = [C.c_char_p, C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_long, C.c_void_p, C.c_long,
                                         C.c_void_p, C.c_long, C.POINTER(C.c_long), C.c_void_p]
    codes = np.ascontiguousarray(codes, dtype=np.int32)
    out = np.empty(len(codes) * 1764 + 8192, dtype=np.float32)
    chunks = np.zeros(len(out) // max(1, chunk_samples) + 64 + 2 * len(codes), dtype=np.int64)
    nch = C.c_long(0)
    prof = np.zeros(5, dtype=np.int64)
    n = L.mio_ref_stream_session(gguf_path.encode(), voice_path.encode(), codes.ctypes.data, len(codes), int(skip_llm), chunk_samples,
                                 out.ctypes.data, out.size, chunks.ctypes.data, chunks.size, C.byref(nch), prof.ctypes.data)
    if n < 0:
        raise RuntimeError("reference stream session failed (%d)" % n)
    stats = dict(decode_calls=int(prof[0]), decoded_codes=int(prof[1]), emitted_samples=int(prof[2]), llm_tokens=int(prof[3]),
                 last_flags=int(prof[4]))
    return out[:n].copy(), chunks[: nch.value].copy(), stats


def ref_tts_skip_llm(gguf_path: str, voice_path: str, text: str, peak_normalize: bool = True, wav_path: str | None = None,
                     max_samples: int = 1764 * 4096):
    """The reference's TestToSpeech::synthesize_to_vector (and synthesize_to_file when wav_path is given) with skip_llm on
    raw token text.  Returns the samples, or None when the reference refuses the request."""
    L = C.CDLL(lib_path())
    L.mio_ref_tts_skip_llm.restype = C.c_long
    L.mio_ref_tts_skip_llm.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_void_p, C.c_long]
    out = np.empty(max_samples, dtype=np.float32)
    n = L.mio_ref_tts_skip_llm(gguf_path.encode(), voice_path.encode(), text.encode(), int(peak_normalize),
                               wav_path.encode() if wav_path else None, out.ctypes.data, out.size)
    if n == -1:
        return None
    if n < 0:
        raise RuntimeError(f"mio_ref_tts_skip_llm failed ({n})")
    return out[:n].copy()


def ref_istft_tables(n_fft: int = 392, win_length: int = 392) -> dict:
    """The tables of the reference's istft_cache (src/istft.cpp:7-41), through its own getters."""
    L = C.CDLL(lib_path())
    L.mio_ref_istft_tables.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    dims = np.zeros(4, dtype=np.int32)
    n_mid = max(0, n_fft // 2 + 1 - 2)
    cos_t = np.empty(n_fft * n_mid, dtype=np.float32)
    sin_t = np.empty_like(cos_t)
    nyq = np.empty(n_fft, dtype=np.float32)
    hann = np.empty(win_length, dtype=np.float32)
    L.mio_ref_istft_tables(n_fft, win_length, cos_t.ctypes.data, sin_t.ctypes.data, nyq.ctypes.data, hann.ctypes.data, dims.ctypes.data)
    return dict(dims=tuple(int(v) for v in dims), cos=cos_t, sin=sin_t, nyq=nyq, hann=hann)


def ref_istft_only(spec: np.ndarray, n_fft: int = 392, hop: int = 98) -> np.ndarray:
    """istft() of the reference without a codec context."""
    p = lib_path()
    if p is None:
        raise RuntimeError("oracle/_ref not built")
    L = C.CDLL(p)
    L.mio_ref_istft.restype = C.c_long
    L.mio_ref_istft.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_long, C.POINTER(C.c_double)]
    spec = np.ascontiguousarray(spec, dtype=np.float32)
    nf = spec.shape[0]
    out = np.empty((nf + 4) * hop + n_fft, dtype=np.float32)
    n = L.mio_ref_istft(spec.ctypes.data, nf, n_fft, hop, out.ctypes.data, out.size, None)
    if n < 0:
        return np.empty(0, dtype=np.float32)
    return out[:n].copy()


def ref_parse_speech_tokens(text) -> np.ndarray:
    """The reference's parse_speech_tokens (src/token-parser.cpp:5-28), compiled unmodified into oracle/_ref."""
    L = C.CDLL(lib_path())
    L.mio_ref_parse_speech_tokens.argtypes = [C.c_char_p, C.c_long, C.c_void_p, C.c_long]
    L.mio_ref_parse_speech_tokens.restype = C.c_long
    raw = text.encode("utf-8") if isinstance(text, str) else bytes(text)
    n = L.mio_ref_parse_speech_tokens(raw, len(raw), None, 0)
    out = np.empty(n, dtype=np.int32)
    if n:
        L.mio_ref_parse_speech_tokens(raw, len(raw), out.ctypes.data, n)
    return out


def ref_wav_write(path: str, samples: np.ndarray, sample_rate: int) -> None:
    """The reference's wav_write (src/wav-writer.cpp:24-45)."""
    L = C.CDLL(lib_path())
    L.mio_ref_wav_write.argtypes = [C.c_char_p, C.c_void_p, C.c_long, C.c_int]
    s = np.ascontiguousarray(samples, dtype=np.float32)
    if L.mio_ref_wav_write(path.encode(), s.ctypes.data, s.size, sample_rate) != 0:
        raise OSError("reference wav_write failed")
