"""ctypes binding of libmiocodec_b200.so (include/miocodec_b200.h) and a Python mirror of the reference's
operator interface for the hot path (same names and argument meaning as src/miocodec.h / src/istft.h of the
reference: ``miocodec_load``, ``miocodec_decode``, ``istft``, ``load_voice_embedding`` ...), so parity tests read
like calls into the reference.

There is no CPU fallback: importing works anywhere (so the symbol table can be checked), but creating a context
without a B200 raises ``MioError``.
"""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmiocodec_b200.so")

MIO_OK, MIO_ERR_ARG, MIO_ERR_IO, MIO_ERR_FORMAT, MIO_ERR_CUDA, MIO_ERR_RANGE, MIO_ERR_STATE = 0, -1, -2, -3, -4, -5, -6

# every symbol include/miocodec_b200.h declares (tests check the library exports exactly these)
EXPORTS = [
    "mio_ctx_create", "mio_istft_ctx_create", "mio_ctx_destroy", "mio_ctx_sample_rate", "mio_ctx_n_fft", "mio_ctx_hop_length",
    "mio_ctx_reserve", "mio_batch_pcm16_device", "mio_batch_pcm16_to_device", "mio_device_alloc", "mio_device_free",
    "mio_ipc_export", "mio_ipc_open", "mio_ipc_close", "mio_ctx_samples_per_token", "mio_ctx_frames_per_code", "mio_ctx_vocab", "mio_ctx_device", "mio_ctx_set_option",
    "mio_ctx_set_stream", "mio_voice_load", "mio_print_tensors", "mio_decode_spec", "mio_istft", "mio_decode_wave",
    "mio_decode_wave_batch", "mio_batch_upload", "mio_batch_run", "mio_batch_fetch", "mio_batch_sync",
    "mio_batch_kernel_launches", "mio_batch_wave_device", "mio_stream_default_config", "mio_stream_open",
    "mio_stream_close", "mio_stream_push_codes", "mio_stream_finish", "mio_streams_step", "mio_streams_step_begin", "mio_streams_step_end", "mio_stream_read",
    "mio_stream_stats", "mio_debug_read", "mio_last_error", "mio_version", "mio_host_alloc", "mio_host_free",
    "mio_profile_enable", "mio_profile_categories", "mio_profile_name", "mio_profile_read", "mio_batch_h2d_bytes",
    "mio_ctx_stream_seconds", "mio_parse_speech_tokens", "mio_decode_pcm16_batch", "mio_batch_fetch_pcm16", "mio_wav_header", "mio_wav_write_pcm16",
    "mio_debug_write", "mio_debug_clear", "mio_debug_gemm", "mio_peak_normalize_device_noalloc",
]


class MioError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"[{code}] {msg}")
        self.code = code


class StreamConfig(C.Structure):
    _fields_ = [("check_interval", C.c_int), ("holdback_codes", C.c_int), ("min_commit_codes", C.c_int),
                ("crossfade_samples", C.c_int), ("chunk_samples", C.c_int)]


_lib = None


def lib() -> C.CDLL:
    """Load the shared library (fails loudly if it has not been built)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MioError(MIO_ERR_IO, f"{LIB_PATH} not built: run miotts-llama.cpp_b200/build.sh (no CPU fallback exists)")
    L = C.CDLL(LIB_PATH)
    vp, i32p, f32p, i64 = C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64
    L.mio_last_error.restype = C.c_char_p
    L.mio_version.restype = C.c_char_p
    L.mio_ctx_create.argtypes = [C.c_char_p, C.c_int, C.POINTER(vp)]
    L.mio_istft_ctx_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.mio_ctx_destroy.argtypes = [vp]
    L.mio_ctx_destroy.restype = None
    for n in ("sample_rate", "n_fft", "hop_length", "samples_per_token", "frames_per_code", "vocab", "device"):
        getattr(L, "mio_ctx_" + n).argtypes = [vp]
    L.mio_ctx_set_option.argtypes = [vp, C.c_char_p, C.c_char_p]
    L.mio_ctx_set_stream.argtypes = [vp, vp]
    L.mio_voice_load.argtypes = [C.c_char_p, f32p, C.c_int]
    L.mio_print_tensors.argtypes = [C.c_char_p]
    L.mio_decode_spec.argtypes = [vp, i32p, C.c_int, f32p, f32p, i64, C.POINTER(C.c_int)]
    L.mio_istft.argtypes = [vp, f32p, C.c_int, f32p, i64, C.POINTER(i64)]
    L.mio_decode_wave.argtypes = [vp, i32p, C.c_int, f32p, C.c_int, f32p, i64, C.POINTER(i64)]
    L.mio_decode_wave_batch.argtypes = [vp, i32p, i32p, C.c_int, f32p, C.c_int, i32p, C.c_int, f32p, i64]
    L.mio_batch_upload.argtypes = [vp, i32p, i32p, C.c_int, f32p, C.c_int, i32p]
    L.mio_batch_run.argtypes = [vp, C.c_int]
    L.mio_batch_fetch.argtypes = [vp, f32p, i64]
    L.mio_batch_sync.argtypes = [vp]
    L.mio_batch_kernel_launches.argtypes = [vp]
    L.mio_batch_kernel_launches.restype = i64
    L.mio_batch_wave_device.argtypes = [vp]
    L.mio_batch_wave_device.restype = vp
    L.mio_ctx_reserve.argtypes = [vp, C.c_int64, C.c_int, C.c_int64]
    L.mio_batch_pcm16_device.argtypes = [vp]
    L.mio_batch_pcm16_device.restype = vp
    L.mio_batch_pcm16_to_device.argtypes = [vp, vp, C.c_int64]
    L.mio_device_alloc.argtypes = [C.c_int, C.c_size_t]
    L.mio_device_alloc.restype = vp
    L.mio_device_free.argtypes = [vp]
    L.mio_device_free.restype = None
    L.mio_ipc_export.argtypes = [vp, C.c_char_p]
    L.mio_ipc_open.argtypes = [C.c_char_p, C.c_int]
    L.mio_ipc_open.restype = vp
    L.mio_ipc_close.argtypes = [vp]
    L.mio_stream_default_config.argtypes = [vp, C.POINTER(StreamConfig)]
    L.mio_stream_default_config.restype = None
    L.mio_stream_open.argtypes = [vp, f32p, C.POINTER(StreamConfig), C.POINTER(vp)]
    L.mio_stream_close.argtypes = [vp]
    L.mio_stream_close.restype = None
    L.mio_stream_push_codes.argtypes = [vp, i32p, C.c_int]
    L.mio_stream_finish.argtypes = [vp]
    L.mio_streams_step.argtypes = [vp, C.POINTER(vp), C.c_int]
    L.mio_streams_step_begin.argtypes = [vp, C.POINTER(vp), C.c_int]
    L.mio_streams_step_end.argtypes = [vp, C.POINTER(vp), C.c_int]
    L.mio_stream_read.argtypes = [vp, f32p, i64, C.POINTER(C.c_int)]
    L.mio_stream_read.restype = i64
    L.mio_stream_stats.argtypes = [vp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]
    L.mio_host_alloc.argtypes = [C.c_size_t]
    L.mio_host_alloc.restype = vp
    L.mio_host_free.argtypes = [vp]
    L.mio_host_free.restype = None
    L.mio_profile_enable.argtypes = [vp, C.c_int]
    L.mio_profile_name.argtypes = [C.c_int]
    L.mio_profile_name.restype = C.c_char_p
    L.mio_profile_read.argtypes = [vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(i64)]
    L.mio_batch_h2d_bytes.argtypes = [vp]
    L.mio_batch_h2d_bytes.restype = i64
    L.mio_ctx_stream_seconds.argtypes = [vp, C.POINTER(i64)]
    L.mio_ctx_stream_seconds.restype = C.c_double
    L.mio_parse_speech_tokens.argtypes = [C.c_char_p, i64, i32p, i64]
    L.mio_parse_speech_tokens.restype = i64
    L.mio_decode_pcm16_batch.argtypes = [vp, i32p, i32p, C.c_int, f32p, C.c_int, i32p, C.c_int, vp, i64]
    L.mio_batch_fetch_pcm16.argtypes = [vp, vp, i64]
    L.mio_wav_header.argtypes = [C.c_int, i64, vp]
    L.mio_wav_write_pcm16.argtypes = [C.c_char_p, vp, i64, C.c_int]
    L.mio_debug_read.argtypes = [vp, C.c_char_p, f32p, i64]
    L.mio_debug_read.restype = i64
    L.mio_debug_write.argtypes = [vp, C.c_char_p, f32p, i64]
    L.mio_debug_clear.argtypes = [vp]
    L.mio_debug_gemm.argtypes = [vp, C.c_int, C.c_int, C.c_int, f32p, f32p, C.c_int, f32p]
    L.mio_peak_normalize_device_noalloc.argtypes = [C.c_int, vp, vp, vp, C.c_int, vp]
    _lib = L
    return L


def peak_normalize_device_noalloc(device: int, wave_in_ptr: int, wave_out_ptr: int, offsets_ptr: int, n_utts: int,
                                  cuda_stream_ptr: int = 0) -> None:
    """mio_peak_normalize_device_noalloc: per-utterance 0.95/max|s| on a waveform already in device memory, for callers that
    may not allocate (no context, no cudaMalloc; at most MIO_NOALLOC_MAX_UTTS = 4096 utterances; calls on one GPU must not
    overlap).  All three pointers are device addresses; offsets is int64[n_utts + 1] in samples.  Only enqueues work."""
    _check(lib().mio_peak_normalize_device_noalloc(device, C.c_void_p(wave_in_ptr), C.c_void_p(wave_out_ptr),
                                                   C.c_void_p(offsets_ptr), n_utts, C.c_void_p(cuda_stream_ptr)))


def device_alloc(device: int, nbytes: int) -> int:
    p = lib().mio_device_alloc(device, nbytes)
    if not p:
        raise MioError(MIO_ERR_CUDA, lib().mio_last_error().decode("utf-8", "replace"))
    return int(p)


def device_free(ptr: int):
    lib().mio_device_free(C.c_void_p(ptr))


def ipc_export(ptr: int) -> bytes:
    buf = C.create_string_buffer(64)
    _check(lib().mio_ipc_export(C.c_void_p(ptr), buf))
    return buf.raw


def ipc_open(handle: bytes, device: int) -> int:
    p = lib().mio_ipc_open(handle, device)
    if not p:
        raise MioError(MIO_ERR_CUDA, lib().mio_last_error().decode("utf-8", "replace"))
    return int(p)


def ipc_close(ptr: int):
    _check(lib().mio_ipc_close(C.c_void_p(ptr)))


def pinned_empty(n: int, dtype) -> np.ndarray:
    """numpy array over cudaMallocHost memory (freed when the array is collected)."""
    dt = np.dtype(dtype)
    ptr = lib().mio_host_alloc(max(1, n * dt.itemsize))
    if not ptr:
        raise MioError(MIO_ERR_CUDA, "cudaMallocHost failed")
    buf = (C.c_char * (n * dt.itemsize)).from_address(ptr)
    arr = np.frombuffer(buf, dtype=dt, count=n)
    import weakref
    weakref.finalize(buf, lib().mio_host_free, C.c_void_p(ptr))
    return arr


def _check(rc: int):
    if rc < 0:
        raise MioError(rc, lib().mio_last_error().decode("utf-8", "replace"))
    return rc


def _i32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.int32)


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def parse_speech_tokens(text) -> np.ndarray:
    """parse_speech_tokens (src/token-parser.h:8): the integer codes of every "<|s_N|>" in the text."""
    raw = text.encode("utf-8") if isinstance(text, str) else bytes(text)
    n = lib().mio_parse_speech_tokens(raw, len(raw), None, 0)
    _check(n)
    out = np.empty(n, dtype=np.int32)
    if n:
        _check(lib().mio_parse_speech_tokens(raw, len(raw), out.ctypes.data, n))
    return out


def wav_header(sample_rate: int, n_samples: int) -> bytes:
    h = (C.c_uint8 * 44)()
    _check(lib().mio_wav_header(sample_rate, n_samples, h))
    return bytes(h)


def wav_write_pcm16(path: str, pcm, sample_rate: int):
    pcm = np.ascontiguousarray(pcm, dtype=np.int16)
    _check(lib().mio_wav_write_pcm16(path.encode(), pcm.ctypes.data, pcm.size, sample_rate))


class MioCodec:
    """One decoder context on one GPU (the reference's ``miocodec_context``)."""

    def __init__(self, gguf_path: str, device: int = 0, gemm: str | None = None, max_chunk_codes: int | None = None):
        L = lib()
        h = C.c_void_p()
        _check(L.mio_ctx_create(gguf_path.encode(), device, C.byref(h)))
        self._h = h
        self.sample_rate = L.mio_ctx_sample_rate(h)
        self.n_fft = L.mio_ctx_n_fft(h)
        self.hop_length = L.mio_ctx_hop_length(h)
        self.samples_per_token = L.mio_ctx_samples_per_token(h)
        self.frames_per_code = L.mio_ctx_frames_per_code(h)
        self.vocab = L.mio_ctx_vocab(h)
        self.n_freq = self.n_fft // 2 + 1
        if gemm:
            self.set_option("gemm", gemm)
        if max_chunk_codes:
            self.set_option("max_chunk_codes", str(max_chunk_codes))

    def close(self):
        if getattr(self, "_h", None):
            lib().mio_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, key: str, value: str):
        _check(lib().mio_ctx_set_option(self._h, key.encode(), str(value).encode()))

    def reserve(self, max_batch_codes: int, max_utts: int, max_emit_samples: int = 0):
        """Pre-size device / pinned buffers (mio_ctx_reserve)."""
        _check(lib().mio_ctx_reserve(self._h, max_batch_codes, max_utts, max_emit_samples))

    def set_stream(self, cuda_stream_ptr: int):
        """Run this context's kernels on an existing CUDA stream (e.g. torch.cuda.current_stream().cuda_stream).  Until this
        is called the context works on a private non-blocking stream, which NOTHING else is ordered against."""
        _check(lib().mio_ctx_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))
        self.stream_ptr = int(cuda_stream_ptr or 0)

    #: the stream given to set_stream, or None while the context still runs on its private stream
    stream_ptr = None

    # ---- reference-shaped calls ------------------------------------------------------------
    def decode(self, codes, global_emb) -> np.ndarray:
        """miocodec_decode (src/miocodec.h:26-32): returns spec [n_frames, n_freq*2] interleaved (re, im)."""
        codes, emb = _i32(codes), _f32(global_emb)
        nf = len(codes) * self.frames_per_code
        out = np.empty((nf, 2 * self.n_freq), dtype=np.float32)
        n_frames = C.c_int(0)
        _check(lib().mio_decode_spec(self._h, codes.ctypes.data, len(codes), emb.ctypes.data, out.ctypes.data, out.size,
                                     C.byref(n_frames)))
        assert n_frames.value == nf
        return out

    def istft(self, spec) -> np.ndarray:
        """istft (src/istft.h:38-42) with the context's n_fft/hop and win_length == n_fft."""
        spec = _f32(spec)
        nf = spec.shape[0]
        out = np.empty(nf * self.hop_length, dtype=np.float32)
        n = C.c_int64(0)
        _check(lib().mio_istft(self._h, spec.ctypes.data, nf, out.ctypes.data, out.size, C.byref(n)))
        return out[: n.value]

    def decode_wave(self, codes, global_emb, peak_normalize: bool = False) -> np.ndarray:
        """decode_codes_to_audio (src/test-to-speech.cpp:248-303)."""
        codes, emb = _i32(codes), _f32(global_emb)
        out = np.empty(len(codes) * self.samples_per_token, dtype=np.float32)
        n = C.c_int64(0)
        _check(lib().mio_decode_wave(self._h, codes.ctypes.data, len(codes), emb.ctypes.data, int(peak_normalize),
                                     out.ctypes.data, out.size, C.byref(n)))
        return out[: n.value]

    # ---- batch ---------------------------------------------------------------------------------
    def batch_upload(self, codes_concat, cu_seqlens, voice_table, voice_ids):
        self._b = (_i32(codes_concat), _i32(cu_seqlens), _f32(voice_table), _i32(voice_ids))
        c, cu, vt, vi = self._b
        _check(lib().mio_batch_upload(self._h, c.ctypes.data, cu.ctypes.data, len(cu) - 1, vt.ctypes.data, vt.shape[0],
                                      vi.ctypes.data))
        self._total = int(cu[-1]) * self.samples_per_token

    def batch_run(self, peak_normalize: bool = False):
        _check(lib().mio_batch_run(self._h, int(peak_normalize)))

    def batch_sync(self):
        _check(lib().mio_batch_sync(self._h))

    def batch_fetch(self, out: np.ndarray | None = None) -> np.ndarray:
        if out is None:
            out = np.empty(self._total, dtype=np.float32)
        _check(lib().mio_batch_fetch(self._h, out.ctypes.data, out.size))
        return out

    def wave_device(self) -> tuple[int, int]:
        """(device pointer, samples) of the resident float waveform of the last batch."""
        return int(lib().mio_batch_wave_device(self._h) or 0), self._total

    def pcm16_device(self) -> tuple[int, int]:
        """(device pointer, samples) of the PCM16 image of the resident waveform, converted ASYNCHRONOUSLY on the context's
        stream: a consumer on another stream (torch ops, a torch.distributed collective) must be ordered behind it -- make
        the streams the same (set_stream), or batch_sync() first; sharding.order_after_codec does whichever is needed."""
        p = lib().mio_batch_pcm16_device(self._h)
        if not p:
            raise MioError(-3, lib().mio_last_error().decode("utf-8", "replace"))
        return int(p), self._total

    def pcm16_to_device(self, dst_ptr: int, cap: int) -> int:
        """PCM16 of the resident waveform stored into device memory at dst_ptr (may be a peer GPU's); returns the samples.
        Asynchronous on the context's stream, like pcm16_device: see the ordering note there."""
        _check(lib().mio_batch_pcm16_to_device(self._h, C.c_void_p(dst_ptr), cap))
        return self._total

    def kernel_launches(self) -> int:
        return int(lib().mio_batch_kernel_launches(self._h))

    def decode_wave_batch(self, codes_concat, cu_seqlens, voice_table, voice_ids, peak_normalize=False, out=None):
        c, cu, vt, vi = _i32(codes_concat), _i32(cu_seqlens), _f32(voice_table), _i32(voice_ids)
        total = int(cu[-1]) * self.samples_per_token
        if out is None:
            out = np.empty(total, dtype=np.float32)
        _check(lib().mio_decode_wave_batch(self._h, c.ctypes.data, cu.ctypes.data, len(cu) - 1, vt.ctypes.data, vt.shape[0],
                                           vi.ctypes.data, int(peak_normalize), out.ctypes.data, out.size))
        return out

    def decode_pcm16_batch(self, codes_concat, cu_seqlens, voice_table, voice_ids, peak_normalize=False, out=None):
        """decode_codes_to_audio + the PCM16 conversion of wav_write (src/wav-writer.cpp:39-42), converted on the device."""
        c, cu, vt, vi = _i32(codes_concat), _i32(cu_seqlens), _f32(voice_table), _i32(voice_ids)
        total = int(cu[-1]) * self.samples_per_token
        if out is None:
            out = np.empty(total, dtype=np.int16)
        _check(lib().mio_decode_pcm16_batch(self._h, c.ctypes.data, cu.ctypes.data, len(cu) - 1, vt.ctypes.data, vt.shape[0],
                                            vi.ctypes.data, int(peak_normalize), out.ctypes.data, out.size))
        return out

    def batch_fetch_pcm16(self, out: np.ndarray | None = None) -> np.ndarray:
        if out is None:
            out = np.empty(self._total, dtype=np.int16)
        _check(lib().mio_batch_fetch_pcm16(self._h, out.ctypes.data, out.size))
        return out

    def profile_enable(self, on: bool = True):
        _check(lib().mio_profile_enable(self._h, int(on)))

    def profile_read(self) -> dict:
        out = {}
        for cat in range(lib().mio_profile_categories()):
            ms, fl, by, n = C.c_double(0), C.c_double(0), C.c_double(0), C.c_int64(0)
            _check(lib().mio_profile_read(self._h, cat, C.byref(ms), C.byref(fl), C.byref(by), C.byref(n)))
            out[lib().mio_profile_name(cat).decode()] = dict(ms=ms.value, flops=fl.value, bytes=by.value, launches=n.value)
        return out

    def h2d_bytes(self) -> int:
        return int(lib().mio_batch_h2d_bytes(self._h))

    def debug_read(self, name: str) -> np.ndarray:
        n = lib().mio_debug_read(self._h, name.encode(), None, 0)
        _check(n)
        out = np.empty(n, dtype=np.float32)
        _check(lib().mio_debug_read(self._h, name.encode(), out.ctypes.data, out.size))
        return out

    def debug_write(self, name: str, data):
        """Teacher forcing: replace the leading floats of the named tap on the next decodes (mio_debug_write)."""
        d = _f32(data).reshape(-1)
        _check(lib().mio_debug_write(self._h, name.encode(), d.ctypes.data, d.size))

    def debug_clear(self):
        _check(lib().mio_debug_clear(self._h))

    def debug_gemm(self, A, B, npass: int = 3) -> np.ndarray:
        """out[M][N] = A[M][K] @ B[N][K].T through the product tcgen05 kernel (mio_debug_gemm)."""
        A, B = _f32(A), _f32(B)
        M, K = A.shape
        N = B.shape[0]
        assert B.shape[1] == K
        out = np.empty((M, N), dtype=np.float32)
        _check(lib().mio_debug_gemm(self._h, M, N, K, A.ctypes.data, B.ctypes.data, npass, out.ctypes.data))
        return out

    # ---- streaming -------------------------------------------------------------------------------
    def default_stream_config(self) -> StreamConfig:
        cfg = StreamConfig()
        lib().mio_stream_default_config(self._h, C.byref(cfg))
        return cfg

    def open_stream(self, global_emb, cfg: StreamConfig | None = None) -> "Stream":
        return Stream(self, global_emb, cfg)

    def stream_seconds(self):
        """(seconds, passes) spent in step_streams so far: StreamProfile.codec_sec + istft_sec of the reference, fused."""
        n = C.c_int64(0)
        return float(lib().mio_ctx_stream_seconds(self._h, C.byref(n))), int(n.value)

    def step_streams(self, streams):
        arr = (C.c_void_p * len(streams))(*[s._h for s in streams])
        _check(lib().mio_streams_step(self._h, arr, len(streams)))

    def step_streams_begin(self, streams):
        """Enqueue the round and return while the GPU works (mio_streams_step_begin); pair with step_streams_end."""
        arr = (C.c_void_p * len(streams))(*[s._h for s in streams])
        return _check(lib().mio_streams_step_begin(self._h, arr, len(streams))) == 1      # True: a round is in flight

    def step_streams_end(self, streams):
        arr = (C.c_void_p * len(streams))(*[s._h for s in streams])
        _check(lib().mio_streams_step_end(self._h, arr, len(streams)))


class Stream:
    def __init__(self, codec: MioCodec, global_emb, cfg: StreamConfig | None):
        self.codec = codec
        emb = _f32(global_emb)
        h = C.c_void_p()
        _check(lib().mio_stream_open(codec._h, emb.ctypes.data, C.byref(cfg) if cfg is not None else None, C.byref(h)))
        self._h = h

    def push(self, codes):
        c = _i32(np.atleast_1d(codes))
        _check(lib().mio_stream_push_codes(self._h, c.ctypes.data, len(c)))

    def finish(self):
        _check(lib().mio_stream_finish(self._h))

    def read(self, max_samples: int = 1 << 22):
        out = np.empty(max_samples, dtype=np.float32)
        last = C.c_int(0)
        n = lib().mio_stream_read(self._h, out.ctypes.data, out.size, C.byref(last))
        _check(n)
        return out[:n].copy(), bool(last.value)

    def stats(self):
        a, b, c = C.c_int64(0), C.c_int64(0), C.c_int64(0)
        _check(lib().mio_stream_stats(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return dict(decode_calls=a.value, decoded_codes=b.value, emitted_samples=c.value)

    def close(self):
        if getattr(self, "_h", None):
            lib().mio_stream_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- free functions with the reference's names (src/miocodec.h, src/istft.h) ---------------------
def miocodec_load(model_path: str, device: int = 0) -> MioCodec | None:
    """Returns None on failure like the reference (nullptr); the error text is in mio_last_error()."""
    try:
        return MioCodec(model_path, device)
    except MioError:
        return None


def miocodec_free(ctx: MioCodec | None):
    if ctx is not None:
        ctx.close()


def miocodec_decode(ctx: MioCodec, codes, n_codes: int, global_emb, audio_length: int = 0):
    """Returns (spec, n_frames). `audio_length` is ignored, as in the reference (src/miocodec.cpp:546-552)."""
    spec = ctx.decode(np.asarray(codes)[:n_codes], global_emb)
    return spec, spec.shape[0]


def istft(ctx: MioCodec, spec, n_frames: int, hop_length: int | None = None) -> np.ndarray:
    return ctx.istft(np.asarray(spec, dtype=np.float32).reshape(n_frames, -1))


def load_voice_embedding(path: str) -> np.ndarray:
    """Empty array on failure like the reference (src/miocodec.cpp:816-853)."""
    out = np.empty(4096, dtype=np.float32)
    n = lib().mio_voice_load(path.encode(), out.ctypes.data, out.size)
    if n < 0:
        return np.empty(0, dtype=np.float32)
    return out[:n].copy()


def miocodec_print_tensors(path: str):
    lib().mio_print_tensors(path.encode())
