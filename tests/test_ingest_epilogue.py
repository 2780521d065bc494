"""The steps either side of the hot path, checked against the UNMODIFIED reference
functions compiled into oracle/_ref: parse_speech_tokens (src/token-parser.cpp:5-28) and wav_write
(src/wav-writer.cpp:24-45).  Integer / byte work: the bar is bit-exact."""
import os

import numpy as np
import pytest

from conftest import ROOT

# (text, expected codes): expectations follow src/token-parser.cpp:9-25 and were confirmed against the compiled reference
PARSER_CASES = [
    ("", []),
    ("<|s_5|>", [5]),                                     # "|>" may end exactly at the end of the text
    ("<|s_1|><|s_22|><|s_12799|>", [1, 22, 12799]),
    ("hello <|s_7|> world <|s_8|>!", [7, 8]),
    ("<|s_|><|s_3|>", [3]),                               # no digits: skipped
    ("<|s_12", []),                                       # unterminated
    ("<|s_12|", []),                                      # "|" but no ">"
    ("<|s_12|x<|s_4|>", [4]),
    ("<|s_-3|>", [-3]),                                   # strtol takes a sign
    ("<|s_ 9|>", [9]),                                    # ... and leading whitespace
    ("<|s_+11|>", [11]),
    ("<|s_007|>", [7]),
    ("<|s_1.5|>", []),                                    # strtol stops at '.', which is not '|'
    ("<|s_<|s_6|>", [6]),                                 # restart one byte after a failed '<'
    ("<|s_99999999999999999999|>", [-1]),                 # strtol saturates at LONG_MAX, int() keeps the low 32 bits
    ("<|s_4294967297|>", [1]),                            # 2^32 + 1 truncated to int
    ("<|S_5|><|s_6|>", [6]),                              # case sensitive
    ("こんにちは<|s_42|>。", [42]),   # multi-byte text around a token
    ("<|s_1|>\0<|s_2|>", [1]),                            # the scan stops at an embedded NUL (strstr on c_str())
]


def _ref():
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    return ref


@pytest.mark.parametrize("text,expected", PARSER_CASES)
def test_parse_speech_tokens_known_answers(text, expected):
    from miocodec_b200 import binding
    got = binding.parse_speech_tokens(text)
    assert got.dtype == np.int32 and got.tolist() == expected


def test_parse_speech_tokens_matches_compiled_reference():
    from miocodec_b200 import binding
    ref = _ref()
    for text, expected in PARSER_CASES:
        assert ref.ref_parse_speech_tokens(text).tolist() == expected, text
    # random soups of the pieces the grammar is made of
    rng = np.random.default_rng(7)
    pieces = ["<|s_", "|>", "<", "|", ">", "s_", "12", "7", "-", " ", "+", "x", "\n", "0", "98765", "<|s_3|>", "あ"]
    for _ in range(300):
        text = "".join(pieces[i] for i in rng.integers(0, len(pieces), size=int(rng.integers(0, 40))))
        assert binding.parse_speech_tokens(text).tolist() == ref.ref_parse_speech_tokens(text).tolist(), repr(text)
    # a long generated sequence, the shape the LLM actually emits
    codes = rng.integers(0, 12800, size=5000)
    text = "".join(f"<|s_{c}|>" for c in codes)
    assert binding.parse_speech_tokens(text).tolist() == codes.tolist()
    assert ref.ref_parse_speech_tokens(text).tolist() == codes.tolist()


def test_parse_speech_tokens_capacity_and_errors():
    from miocodec_b200 import binding
    L = binding.lib()
    raw = b"<|s_1|><|s_2|><|s_3|>"
    out = np.full(2, -7, dtype=np.int32)
    assert L.mio_parse_speech_tokens(raw, len(raw), out.ctypes.data, 2) == 3      # count may exceed cap
    assert out.tolist() == [1, 2]
    assert L.mio_parse_speech_tokens(None, 0, None, 0) == binding.MIO_ERR_ARG
    assert L.mio_parse_speech_tokens(raw, -1, None, 0) == binding.MIO_ERR_ARG


def _pcm16_numpy(x):
    """int16(clamp(s * 32767.0f, -32768.0f, 32767.0f)) in float32 arithmetic, truncation (src/wav-writer.cpp:39-42)."""
    y = np.clip(x.astype(np.float32) * np.float32(32767.0), np.float32(-32768.0), np.float32(32767.0))
    return np.trunc(y).astype(np.int16)


def test_wav_header_and_file_match_reference_writer(tmp_path):
    from miocodec_b200 import binding
    ref = _ref()
    rng = np.random.default_rng(11)
    for n, sr in [(0, 44100), (1, 44100), (1764 * 20, 44100), (12345, 24000)]:
        x = (rng.standard_normal(n) * 0.7).astype(np.float32)        # includes |s| > 1 (clamped)
        if n > 4:
            x[:4] = [1.0, -1.0, 1.00002, -1.00004]
        a, b = tmp_path / f"ref_{n}.wav", tmp_path / f"ours_{n}.wav"
        ref.ref_wav_write(str(a), x, sr)
        binding.wav_write_pcm16(str(b), _pcm16_numpy(x), sr)
        ra, rb = a.read_bytes(), b.read_bytes()
        assert ra[:44] == binding.wav_header(sr, n)
        assert ra == rb


@pytest.mark.gpu
def test_pcm16_batch_is_the_reference_conversion_of_the_float_batch(codec, assets, tmp_path):
    from miocodec_b200 import binding
    rng = np.random.default_rng(5)
    lens = [33, 1, 7, 120, 64]                                       # odd sample offsets: unaligned chunk starts
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    codes = rng.integers(0, codec.vocab, size=int(cu[-1])).astype(np.int32)
    voices = np.stack([binding.load_voice_embedding(assets[1][v]) for v in range(3)])
    vids = np.arange(len(lens), dtype=np.int32) % 3
    for peak in (False, True):
        wave = codec.decode_wave_batch(codes, cu, voices, vids, peak_normalize=peak)
        pcm = codec.decode_pcm16_batch(codes, cu, voices, vids, peak_normalize=peak)
        assert pcm.dtype == np.int16 and pcm.shape == wave.shape
        assert np.array_equal(pcm, _pcm16_numpy(wave))
        # the three-call form delivers the same bytes
        codec.batch_upload(codes, cu, voices, vids)
        codec.batch_run(peak)
        assert np.array_equal(codec.batch_fetch_pcm16(), pcm)
    # small chunks: every chunk boundary lands on an utterance boundary, not on an 8-sample boundary
    codec.set_option("fetch_chunk_codes", "64")
    try:
        wave_s = codec.decode_wave_batch(codes, cu, voices, vids, peak_normalize=True)
        pcm_s = codec.decode_pcm16_batch(codes, cu, voices, vids, peak_normalize=True)
        bad = np.flatnonzero(pcm_s != _pcm16_numpy(wave_s))
        assert bad.size == 0, (bad[:8], pcm_s[bad[:8]], _pcm16_numpy(wave_s)[bad[:8]])
        # another chunking regroups the fp32 partial sums of the GroupNorm statistics: same tolerance as
        # test_gpu_parity.test_multi_chunk_batches
        from conftest import snr_db
        assert snr_db(wave_s, wave) > 55.0
    finally:
        codec.set_option("fetch_chunk_codes", str(1 << 20))
    from oracle import ref
    if ref.available():                                              # file == the reference writer's file for the same floats
        a, b = tmp_path / "ref.wav", tmp_path / "ours.wav"
        ref.ref_wav_write(str(a), wave, codec.sample_rate)
        binding.wav_write_pcm16(str(b), pcm, codec.sample_rate)
        assert a.read_bytes() == b.read_bytes()
