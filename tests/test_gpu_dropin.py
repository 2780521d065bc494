"""The drop-in proof (VERDICT r1 "what's missing" item 2): the reference's OWN executables — src/main.cpp (miotts),
examples/stream-compare.cpp, examples/stream-benchmark.cpp, with the reference's own src/test-to-speech.cpp,
token-parser.cpp, wav-writer.cpp and text-normalize.cpp, all unmodified and compiled against the reference's headers —
linked against libmiocodec_shim.so + libmiocodec_b200.so INSTEAD of src/miocodec.cpp + src/istft.cpp + ggml
(oracle/build_dropin.sh; libllama replaced by the scripted oracle/stub_llama.cpp), run on the B200 and compared with the
same executables linked the stock way (oracle/_ref/bin, the reference's CPU codec)."""
import os
import re
import subprocess
import wave

import numpy as np
import pytest

from conftest import ROOT, snr_db

pytestmark = pytest.mark.gpu

OURS = os.path.join(ROOT, "miotts-llama.cpp_b200", "miocodec_b200", "dropin")
STOCK = os.path.join(ROOT, "oracle", "_ref", "bin")
NAMES = ("miotts", "miotts-stream-compare", "miotts-stream-benchmark")


def _need_binaries():
    for d in (OURS, STOCK):
        for n in NAMES:
            if not os.path.exists(os.path.join(d, n)):
                pytest.skip(f"{os.path.join(d, n)} not built (oracle/build_dropin.sh needs /root/reference)")


def _run(args, **kw):
    r = subprocess.run(args, capture_output=True, text=True, timeout=600, **kw)
    assert r.returncode == 0, (args[0], r.stdout[-2000:], r.stderr[-2000:])
    return r.stdout


def _pcm(path):
    with wave.open(path, "rb") as w:
        assert (w.getnchannels(), w.getsampwidth(), w.getframerate()) == (1, 2, 44100)
        return np.frombuffer(w.readframes(w.getnframes()), dtype=np.int16)


def _prompt(codes):
    return "".join(f"<|s_{int(c)}|>" for c in codes)


def test_miotts_skip_llm_wav_matches_the_stock_binary(assets, tmp_path):
    """miotts --skip-llm (src/main.cpp:91-157 -> synthesize_to_file -> decode_tokens_to_audio -> miocodec_decode + istft ->
    peak normalise -> wav_write), ours on the GPU vs the stock link on the host CPU: same header, same sample count,
    PCM at the parity floor."""
    _need_binaries()
    from miocodec_b200 import synth
    codes = synth.synth_codes(60, 9)
    outs = {}
    for tag, d in (("ours", OURS), ("stock", STOCK)):
        out = str(tmp_path / f"{tag}.wav")
        _run([os.path.join(d, "miotts"), "--skip-llm", "-c", assets[0], "-v", assets[1][3], "-p", _prompt(codes), "-o", out])
        outs[tag] = out
    a, b = _pcm(outs["ours"]), _pcm(outs["stock"])
    assert a.shape == b.shape == (60 * 1764,)
    assert open(outs["ours"], "rb").read(44) == open(outs["stock"], "rb").read(44)      # RIFF header byte-identical
    s = snr_db(a.astype(np.float64), b.astype(np.float64))
    print(f"miotts --skip-llm, 60 codes: PCM16 SNR ours vs stock {s:.1f} dB, max |diff| {np.abs(a.astype(int) - b.astype(int)).max()} LSB")
    assert s > 58.0
    assert abs(int(np.abs(a).max()) - int(np.abs(b).max())) <= 1 and 31100 <= int(np.abs(a).max()) <= 31130   # 0.95 * 32767


def test_dump_tensors_runs_through_the_shim(assets):
    _need_binaries()
    for d in (OURS, STOCK):
        out = _run([os.path.join(d, "miotts"), "--dump-tensors", "-c", assets[0]])
        assert "token_embd" in out and "istft_head.out.weight" in out


def test_stream_benchmark_llm_branch_counters_match(assets, tmp_path):
    """miotts-stream-benchmark with the scripted LLM (holdback / cadence / re-decode policy of the reference's own
    synthesize_stream_profiled, the codec underneath is ours): every StreamProfile counter equals the stock binary's."""
    _need_binaries()
    from miocodec_b200 import synth
    script = tmp_path / "script.txt"
    script.write_text("\n".join(str(int(c)) for c in synth.synth_codes(130, 8)))
    got = {}
    for tag, d in (("ours", OURS), ("stock", STOCK)):
        out = _run([os.path.join(d, "miotts-stream-benchmark"), "-m", str(script), "-c", assets[0], "-v", assets[1][0], "-p", "hello"])
        got[tag] = {k: v for k, v in re.findall(r"stream_bench\.(llm_tokens|decode_calls|decoded_codes|emitted_samples)=(\d+)", out)}
        print(tag, got[tag], re.findall(r"stream_bench\.(x_realtime|stage\.codec_sec|stage\.istft_sec)=([0-9.]+)", out))
    assert got["ours"] == got["stock"] == {"llm_tokens": "130", "decode_calls": "3", "decoded_codes": "290", "emitted_samples": str(130 * 1764)}


def test_stream_compare_offline_equals_stream(assets, tmp_path):
    """miotts-stream-compare --skip-llm (examples/stream-compare.cpp:166-278): offline vs streamed synthesis through the
    shim are the same samples (sample_diff 0, max_abs 0), and the offline WAV is at the parity floor of the stock one."""
    _need_binaries()
    from miocodec_b200 import synth
    codes = synth.synth_codes(45, 4)
    wavs = {}
    for tag, d in (("ours", OURS), ("stock", STOCK)):
        off, st = str(tmp_path / f"{tag}_off.wav"), str(tmp_path / f"{tag}_str.wav")
        out = _run([os.path.join(d, "miotts-stream-compare"), "--skip-llm", "-c", assets[0], "-v", assets[1][1], "-p", _prompt(codes),
                    "--out-offline", off, "--out-stream", st])
        assert "sample_diff=0" in out and "best_lag_samples=0" in out, out
        assert re.search(r"compare\.max_abs=0\.0+\b", out), out
        wavs[tag] = (_pcm(off), _pcm(st))
        assert np.array_equal(*wavs[tag])
    s = snr_db(wavs["ours"][0].astype(np.float64), wavs["stock"][0].astype(np.float64))
    print(f"stream-compare offline WAV, 45 codes: PCM16 SNR ours vs stock {s:.1f} dB")
    assert s > 57.5
