"""BASELINE config C4 at full scale: 1024 concurrent streams, each a 700-code sequence revealed one code per tick with
random phase offsets, decoded with the reference's streaming policy (commit holdback 32, min step 24, cadence 20,
full-prefix re-decode, llround indexing, 1323-sample crossfade; src/test-to-speech.cpp:364-417, :499-571).

Checked for every stream: emitted sample count == 1764 * codes (stream-compare's sample_diff == 0,
examples/stream-compare.cpp:258-262) and the StreamProfile counters (decode_calls, decoded_codes, emitted_samples,
src/test-to-speech.h:28-38) equal the restated policy's.  Checked for a sample of streams: the emitted waveform against
the policy port driven by single-utterance decodes (MAE / max-abs like stream-compare.cpp:100-124)."""
import numpy as np
import pytest

from conftest import snr_db

pytestmark = pytest.mark.gpu


def _policy_counters(n_codes, check=20, hold=32, min_commit=24):
    """decode_calls / decoded_codes of maybe_emit for a stream of n_codes codes (one per token)."""
    committed, calls, codes = 0, 0, 0
    for n in range(1, n_codes + 1):
        if n % check == 0:
            target = n - hold if n > hold else 0
            if target > committed and target - committed >= min_commit:
                calls += 1; codes += n; committed = target
    if n_codes > committed:                       # final flush
        calls += 1; codes += n_codes
    return calls, codes


def test_c4_1024_streams_700_codes(codec):
    from miocodec_b200 import synth
    from oracle import port
    S, N = 1024, 700
    rng = np.random.default_rng(44)
    phase = rng.integers(0, 20, size=S)
    embs = [synth.synth_voice(i % 16) for i in range(S)]
    codes = [synth.synth_codes(N, 4000 + i) for i in range(S)]
    streams = [codec.open_stream(embs[i]) for i in range(S)]
    got = [[] for _ in range(S)]
    done = [False] * S
    for tick in range(N + 20):
        for i, s in enumerate(streams):
            k = tick - int(phase[i])
            if 0 <= k < N:
                s.push(codes[i][k])
                if k == N - 1:
                    s.finish()
        codec.step_streams(streams)
        if tick % 20 == 19 or tick == N + 19:
            for i, s in enumerate(streams):
                if not done[i]:
                    a, last = s.read()
                    if a.size:
                        got[i].append(a)
                    done[i] = last
    calls, dcodes = _policy_counters(N)
    for i, s in enumerate(streams):
        assert done[i], i
        st = s.stats()
        n = sum(a.size for a in got[i])
        assert n == 1764 * N, (i, n)
        assert st == dict(decode_calls=calls, decoded_codes=dcodes, emitted_samples=1764 * N), (i, st)
    # waveform parity for a sample of streams against the policy port (same decoder, single-utterance calls)
    for i in (0, 517, 1023):
        out = np.concatenate(got[i])
        ref_out, st = port.stream_policy(lambda c: codec.decode_wave(c, embs[i]), codes[i])
        assert st["decode_calls"] == calls
        err = np.abs(out - ref_out)
        print(f"stream {i}: SNR {snr_db(out, ref_out):.1f} dB, MAE {err.mean():.2e}, max-abs {err.max():.2e}")
        assert snr_db(out, ref_out) > 55.0
    for s in streams:
        s.close()
