"""BASELINE config C4 at full scale: 1024 concurrent streams, each a 700-code sequence revealed one code per tick with
random phase offsets, decoded with the reference's streaming policy (commit holdback 32, min step 24, cadence 20,
full-prefix re-decode, llround indexing, 1323-sample crossfade; src/test-to-speech.cpp:364-417, :499-571).

Checked for every stream: emitted sample count == 1764 * codes (stream-compare's sample_diff == 0,
examples/stream-compare.cpp:258-262) and the StreamProfile counters (decode_calls, decoded_codes, emitted_samples,
src/test-to-speech.h:28-38) equal the reference's own (golden of the UNMODIFIED TestToSpeech::synthesize_stream_profiled
run with the scripted LLM stub, tests/golden/stream700.npz).  Checked for a sample of streams: the emitted waveform is
BIT-IDENTICAL to the policy port (oracle/port.stream_policy, itself pinned bit for bit against the reference's
orchestrator in tests/test_oracle.py) driven by single-utterance decodes of the same decoder.

test_streams_against_the_reference_orchestrator runs the reference's own streaming code live on the host (its CPU codec)
next to GPU streams fed the same codes: counters exact, waveform at the parity floor."""
import numpy as np
import pytest

from conftest import load_golden, snr_db

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
    g = load_golden("stream700")                 # the reference's own orchestrator on a 700-code stream
    assert (calls, dcodes) == (int(g["decode_calls"]), int(g["decoded_codes"])) == (18, 7160)
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
        assert np.array_equal(out, ref_out)      # batched streaming == the policy over single decodes, bit for bit
    for s in streams:
        s.close()


@pytest.mark.parametrize("N", [7, 47, 100, 230])
def test_streams_against_the_reference_orchestrator(codec, assets, N):
    """GPU stream vs the reference's OWN streaming session (oracle/_ref: unmodified src/test-to-speech.cpp + stub LLM +
    the reference's CPU codec) on the same codes: StreamProfile counters exact, sample count exact, waveform at the
    parity floor (the two codecs differ, the policy must not)."""
    from miocodec_b200 import synth
    from oracle import ref
    g = load_golden(f"stream{N}")
    codes, vid = g["codes"], int(g["voice"])
    s = codec.open_stream(synth.synth_voice(vid))
    got = []
    for k in range(N):
        s.push(codes[k])
        if k == N - 1:
            s.finish()
        codec.step_streams([s])
        a, last = s.read()
        got.append(a)
    assert last
    out = np.concatenate(got)
    st = s.stats()
    s.close()
    assert st == dict(decode_calls=int(g["decode_calls"]), decoded_codes=int(g["decoded_codes"]), emitted_samples=int(g["emitted_samples"]))
    assert out.shape == (1764 * N,)
    if ref.available():
        want, chunks, rst = ref.ref_stream_session(assets[0], assets[1][vid], codes)
        assert rst["decode_calls"] == st["decode_calls"] and rst["decoded_codes"] == st["decoded_codes"]
    else:
        if N > 100:
            pytest.skip("golden audio is stored for N <= 100 only")
        want = g["audio"]
    err = np.abs(out - want)
    sn = snr_db(out, want)
    print(f"stream N={N}: vs the reference orchestrator SNR {sn:.2f} dB, MAE {err.mean():.2e}, max-abs {err.max():.2e}")
    assert sn > 58.0 and err.max() < 4.5e-3


def test_unread_rounds_spill_out_of_the_pinned_pool(codec):
    """The emitted samples of a round stay in a pinned block that the stream queues reference (no second host copy); at
    most EMIT_POOL_MAX = 8 blocks may be referenced at once, later rounds are copied into the stream.  A consumer that
    reads nothing until the end crosses that limit: 12 decode rounds for 480 codes (one every 40 codes, plus the flush).  Same samples, bit for bit, as the
    policy port over single decodes, in both regimes, and a second stream that reads every round is not disturbed."""
    from miocodec_b200 import synth
    from oracle import port
    N = 480
    emb_a, emb_b = synth.synth_voice(3), synth.synth_voice(4)
    ca, cb = synth.synth_codes(N, 901), synth.synth_codes(N, 902)
    lazy, eager = codec.open_stream(emb_a), codec.open_stream(emb_b)
    got_b = []
    for k in range(N):
        lazy.push(ca[k]); eager.push(cb[k])
        if k == N - 1:
            lazy.finish(); eager.finish()
        codec.step_streams([lazy, eager])
        a, last_b = eager.read()
        got_b.append(a)
    assert last_b
    got_a, last_a = [], False
    while not last_a:
        a, last_a = lazy.read(1 << 16)             # small reads: slices are consumed across segment boundaries
        got_a.append(a)
    st = lazy.stats()
    lazy.close(); eager.close()
    assert st["decode_calls"] >= 11
    for got, codes, emb in ((np.concatenate(got_a), ca, emb_a), (np.concatenate(got_b), cb, emb_b)):
        want, _ = port.stream_policy(lambda c: codec.decode_wave(c, emb), codes)
        assert got.shape == want.shape == (1764 * N,)
        assert np.array_equal(got, want)


def test_random_stream_sessions_match_the_policy_port():
    """tests/fuzz_streams.py: random concurrent sessions (1-12 streams of 1-420 codes, push bursts of 1-37 codes, random read
    sizes, some consumers that read nothing until the end): every stream bit-identical to the policy port over single
    decodes, StreamProfile counters equal."""
    import os
    import subprocess
    import sys
    from conftest import ROOT
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "fuzz_streams.py"), "6", "9"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "STREAM FUZZ PASSED" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_split_rounds_give_the_same_stream(codec):
    """mio_streams_step_begin / _end with pushes and reads in between (the host works while the GPU decodes) produce the same
    samples, bit for bit, and the same counters as the synchronous step; a stream of the round in flight can be closed
    (the library drains the round first)."""
    from miocodec_b200 import synth
    from oracle import port
    N = 230
    embs = [synth.synth_voice(2), synth.synth_voice(6)]
    cs = [synth.synth_codes(N, 811), synth.synth_codes(N - 37, 812)]
    streams = [codec.open_stream(e) for e in embs]
    got = [[], []]
    last = [False, False]
    for k in range(N):
        for i, s in enumerate(streams):
            if k < len(cs[i]):
                s.push(cs[i][k])
                if k == len(cs[i]) - 1:
                    s.finish()
        codec.step_streams_begin(streams)
        for i, s in enumerate(streams):                 # reads while the round is in flight: earlier rounds only
            if not last[i]:
                a, last[i] = s.read(50000)
                got[i].append(a)
        codec.step_streams_end(streams)
    for i, s in enumerate(streams):
        while not last[i]:
            a, last[i] = s.read()
            got[i].append(a)
        want, st = port.stream_policy(lambda c, e=embs[i]: codec.decode_wave(c, e), cs[i])
        out = np.concatenate(got[i])
        mine = s.stats()
        assert out.shape == want.shape and np.array_equal(out, want)
        assert mine["decode_calls"] == st["decode_calls"] and mine["decoded_codes"] == st["decoded_codes"]
    # closing a stream whose round is in flight
    t = codec.open_stream(embs[0])
    t.push(cs[0][:60])
    codec.step_streams_begin([t])
    t.close()
    u = codec.open_stream(embs[0])                      # the context is still usable
    u.push(cs[0][:60]); u.finish()
    codec.step_streams([u])
    a, fin = u.read()
    assert fin and a.size == 60 * 1764
    u.close()
    for s in streams:
        s.close()


def test_context_destroyed_before_its_streams(assets):
    """mio_ctx_destroy with streams still open (Python: codec.close() before the Stream objects are closed or collected):
    the streams are detached, keep the samples they had not been read yet as their own copies, and can be read, queried
    and closed afterwards -- no use of freed engine state, no read from freed pinned memory.  Also with a split round still
    in flight when the context goes."""
    from miocodec_b200 import binding, synth
    emb = synth.synth_voice(3)
    codes = synth.synth_codes(120, 909)

    def session(destroy_first, in_flight):
        c = binding.MioCodec(assets[0], 0)
        s, t = c.open_stream(emb), c.open_stream(emb)
        for st in (s, t):
            st.push(codes)
            st.finish()
        if in_flight:
            assert c.step_streams_begin([s, t]) == 1       # the round is enqueued, not waited for
        else:
            c.step_streams([s, t])
        if destroy_first:
            c.close()                                       # drains the round, detaches s and t
        a, last = s.read()
        if in_flight and not destroy_first:
            assert a.size == 0
            c.step_streams_end([s, t])
            a, last = s.read()
        st_ = s.stats()
        s.close(); t.close()                                # t still holds unread samples when it is closed
        if not destroy_first:
            c.close()
        return a, st_

    want, stats = session(False, False)
    assert want.size == 120 * 1764 and stats["emitted_samples"] == want.size
    got, st_ = session(True, False)
    assert np.array_equal(got, want) and st_ == stats
    # destroyed with the FIRST pass of the round in flight: that pass is drained and delivered (the commit of the cadence at
    # 60 codes: samples of codes [0, 28)); the later passes never ran
    got, st_ = session(True, True)
    assert got.size == 28 * 1764 and np.array_equal(got, want[:got.size])
    assert st_["decode_calls"] == 1 and st_["decoded_codes"] == 60 and st_["emitted_samples"] == got.size
