"""GPU parity tests: every call goes through the C ABI (miocodec_b200.binding -> libmiocodec_b200.so) and is
compared with the oracle: golden vectors produced by the unmodified reference, the compiled reference itself when
oracle/_ref loads on this host, and the torch restatement (oracle/port.py) for per-stage checks.

Tolerances (DESIGN.md "parity floor", tests/test_parity_floor.py): the reference rounds conv1d operands to FP16
(ggml.c:4416), which makes its output discontinuous in fp32 round-off; the reference built for another ISA, and an exact
(fp64) evaluation of the same graph, only reach 59.3-59.9 dB SNR / 1.9-2.9e-3 max-abs against it.  That measured floor is
the bar: over 32 utterances the mean SNR must be within 0.6 dB of the mean floor and the mean max-abs within 1.25 x
(measured: GPU 59.43 +- 0.25 dB / 2.41e-3, floor 59.85 +- 0.24 dB / 2.32e-3, an exact fp64 evaluation 59.6-59.85 dB); a
single utterance must be within 1.0 dB of ITS floor (conftest.assert_at_parity_floor: both builds of the reference are
run live when the host has AVX-512) and within 1.25 x its max-abs.  Stages in front of the
first conv1d are continuous and are held to >= 105 dB (tcgen05 with chunked accumulation) and the ISTFT to >= 88 dB; every
later stage is held to fp32 accuracy by the teacher-forced tests (tests/test_gpu_stages.py).  Frame counts, sample counts
and indexing are exact; results are bit-identical across batching, chunking and delivery splits."""
import os

import numpy as np
import pytest

from conftest import assert_at_parity_floor, load_golden, snr_db

pytestmark = pytest.mark.gpu

VALIDATOR_SNR_MIN = 58.5     # product path vs a CUDA-core validation path: two independent evaluations, i.e. the floor again


def _voice(v):
    from miocodec_b200 import synth
    return synth.synth_voice(v)


@pytest.mark.parametrize("case", ["ref20", "rand33", "one"])
def test_wave_and_spec_match_reference_golden(codec, weights, assets, case):
    from oracle import port
    g = load_golden(case)
    codes, emb = g["codes"], _voice(int(g["voice"]))
    spec = codec.decode(codes, emb)                                   # miocodec_decode
    assert spec.shape == g["spec"].shape == (18 * len(codes), 394)    # exact frame indexing
    wave = codec.istft(spec)                                          # istft
    assert wave.shape == g["wave"].shape == (1764 * len(codes),)
    fused = codec.decode_wave(codes, emb)                             # decode_codes_to_audio
    assert fused.shape == wave.shape
    assert np.array_equal(fused, wave)                                # same kernels, deterministic: same bits
    hp, W = weights
    port_snr = snr_db(port.istft(port.decode(W, hp, codes, emb)), g["wave"])
    print(f"{case}: the fp32 port reaches {port_snr:.2f} dB, spec SNR {snr_db(spec, g['spec']):.2f} dB")
    if case == "one":      # a single code: 1764 samples, the statistic is too noisy for a 0.5 dB band
        assert snr_db(wave, g["wave"]) > 58.5 and np.abs(wave - g["wave"]).max() < 3.6e-3
    else:
        assert_at_parity_floor(wave, g["wave"], assets[0], codes, emb, case)
    assert snr_db(spec, g["spec"]) > 58.3


def test_istft_matches_reference_golden(codec):
    for case in ("istft50", "ref20", "one"):
        g = load_golden(case)
        wave = codec.istft(g["spec"])
        assert wave.shape == g["wave"].shape
        # the reference's DFT tables are built from f32 angles (src/istft.cpp:19-25, error ~6e-5 rad at k*n ~ 7e4),
        # an exact FFT therefore differs from it at the 1e-5 level
        assert snr_db(wave, g["wave"]) > 88.0
        assert np.abs(wave - g["wave"]).max() < 3e-5 * max(1.0, np.abs(g["wave"]).max())


def test_istft_is_linear_at_full_size(codec):
    """Size-independent property at BASELINE config C2 size (12600 frames = 700 codes)."""
    rng = np.random.default_rng(0)
    X = rng.standard_normal((12600, 394)).astype(np.float32)
    Y = rng.standard_normal((12600, 394)).astype(np.float32)
    wx, wy = codec.istft(X), codec.istft(Y)
    assert wx.shape == (12600 * 98,)
    wz = codec.istft((0.5 * X - 2.0 * Y).astype(np.float32))
    assert np.abs(wz - (0.5 * wx - 2.0 * wy)).max() < 2e-5
    # frame locality: changing frame t only changes samples [98 t - 147, 98 t + 245)
    X2 = X.copy(); X2[6000] += 1.0
    d = np.nonzero(codec.istft(X2) != wx)[0]
    assert d.min() >= 98 * 6000 - 147 and d.max() < 98 * 6000 + 245


@pytest.mark.parametrize("mode", ["tc3", "simt"])
def test_stages_before_first_conv_are_fp32_accurate(assets, weights, mode):
    from miocodec_b200 import binding, synth
    from oracle import port
    hp, W = weights
    codes, emb = synth.REFERENCE_TEST_CODES, _voice(0)
    _, inter = port.decode(W, hp, codes, emb, want_inter=True)
    c = binding.MioCodec(assets[0], 0, gemm=mode)
    c.set_option("debug_snapshots", "1")
    c.decode_wave(codes, emb)
    T = len(codes)
    lim = {"tc3": 105.0, "simt": 105.0}[mode]       # tcgen05 with chunked accumulation is as accurate as fp32 FMAs
    for name, r in [("embed", 1), ("pre0_qkv_rope", 1), ("pre0_attn", 1), ("pre0_out", 1), ("prenet_out", 1), ("upsample", 2)]:
        got = c.debug_read(name).reshape(-1, inter[name].shape[1])[: T * r]
        s = snr_db(got, inter[name])
        print(f"{mode} {name}: {s:.1f} dB")
        assert s > lim, (name, s)
    # after the convs the reference's own fp16 rounding dominates
    for name, r in [("prior", 2), ("decoder_x", 2), ("post", 2), ("up0", 6), ("up1", 18), ("spec", 18)]:
        ref = inter[name] if name != "spec" else port.decode(W, hp, codes, emb)
        got = c.debug_read(name).reshape(-1, ref.shape[1])[: T * r]
        assert np.isfinite(got).all()
        assert snr_db(got, ref) > 54.0, name
    c.close()


def test_batch_ragged_equals_single(codec):
    from miocodec_b200 import synth
    lens = [1, 7, 33, 64, 100, 2, 129]
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    codes = synth.synth_codes(int(cu[-1]), 21)
    vt = np.stack([_voice(v) for v in range(4)])
    vids = np.array([0, 1, 2, 3, 0, 1, 2], dtype=np.int32)
    out = codec.decode_wave_batch(codes, cu, vt, vids)
    assert out.shape == (1764 * int(cu[-1]),)
    for b, n in enumerate(lens):
        single = codec.decode_wave(codes[cu[b]:cu[b + 1]], vt[vids[b]])
        got = out[1764 * cu[b]: 1764 * cu[b + 1]]
        assert got.shape == single.shape
        assert np.array_equal(got, single), (b, n, snr_db(got, single))   # batch-invariant like the reference
    # permutation invariance: utterances are independent
    perm = [3, 0, 6, 2, 5, 1, 4]
    cu2 = np.concatenate([[0], np.cumsum([lens[p] for p in perm])]).astype(np.int32)
    codes2 = np.concatenate([codes[cu[p]:cu[p + 1]] for p in perm])
    out2 = codec.decode_wave_batch(codes2, cu2, vt, vids[perm])
    for j, p in enumerate(perm):
        assert np.array_equal(out2[1764 * cu2[j]: 1764 * cu2[j + 1]], out[1764 * cu[p]: 1764 * cu[p + 1]])


def test_multi_chunk_batches(assets, codec):
    """A batch larger than the workspace is split into chunks of whole utterances."""
    from miocodec_b200 import binding, synth
    small = binding.MioCodec(assets[0], 0, max_chunk_codes=128)
    lens = [60, 50, 100, 10, 90, 40]
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    codes = synth.synth_codes(int(cu[-1]), 33)
    vt = np.stack([_voice(v) for v in range(2)])
    vids = np.array([0, 1, 0, 1, 0, 1], dtype=np.int32)
    a = small.decode_wave_batch(codes, cu, vt, vids)
    b = codec.decode_wave_batch(codes, cu, vt, vids)
    assert np.array_equal(a, b)                                       # chunking never changes an utterance's bits
    with pytest.raises(binding.MioError) as ei:
        small.decode_wave(synth.synth_codes(400, 1), vt[0])          # one utterance larger than the workspace
    assert ei.value.code == binding.MIO_ERR_ARG
    small.close()


def test_edge_cases_and_errors(codec):
    from miocodec_b200 import binding, synth
    emb = _voice(0)
    assert codec.decode_wave(np.zeros(0, dtype=np.int32), emb).shape == (0,)      # decode_codes_to_audio: empty, true
    bad = synth.synth_codes(10, 3)
    bad[4] = codec.vocab                                                          # reference: GGML_ASSERT abort
    with pytest.raises(binding.MioError) as ei:
        codec.decode_wave(bad, emb)
    assert ei.value.code == binding.MIO_ERR_RANGE
    bad[4] = -7
    with pytest.raises(binding.MioError):
        codec.decode_wave(bad, emb)
    # the context stays usable after an error
    w = codec.decode_wave(synth.synth_codes(5, 4), emb, peak_normalize=True)
    assert abs(np.abs(w).max() - 0.95) < 1e-5                                     # 0.95 / max|s| (test-to-speech.cpp:289-300)
    # boundary codes 0 and vocab-1 decode
    w = codec.decode_wave(np.array([0, codec.vocab - 1, 0], dtype=np.int32), emb)
    assert np.isfinite(w).all() and w.shape == (3 * 1764,)
    # accessors
    assert (codec.sample_rate, codec.n_fft, codec.hop_length, codec.samples_per_token, codec.frames_per_code) == (44100, 392, 98, 1764, 18)


def test_simt_and_tensor_core_paths_agree(assets, codec):
    from miocodec_b200 import binding, synth
    codes, emb = synth.synth_codes(50, 8), _voice(2)
    simt = binding.MioCodec(assets[0], 0, gemm="simt")
    a, b = simt.decode_wave(codes, emb), codec.decode_wave(codes, emb)
    assert snr_db(b, a) > VALIDATOR_SNR_MIN
    simt.close()


def test_attention_validation_kernel_agrees(assets, codec):
    """MIO_ATTN=simt selects the fp32 CUDA-core attention; the TMA + ldmatrix + warp-MMA kernel must agree with it."""
    import os
    from miocodec_b200 import binding, synth
    codes, emb = synth.synth_codes(150, 9), _voice(1)
    os.environ["MIO_ATTN"] = "simt"
    try:
        simt = binding.MioCodec(assets[0], 0)
    finally:
        del os.environ["MIO_ATTN"]
    a, b = simt.decode_wave(codes, emb), codec.decode_wave(codes, emb)
    assert snr_db(b, a) > VALIDATOR_SNR_MIN
    simt.close()


def test_c2_full_size_against_live_reference(assets, codec):
    """BASELINE config C2: one 700-code utterance; compared with the compiled reference when it loads here."""
    from miocodec_b200 import synth
    from oracle import ref
    codes, emb = synth.synth_codes(700, 1234), _voice(0)
    wave = codec.decode_wave(codes, emb)
    assert wave.shape == (700 * 1764,) and np.isfinite(wave).all()
    if not ref.available():
        pytest.skip("oracle/_ref does not load on this host")
    R = ref.Reference(assets[0])
    w_ref = R.decode_wave(codes, emb)
    R.close()
    assert w_ref.shape == wave.shape
    assert_at_parity_floor(wave, w_ref, assets[0], codes, emb, "C2 T=700")


def test_streaming_matches_reference_policy(codec):
    """Streams fed one code per tick (the reference emits one code per LLM token) against the restated policy of
    src/test-to-speech.cpp:364-417, :499-571 driven by the same decoder."""
    from miocodec_b200 import synth
    from oracle import port
    n_streams, N = 3, [230, 100, 47]
    embs = [_voice(v) for v in range(n_streams)]
    codes = [synth.synth_codes(n, 70 + i) for i, n in enumerate(N)]
    streams = [codec.open_stream(e) for e in embs]
    got = [[] for _ in streams]
    sec0, passes0 = codec.stream_seconds()
    for tick in range(max(N)):
        for i, s in enumerate(streams):
            if tick < N[i]:
                s.push(codes[i][tick])
            if tick == N[i] - 1:
                s.finish()
        codec.step_streams(streams)
        for i, s in enumerate(streams):
            a, last = s.read()
            got[i].append(a)
    for i, s in enumerate(streams):
        a, last = s.read()
        assert last
        out = np.concatenate(got[i] + [a])
        ref_out, st = port.stream_policy(lambda c: codec.decode_wave(c, embs[i]), codes[i])
        mine = s.stats()
        assert mine == st, (mine, st)                                  # decode_calls / decoded_codes / emitted_samples exact
        assert out.shape == ref_out.shape == (N[i] * 1764,)            # sample_diff == 0 (stream-compare.cpp:258-262)
        assert np.array_equal(out, ref_out)                            # same decoder under both policies: same bits
        s.close()
    # StreamProfile's codec_sec + istft_sec: one timed device pass per round in which a cadence fired
    sec1, passes1 = codec.stream_seconds()
    assert sec1 > sec0 and 1 <= passes1 - passes0 <= max(N)


def test_c3_and_c5_full_batches_sampled_against_live_reference(assets, codec):
    """BASELINE configs C3 (256 x 250 codes, 16 voices) and a C5-shaped ragged batch (50-750 codes): the full batch runs
    on the GPU, a sample of its utterances is decoded by the compiled reference on the host.  Size-independent properties
    at full size: exact sample counts and offsets, finiteness, and independence of an utterance from its batch."""
    import bench
    from oracle import ref
    for name, picks in (("c3", [0, 100, 255]), ("c5", None)):
        wl = bench.make_workload(name, 0, 1)
        cu, codes, vt, vids = wl["cu"], wl["codes"], wl["voice_table"], wl["voice_ids"]
        if name == "c5":                                         # the first 256 utterances of the sweep keep the test short
            n = 256
            cu, vids = cu[: n + 1], vids[:n]
            codes = codes[: int(cu[-1])]
            lens = np.diff(cu)
            picks = [int(np.argmin(lens)), int(np.argmax(lens)), 7]
        wave = codec.decode_wave_batch(codes, cu, vt, vids)
        assert wave.shape == (int(cu[-1]) * 1764,) and np.isfinite(wave).all()
        R = ref.Reference(assets[0]) if ref.available() else None
        for b in picks:
            got = wave[1764 * int(cu[b]): 1764 * int(cu[b + 1])]
            alone = codec.decode_wave(codes[cu[b]:cu[b + 1]], vt[vids[b]])
            assert got.shape == alone.shape and np.array_equal(got, alone), (name, b)
            if R is not None:
                w_ref = R.decode_wave(codes[cu[b]:cu[b + 1]], vt[vids[b]])
                assert w_ref.shape == got.shape
                assert_at_parity_floor(got, w_ref, assets[0], codes[cu[b]:cu[b + 1]], vt[vids[b]], f"{name} utterance {b} ({int(cu[b + 1] - cu[b])} codes)")
        if R is not None:
            R.close()


def test_delivery_pipeline_splits_do_not_change_the_result(codec):
    """mio_decode_wave_batch cuts the batch for the overlapped D2H (one tail split whose fraction is fixed or
    auto-tuned): every variant must deliver the same waveform in the same layout."""
    from miocodec_b200 import synth
    lens = [250] * 80                                              # 20 080 rows: large enough for the tail split
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    codes = synth.synth_codes(int(cu[-1]), 77)
    vt = np.stack([_voice(v) for v in range(4)])
    vids = (np.arange(len(lens)) % 4).astype(np.int32)
    outs = {}
    try:
        for pct in (0, 12, 30, -1):
            codec.set_option("fetch_tail_percent", str(pct))
            outs[pct] = codec.decode_wave_batch(codes, cu, vt, vids)
            assert outs[pct].shape == (int(cu[-1]) * 1764,) and np.isfinite(outs[pct]).all()
    finally:
        codec.set_option("fetch_tail_percent", "-1")
    for pct in (12, 30, -1):
        assert np.array_equal(outs[pct], outs[0]), pct               # bit-identical under any split


def test_mean_parity_over_32_utterances(assets, codec):
    """the GPU path against the live reference on 32 utterances (40-120 codes, 16 voices), next to the
    reference's own cross-build floor on the SAME utterances.  Bars: mean SNR within 0.5 dB of the mean floor (measured gap
    0.42 dB with a standard error of 0.05, profiles/r3/summary.md section 3: a chunk length of 4 instead of 3 in the chunked
    accumulation, gap 0.55, already fails), mean max-abs within 1.25 x, no utterance more than 1 dB under its own floor."""
    from conftest import reference_floor
    from miocodec_b200 import synth
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref does not load on this host")
    rng = np.random.default_rng(77)
    lens = rng.integers(40, 121, size=32)
    R = ref.Reference(assets[0])
    g_snr, g_max, f_snr, f_max = [], [], [], []
    for i, L in enumerate(lens):
        codes, emb = synth.synth_codes(int(L), 1000 + i), synth.synth_voice(i % 16)
        w, w_ref = codec.decode_wave(codes, emb), R.decode_wave(codes, emb)
        fs, fm = reference_floor(assets[0], codes, emb)
        g_snr.append(snr_db(w, w_ref)); g_max.append(float(np.abs(w - w_ref).max())); f_snr.append(fs); f_max.append(fm)
    R.close()
    g_snr, g_max, f_snr, f_max = map(np.array, (g_snr, g_max, f_snr, f_max))
    print(f"32 utterances: GPU {g_snr.mean():.2f} +- {g_snr.std():.2f} dB [min {g_snr.min():.2f}], max-abs {g_max.mean():.2e} [{g_max.max():.2e}]; "
          f"floor {f_snr.mean():.2f} +- {f_snr.std():.2f} dB [min {f_snr.min():.2f}], max-abs {f_max.mean():.2e} [{f_max.max():.2e}]")
    assert g_snr.mean() > f_snr.mean() - 0.5
    assert g_max.mean() < 1.25 * f_max.mean()
    assert (g_snr > np.minimum(f_snr, 60.0) - 1.0).all() and g_snr.min() > 58.5


def test_fused_resblock_tail_is_bit_identical(assets):
    """The last resblock of a stage writes x + conv2(...) straight into the next product's (hi, lo) operand (EPI_SPLIT with a
    residual input) instead of a reduce-add into x followed by a split pass: the same fp32 add, so the waveform must not
    change by a single bit -- ragged batch, pad rows included."""
    from miocodec_b200 import binding, synth
    lens = [33, 1, 250, 17]
    cu = np.cumsum([0] + lens).astype(np.int32)
    codes = synth.synth_codes(int(cu[-1]), 321)
    vt = np.stack([synth.synth_voice(v) for v in range(3)]).astype(np.float32)
    vids = np.array([0, 1, 2, 1], dtype=np.int32)
    out = {}
    for tag in ("1", "0"):
        c = binding.MioCodec(assets[0], 0)
        c.set_option("fuse_tail", tag)
        out[tag] = (c.decode_wave_batch(codes, cu, vt, vids), c.kernel_launches())
        c.close()
    assert np.array_equal(out["1"][0], out["0"][0])
    assert out["0"][1] - out["1"][1] == 3          # three split launches gone
