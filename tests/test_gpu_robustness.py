"""Host-side robustness of the engine on the GPU: the TMA-descriptor cache
under thousands of distinct batch shapes, utterances longer than the initial RoPE table, two devices in one process."""
import numpy as np
import pytest

from conftest import snr_db

pytestmark = pytest.mark.gpu


def test_descriptor_cache_survives_thousands_of_shapes(codec):
    """Every distinct row count adds ~30 tensor-map keys; the cache is emptied at 4096.  Results before, across and
    after the eviction must be the same bits (the old cache handed out pointers that an eviction left dangling)."""
    from miocodec_b200 import synth
    emb = synth.synth_voice(1)
    probe = synth.synth_codes(37, 5)
    want = codec.decode_wave(probe, emb)
    seen_clear = 0
    for T in list(range(1, 330, 2)):                                  # 165 shapes x ~30 keys: crosses 4096 at least once
        w = codec.decode_wave(synth.synth_codes(T, 100 + T), emb)
        assert w.shape == (1764 * T,) and np.isfinite(w).all()
        if T % 41 == 0:
            assert np.array_equal(codec.decode_wave(probe, emb), want)
            seen_clear += 1
    assert seen_clear >= 3
    assert np.array_equal(codec.decode_wave(probe, emb), want)


def test_utterance_longer_than_the_initial_rope_table(assets, weights):
    """The RoPE table starts at 8192 positions and must grow (the reference computes the phases for any position,
    ggml-cpu/ops.cpp:5612-5626).  T = 4200 codes has decoder positions up to 8399.  Checked on the rotated q|k of prenet
    layer 0 against the oracle's RoPE (positions < 8192 at this resolution) and, for the decoder, by shift-consistency:
    rotating by position p must match the oracle's table at p for the rows past 8192."""
    import torch
    import torch.nn.functional as F
    from miocodec_b200 import binding, synth
    from oracle import port
    hp, W = weights
    Tn = 4200
    codes, emb = synth.synth_codes(Tn, 42), synth.synth_voice(0)
    c = binding.MioCodec(assets[0], 0)
    c.set_option("debug_snapshots", "1")
    c.set_option("debug_snapshot_filter", "prior,dec0_qkv_rope")
    w = c.decode_wave(codes, emb)
    assert w.shape == (1764 * Tn,) and np.isfinite(w).all()
    got = c.debug_read("dec0_qkv_rope").reshape(-1, 1536)[: 2 * Tn]
    x = torch.from_numpy(c.debug_read("prior").reshape(-1, 512)[: 2 * Tn].copy())
    c.close()
    # oracle side of decoder layer 0's q|k|v (decoder_layer, src/miocodec.cpp:333-343) from the GPU's own layer input
    C = 512
    t = lambda n: torch.from_numpy(np.ascontiguousarray(W[n]))
    cond = F.silu(torch.from_numpy(emb))
    m = t("wave_decoder.blk.0.attn_cond.weight") @ cond + t("wave_decoder.blk.0.attn_cond.bias")
    h = F.layer_norm(x, (C,), None, None, hp["norm_eps"]) * (1.0 + m[C:2 * C]) + m[:C]
    cos, sin = port.rope_table(2 * Tn, 64, hp["rope_theta"])
    q = port.rope_pairs((h @ t("wave_decoder.blk.0.attn_q.weight").T).view(2 * Tn, 8, 64), cos, sin).reshape(2 * Tn, C) * 0.125
    k = port.rope_pairs((h @ t("wave_decoder.blk.0.attn_k.weight").T).view(2 * Tn, 8, 64), cos, sin).reshape(2 * Tn, C)
    v = h @ t("wave_decoder.blk.0.attn_v.weight").T
    want = torch.cat([q, k, v], dim=1).numpy()
    for lo, hi in ((0, 8192), (8192, 2 * Tn)):
        s = snr_db(got[lo:hi], want[lo:hi])
        print(f"decoder layer 0 q|k|v, positions [{lo}, {hi}): {s:.1f} dB")
        assert s > 90.0, (lo, hi, s)


def test_two_devices_in_one_process(assets):
    """> 48 KB dynamic shared memory is a per-device attribute and the ISTFT tables are per context: a second context
    on another GPU of the same process must work and give the same bits."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    from miocodec_b200 import binding, synth
    codes, emb = synth.synth_codes(60, 3), synth.synth_voice(2)
    c0 = binding.MioCodec(assets[0], 0)
    c1 = binding.MioCodec(assets[0], 1)
    a, b = c0.decode_wave(codes, emb), c1.decode_wave(codes, emb)
    assert np.array_equal(a, b)
    spec = np.random.default_rng(0).standard_normal((90, 394)).astype(np.float32)
    assert np.array_equal(c0.istft(spec), c1.istft(spec))
    # interleaved use keeps working (each call selects its own device)
    assert np.array_equal(c0.decode_wave(codes, emb), a) and np.array_equal(c1.decode_wave(codes, emb), b)
    c0.close(); c1.close()


def test_pcm16_into_a_caller_supplied_device_buffer(codec):
    """mio_batch_pcm16_to_device (the store side of sharding.P2PGather): same bytes as the host delivery, written into a
    caller-owned device allocation at a 16-byte aligned offset; a misaligned destination is refused, not crashed on; the
    allocation exports a 64-byte IPC handle.  (Opening the handle needs a second process: bench.py --gpus >= 2 does that and
    checks every rank's bytes.)"""
    import torch
    from miocodec_b200 import binding, sharding, synth
    lens = [33, 17, 60]
    cu = np.cumsum([0] + lens).astype(np.int32)
    codes = synth.synth_codes(int(cu[-1]), 77)
    vt = np.stack([synth.synth_voice(1), synth.synth_voice(2)]).astype(np.float32)
    vids = np.array([0, 1, 0], dtype=np.int32)
    want = codec.decode_pcm16_batch(codes, cu, vt, vids)
    codec.batch_upload(codes, cu, vt, vids)
    codec.batch_run(False)
    n = int(cu[-1]) * codec.samples_per_token
    ptr = binding.device_alloc(0, 2 * (n + 64))
    try:
        assert len(binding.ipc_export(ptr)) == 64
        assert codec.pcm16_to_device(ptr + 2 * 24, n) == n          # 24 samples = 48 bytes in: 16-byte aligned
        codec.batch_sync()
        got = sharding.device_tensor(ptr + 2 * 24, n, "int16", 0).cpu().numpy()
        assert np.array_equal(got, want)
        with pytest.raises(binding.MioError):
            codec.pcm16_to_device(ptr + 2 * 3, n)                   # 6 bytes in
        with pytest.raises(binding.MioError):
            codec.pcm16_to_device(ptr, n - 1)                       # too small
    finally:
        torch.cuda.synchronize()
        binding.device_free(ptr)


def test_random_ragged_batches_are_batch_invariant():
    """tools/fuzz_batches.py: random batches (1-40 utterances of 1-400 codes, random voices, workspace sizes 512 .. 65 536
    codes, three delivery calls incl. PCM16): every utterance equals its single-utterance decode bit for bit."""
    import subprocess
    import sys
    from conftest import ROOT
    import os
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_batches.py"), "12", "5"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "FUZZ PASSED" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_reserve_presizes_without_changing_results(assets, codec):
    """mio_ctx_reserve: a dummy upload of the requested shape plus the streaming emit buffers; results afterwards are the
    same bits as on an unreserved context, for a batch and for a stream."""
    from miocodec_b200 import binding, synth
    c = binding.MioCodec(assets[0], 0)
    c.reserve(30000, 48, 1764 * 20 * 8)
    with pytest.raises(binding.MioError):
        c.reserve(-1, 1, 0)
    lens = [120, 7, 64]
    cu = np.cumsum([0] + lens).astype(np.int32)
    codes = synth.synth_codes(int(cu[-1]), 55)
    vt = np.stack([synth.synth_voice(1), synth.synth_voice(2)]).astype(np.float32)
    vids = np.array([1, 0, 1], dtype=np.int32)
    assert np.array_equal(c.decode_wave_batch(codes, cu, vt, vids), codec.decode_wave_batch(codes, cu, vt, vids))
    outs = []
    for ctx in (c, codec):
        s = ctx.open_stream(synth.synth_voice(3))
        got = []
        sc = synth.synth_codes(90, 56)
        for k in range(90):
            s.push(sc[k])
            if k == 89:
                s.finish()
            ctx.step_streams([s])
            a, last = s.read()
            got.append(a)
        assert last
        s.close()
        outs.append(np.concatenate(got))
    assert np.array_equal(outs[0], outs[1])
    c.close()


def test_two_contexts_from_two_threads(assets, codec):
    """Different contexts are independent: two host threads, each hammering its own context on the same GPU with a different
    batch, get the same bits as the shared context does alone."""
    import threading
    from miocodec_b200 import binding, synth
    jobs = []
    for seed, lens in ((71, [40, 3, 111]), (72, [64, 64])):
        cu = np.cumsum([0] + lens).astype(np.int32)
        codes = synth.synth_codes(int(cu[-1]), seed)
        vt = np.stack([synth.synth_voice(seed % 5), synth.synth_voice(1)]).astype(np.float32)
        vids = (np.arange(len(lens)) % 2).astype(np.int32)
        jobs.append((codes, cu, vt, vids, codec.decode_wave_batch(codes, cu, vt, vids).copy()))
    ctxs = [binding.MioCodec(assets[0], 0) for _ in jobs]
    bad = []

    def work(c, job):
        codes, cu, vt, vids, want = job
        for _ in range(12):
            if not np.array_equal(c.decode_wave_batch(codes, cu, vt, vids), want):
                bad.append(1)

    th = [threading.Thread(target=work, args=(c, j)) for c, j in zip(ctxs, jobs)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for c in ctxs:
        c.close()
    assert not bad
