"""CPU tests of the host side: the C-ABI library loads without a GPU and exports every declared symbol, the GGUF
reader / voice loader behave like the reference's, errors are reported (never a silent CPU fallback), and the FFT
arithmetic of the ISTFT kernel (shared host/device header) is exact."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

PKG = os.path.join(ROOT, "miotts-llama.cpp_b200")


def _lib():
    from miocodec_b200 import binding
    if not os.path.exists(binding.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return binding.lib()


def test_library_exports_every_declared_symbol():
    from miocodec_b200 import binding
    L = _lib()
    header = open(os.path.join(ROOT, "include", "miocodec_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(mio_[a-z0-9_]+)\s*\(", header)))
    assert declared == sorted(binding.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name
    assert b"sm_100a" in L.mio_version()


def test_no_cpu_fallback_and_error_codes(assets, tmp_path):
    import torch
    from miocodec_b200 import binding
    L = _lib()
    h = C.c_void_p()
    rc = L.mio_ctx_create(b"/nonexistent/miocodec.gguf", 0, C.byref(h))
    assert rc == binding.MIO_ERR_IO and not h.value
    bad = tmp_path / "bad.gguf"
    bad.write_bytes(b"NOPE" + b"\0" * 64)
    assert L.mio_ctx_create(str(bad).encode(), 0, C.byref(h)) == binding.MIO_ERR_FORMAT
    if not torch.cuda.is_available():
        rc = L.mio_ctx_create(assets[0].encode(), 0, C.byref(h))
        assert rc == binding.MIO_ERR_CUDA and not h.value
        assert b"no CPU fallback" in L.mio_last_error() or b"CUDA" in L.mio_last_error()
        assert binding.miocodec_load(assets[0]) is None          # reference contract: nullptr on failure
    assert L.mio_ctx_sample_rate(None) == 0                      # null-safe accessors


def test_voice_loader_matches_reference_layout(assets, tmp_path):
    from miocodec_b200 import binding, synth
    _lib()
    for v in (0, 5):
        emb = binding.load_voice_embedding(assets[1][v])
        assert emb.dtype == np.float32 and emb.shape == (128,)
        assert np.array_equal(emb, synth.synth_voice(v))
    # first tensor whatever its name; non-F32 rejected; missing file -> empty (src/miocodec.cpp:816-853)
    w = synth.GGUFWriter("mio-embedding")
    w.add_u32("mio.embedding.dim", 4)
    w.add_tensor("anything", np.arange(4, dtype=np.float32))
    w.add_tensor("second", np.ones(4, dtype=np.float32))
    p = str(tmp_path / "v.gguf")
    w.write(p)
    assert np.array_equal(binding.load_voice_embedding(p), np.arange(4, dtype=np.float32))
    w = synth.GGUFWriter("mio-embedding")
    w.add_tensor("h", np.arange(4, dtype=np.float16))
    w.write(p)
    assert binding.load_voice_embedding(p).size == 0
    assert binding.load_voice_embedding(str(tmp_path / "missing.gguf")).size == 0
    from oracle import ref
    if ref.available():
        R = ref.Reference(assets[0])
        assert np.array_equal(R.load_voice(assets[1][3]), binding.load_voice_embedding(assets[1][3]))
        R.close()


def test_print_tensors_lists_all_227(assets, capfd):
    from miocodec_b200 import binding
    _lib()
    binding.miocodec_print_tensors(assets[0])
    out = capfd.readouterr().out
    assert "Tensors in" in out and ": 227" in out
    assert "token_embd" in out and "istft_head.out.weight" in out and "miocodec.wave_upsampler.factors" in out


def test_fft_arithmetic_of_the_istft_kernel(tmp_path):
    """csrc/istft_math.h is compiled for the host and checked against the reference's irfft formula."""
    exe = str(tmp_path / "istft_math_test")
    subprocess.check_call(["g++", "-O2", "-std=c++17", os.path.join(PKG, "csrc", "istft_math_test.cpp"), "-o", exe])
    for seed in ("1", "2"):
        out = subprocess.run([exe, seed], capture_output=True, text=True)
        assert out.returncode == 0, out.stdout
        err = float(out.stdout.split()[1])
        assert err < 5e-7


def test_sharding_balances_and_covers():
    from miocodec_b200 import sharding
    rng = np.random.default_rng(5)
    lens = rng.integers(50, 751, size=4096)
    for world in (1, 2, 4, 8):
        parts = [sharding.shard_utterances(lens, world, r) for r in range(world)]
        allidx = np.sort(np.concatenate(parts))
        assert np.array_equal(allidx, np.arange(len(lens)))          # every utterance exactly once
        loads = np.array([lens[p].sum() for p in parts], dtype=np.float64)
        assert loads.max() / loads.mean() < 1.01                      # longest-first packing: <1% imbalance
    assert sharding.shard_utterances(np.array([], dtype=np.int64), 4, 2).size == 0
    assert sharding.shard_utterances(np.array([7]), 4, 3).size + sharding.shard_utterances(np.array([7]), 4, 0).size == 1


def test_corrupt_tensor_extents_are_rejected_at_open(tmp_path):
    """Sizes and offsets come from the file: a tensor whose extent or data range does not fit the file must be refused
    with MIO_ERR_FORMAT when the file is opened (no bad_alloc / wild seek across the C ABI)."""
    import struct
    from miocodec_b200 import binding

    def gguf(ne, offset, payload=64):
        def s(b):
            e = b.encode()
            return struct.pack("<Q", len(e)) + e
        head = b"GGUF" + struct.pack("<IQQ", 3, 1, 1) + s("general.architecture") + struct.pack("<I", 8) + s("miocodec")
        head += s("token_embd") + struct.pack("<I", len(ne)) + struct.pack("<%dQ" % len(ne), *ne) + struct.pack("<IQ", 0, offset)
        head += b"\0" * ((-len(head)) % 32)
        return head + b"\0" * payload

    L = binding.lib()
    import ctypes as C
    for name, blob in (("huge", gguf([1 << 40, 1 << 30], 0)), ("negative", gguf([(1 << 64) - 8, 4], 0)), ("past_end", gguf([4, 4], 1 << 20)),
                       ("too_long", gguf([768, 12800], 0))):
        p = tmp_path / f"{name}.gguf"
        p.write_bytes(blob)
        h = C.c_void_p()
        rc = L.mio_ctx_create(str(p).encode(), 0, C.byref(h))
        assert rc == binding.MIO_ERR_FORMAT, (name, rc, L.mio_last_error())
        assert b"corrupt GGUF" in L.mio_last_error()
        out = np.zeros(8, dtype=np.float32)
        assert L.mio_voice_load(str(p).encode(), out.ctypes.data, 8) == binding.MIO_ERR_FORMAT


def test_p2p_gather_layout_covers_every_sample_once():
    """sharding.p2p_layout (the job buffer of the P2P gather): group-major, every (rank, group) slice on a multiple of 8
    samples, slices disjoint, every utterance exactly once, groups of a rank contiguous in its shard order, and each
    rank's groups concatenated give its shard back (what lets a rank check its own bytes against its plain decode)."""
    from miocodec_b200 import sharding
    rng = np.random.default_rng(11)
    for world, G, n in ((8, 8, 4096), (2, 4, 37), (3, 5, 4), (1, 3, 10)):
        lens = rng.integers(1, 751, size=n)
        L = sharding.p2p_layout(lens, world, 1764, G)
        seen = np.zeros(L["total"], dtype=np.int8)
        for i, (o, cnt) in enumerate(L["slices"]):
            assert cnt == int(lens[i]) * 1764
            seen[o:o + cnt] += 1
        assert seen.max() == 1 and int(seen.sum()) == int(lens.sum()) * 1764
        for r in range(world):
            shard = sharding.shard_utterances(lens, world, r)
            assert np.array_equal(np.concatenate([np.asarray(g, dtype=np.int64) for g in L["groups"][r]]) if len(shard) else np.arange(0), shard)
            for g in range(G):
                assert L["off"][r][g] % 8 == 0
                lo, hi = L["group_range"][g]
                assert lo <= L["off"][r][g] and L["off"][r][g] + L["count"][r][g] <= hi
        for g in range(G - 1):
            assert L["group_range"][g][1] <= L["group_range"][g + 1][0] + 0 or L["group_range"][g][1] == L["group_range"][g + 1][0]
        # balance: no group of a rank carries more than twice the mean once there is enough work
        if n >= 1000:
            for r in range(world):
                c = np.array(L["count"][r], dtype=np.float64)
                assert c.max() < 1.5 * c.mean()


def test_bench_arms_share_one_config_object():
    """bench.py prints the SAME `config` in the GPU arm and in --impl reference (the driver compares them): both are built by
    bench_config() from the workload alone."""
    import importlib.util
    import os
    from conftest import ROOT
    spec = importlib.util.spec_from_file_location("bench_module", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    for name in ("c3", "c2", "c5"):
        a = bench.bench_config(bench.make_workload(name, 0, 1), 1)
        b = bench.bench_config(bench.make_workload(name, 0, 1), 1)
        assert a == b and set(a) == {"workload", "codes_per_gpu_per_step", "audio_s_per_gpu_per_step", "weights", "parallelism", "l2"}
    c3 = bench.bench_config(bench.make_workload("c3", 0, 1), 1)
    assert c3["codes_per_gpu_per_step"] == 64000 and c3["audio_s_per_gpu_per_step"] == 2560.0
    # C5 is ONE sweep sharded over the ranks: the shards of all ranks add up to the 4096 utterances
    lens = [bench.make_workload("c5", r, 4)["lens"] for r in range(4)]
    assert sum(len(l) for l in lens) == 4096 and sum(int(l.sum()) for l in lens) == int(bench.make_workload("c5", 0, 1)["lens"].sum())


def test_oracle_is_reached_from_test_infrastructure_only():
    """oracle/ is the checker: the product package and tools/ must not import, link or execute it (bench.py may, in its
    cpu_baseline / --impl reference legs; __graft_entry__.smoke() may, as the checker)."""
    import os
    import re
    from conftest import ROOT
    pat = re.compile(r"(from\s+oracle\b|import\s+oracle\b|oracle/_ref|libmio_ref)")
    offenders = []
    for top in ("miotts-llama.cpp_b200", "tools", "include"):
        for d, _, files in os.walk(os.path.join(ROOT, top)):
            if os.sep + "build" in d or "dropin" in d:
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh", ".sh")):
                    p = os.path.join(d, f)
                    if f.startswith("gpu_job_") or f in ("profile.sh", "profile_gemm_layers.sh"):
                        continue                                    # GPU job scripts only invoke pytest / bench.py
                    if pat.search(open(p, errors="replace").read()):
                        offenders.append(os.path.relpath(p, ROOT))
    assert offenders == [], offenders
