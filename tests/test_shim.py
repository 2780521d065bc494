"""The C++ drop-in layer (miotts-llama.cpp_b200/shim): reference-named entry points on top of the C ABI.

CPU: the shim library exports every symbol the reference's callers need — checked against the undefined symbols of
the reference's OWN src/test-to-speech.cpp compiled with the reference's OWN headers (so signatures and the
istft_cache class are the reference's, not ours).  GPU: tests/shim_demo.cpp (built by build.sh) runs the
decode_codes_to_audio sequence through the shim and is compared with the golden vectors of the unmodified reference."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

from conftest import ROOT, load_golden, snr_db

PKG = os.path.join(ROOT, "miotts-llama.cpp_b200", "miocodec_b200")
SHIM = os.path.join(PKG, "libmiocodec_shim.so")
REF = "/root/reference"

NEEDED = ["miocodec_load(", "miocodec_free(", "miocodec_decode(", "miocodec_sample_rate(", "miocodec_n_fft(",
          "miocodec_hop_length(", "miocodec_samples_per_token(", "load_voice_embedding(", "miocodec_print_tensors(",
          "istft_cache::istft_cache(int, int)", "istft_cache::n_fft() const", "istft_cache::cos_table() const",
          "istft_cache::hann_window() const", "istft(float const*, int, int, istft_cache const&)"]


def _defined(lib):
    out = subprocess.check_output(["nm", "-D", "-C", "--defined-only", lib], text=True)
    return [l.split(" ", 2)[2] for l in out.splitlines() if l.count(" ") >= 2]


def test_shim_exports_reference_entry_points():
    assert os.path.exists(SHIM), "run build() first"
    syms = _defined(SHIM)
    for n in NEEDED:
        assert any(s.startswith(n) or s == n for s in syms), n


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="reference tree not present on this host")
def test_reference_callers_link_against_the_shim():
    """Every miocodec/istft symbol that the reference's test-to-speech.cpp and main.cpp leave undefined is defined,
    with the identical mangled name, by the shim."""
    inc = ["-I", os.path.join(REF, "src"), "-I", os.path.join(REF, "third_party/llama.cpp/include"),
           "-I", os.path.join(REF, "third_party/llama.cpp/ggml/include")]
    mangled = set(subprocess.check_output(["nm", "-D", "--defined-only", SHIM], text=True).split())
    with tempfile.TemporaryDirectory() as td:
        for src in ("test-to-speech.cpp", "main.cpp"):
            obj = os.path.join(td, src + ".o")
            subprocess.check_call(["g++", "-std=c++17", "-O0", "-c", os.path.join(REF, "src", src), "-o", obj] + inc)
            und = subprocess.check_output(["nm", "-u", obj], text=True).split()
            want = [s for s in und if ("miocodec" in s or "istft" in s or "load_voice_embedding" in s) and s != "U"]
            if src == "test-to-speech.cpp":
                assert len(want) >= 8, want
            for s in want:
                assert s in mangled, f"{src}: {s} not exported by the shim"


@pytest.mark.parametrize("n_fft,win", [(392, 392), (16, 16), (30, 20)])
def test_istft_cache_tables_equal_the_reference(n_fft, win):
    """the drop-in istft_cache must hold the SAME numbers as the reference's (callers may read them through
    the eight getters): cos/sin tables built from f32 angles, Nyquist signs, periodic Hann — compared value by value with
    the reference's class compiled unmodified into oracle/_ref."""
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    exe = os.path.join(PKG, "shim_demo")
    assert os.path.exists(exe), "run build() first"
    want = ref.ref_istft_tables(n_fft, win)
    with tempfile.TemporaryDirectory() as td:
        pre = os.path.join(td, "t")
        r = subprocess.run([exe, "--dump-tables", str(n_fft), str(win), pre], capture_output=True, text=True, timeout=60)
        assert r.returncode == 0, r.stdout + r.stderr
        assert r.stdout.strip() == "n_fft=%d win_length=%d n_freq=%d n_mid=%d" % want["dims"]
        for key in ("cos", "sin", "nyq", "hann"):
            got = np.fromfile(pre + f".{key}.f32", dtype=np.float32)
            assert got.shape == want[key].shape, key
            assert np.array_equal(got.view(np.uint32), want[key].view(np.uint32)), key     # bit for bit


@pytest.mark.gpu
def test_shim_demo_matches_reference_golden(assets):
    from miocodec_b200 import synth
    g = load_golden("ref20")
    exe = os.path.join(PKG, "shim_demo")
    assert os.path.exists(exe)
    with tempfile.TemporaryDirectory() as td:
        codes_path = os.path.join(td, "codes.i32")
        g["codes"].astype(np.int32).tofile(codes_path)
        prefix = os.path.join(td, "out")
        r = subprocess.run([exe, assets[0], assets[1][int(g["voice"])], codes_path, prefix], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout + r.stderr
        assert "frames=360 samples=35280 sample_rate=44100 n_fft=392 hop=98 spt=1764" in r.stdout
        spec = np.fromfile(prefix + ".spec.f32", dtype=np.float32).reshape(360, 394)
        wave = np.fromfile(prefix + ".wave.f32", dtype=np.float32)
        batch0 = np.fromfile(prefix + ".batch0.f32", dtype=np.float32)
        assert snr_db(spec, g["spec"]) > 58.5 and snr_db(wave, g["wave"]) > 58.5
        assert np.array_equal(batch0, wave)
        # standalone istft (no codec loaded) against the oracle port
        from oracle import port
        x = np.fromfile(prefix + ".istft_in.f32", dtype=np.float32).reshape(50, 394)
        y = np.fromfile(prefix + ".istft_out.f32", dtype=np.float32)
        assert snr_db(y, port.istft(x)) > 88.0
