"""SURVEY.md row A16 and the storage types of the matrices: files other than the default synthetic one.

The reference reads the upsampler geometry (factors, kernel sizes, channel widths) from the file at run time and accepts
any kernel with (k - f) even (/root/reference/src/miocodec.cpp:476-481, :677-697); the real miocodec.gguf's values are not
known offline (SURVEY.md 8c "Unknowns").  Matrices stored as F16 / BF16 change the reference's ARITHMETIC: ggml rounds the
activations to the storage type and multiplies once (ggml-cpu.c:1291-1326, ops.cpp:5909-5995 for ConvTranspose1d).

Each case writes a GGUF with that geometry / storage, decodes it on the GPU and with the reference itself (oracle/_ref,
live) and holds the GPU to the reference's own cross-build floor FOR THAT FILE (v3 vs v4 build of the reference: ~59.5 dB
for F32 files, ~49.5 dB for F16, ~34.5 dB for BF16 -- every rounding point is a discontinuity the two builds hit
differently).  A product that ignored the storage type would sit at the F32-vs-F16 distance instead.
"""
import os

import numpy as np
import pytest

from conftest import both_reference_builds_available, reference_floor, snr_db

pytestmark = pytest.mark.gpu

GEOMETRY = {
    "k=f [3,3]": dict(hp=dict(up_kernels=(3, 3))),
    "k [7,5]": dict(hp=dict(up_kernels=(7, 5))),
    "k [5,9]": dict(hp=dict(up_kernels=(5, 9))),
    "k=5f [15,15]": dict(hp=dict(up_kernels=(15, 15))),
    "factors [9,1]": dict(hp=dict(up_factors=(9, 1), up_kernels=(9, 3))),
    "factors [1,9]": dict(hp=dict(up_factors=(1, 9), up_kernels=(5, 27))),
    "factors [2,2] (784 samples per code)": dict(hp=dict(up_factors=(2, 2), up_kernels=(4, 10))),
    "widths 384/192, out_proj 448": dict(hp=dict(up_channels=(384, 192), out_proj_dim=448)),
    "widths 128/64": dict(hp=dict(up_channels=(128, 64))),
    "widths 512/512": dict(hp=dict(up_channels=(512, 512))),
}
STORAGE = {
    "F16 matrices and conv kernels": dict(matrices="f16"),
    "F16 matrices, F32 conv kernels": dict(matrices="f16", convs="f32"),
    "BF16 matrices": dict(matrices="bf16"),
}
# v3-v4 floors recorded on an AVX-512 host (profiles/r2f), used when this host cannot run both reference builds
RECORDED_GEOMETRY_FLOOR = {"k=f [3,3]": (59.81, 1.8e-3), "k [7,5]": (59.20, 2.5e-3), "k [5,9]": (59.56, 2.5e-3), "k=5f [15,15]": (60.23, 2.7e-3),
                           "factors [9,1]": (58.04, 3.9e-3), "factors [1,9]": (60.24, 1.6e-3),
                           "factors [2,2] (784 samples per code)": (57.63, 3.4e-3), "widths 384/192, out_proj 448": (59.74, 1.8e-3),
                           "widths 128/64": (58.58, 3.8e-3), "widths 512/512": (61.21, 6.0e-4)}
RECORDED_FLOOR = {"F16 matrices and conv kernels": (49.38, 7.4e-3), "F16 matrices, F32 conv kernels": (50.45, 5.0e-3),
                  "BF16 matrices": (34.06, 5.0e-2)}


def _decode_pair(tmp_path, kw, T=60, seed=9):
    from miocodec_b200 import binding, synth
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref is not built on this host")
    path = str(tmp_path / "variant.gguf")
    hp = synth.write_codec_gguf(path, seed=1, **kw)
    codes, emb = synth.synth_codes(T, seed), synth.synth_voice(5)
    R = ref.Reference(path)
    want = R.decode_wave(codes, emb)
    R.close()
    c = binding.MioCodec(path, 0)
    try:
        assert c.samples_per_token == hp["samples_per_token"]
        got = c.decode_wave(codes, emb)
        # two more utterances in one batch must not change the first one (row space at the new resolutions)
        others = [synth.synth_codes(17, 1), synth.synth_codes(33, 2)]
        cu = np.cumsum([0, T, 17, 33]).astype(np.int32)
        batch = c.decode_wave_batch(np.concatenate([codes] + others), cu, emb[None, :], np.zeros(3, np.int32))
        assert np.array_equal(batch[: got.size], got)
    finally:
        c.close()
    assert got.shape == want.shape == (T * hp["samples_per_token"],)
    f_snr, f_max = reference_floor(path, codes, emb)
    s, m = snr_db(got, want), float(np.abs(got - want).max())
    return s, m, f_snr, f_max


@pytest.mark.parametrize("name", list(GEOMETRY))
def test_upsampler_geometry(tmp_path, name):
    s, m, f_snr, f_max = _decode_pair(tmp_path, GEOMETRY[name])
    if not both_reference_builds_available():
        f_snr, f_max = RECORDED_GEOMETRY_FLOOR[name]
    print(f"{name}: SNR {s:.2f} dB (reference v3-v4 floor {f_snr:.2f}), max-abs {m:.2e} (floor {f_max:.2e})")
    # relative to THIS file's floor (a random file with a 1-tap second stage sits at 58.0 dB reference-to-reference)
    assert s > min(f_snr, 60.0) - 1.0 and m < 1.5 * max(f_max, 2.9e-3)


@pytest.mark.parametrize("name", list(STORAGE))
def test_stored_matrix_types_follow_the_reference_arithmetic(tmp_path, name):
    s, m, f_snr, f_max = _decode_pair(tmp_path, STORAGE[name])
    if not both_reference_builds_available():
        f_snr, f_max = RECORDED_FLOOR[name]
    print(f"{name}: SNR {s:.2f} dB (reference v3-v4 floor {f_snr:.2f}), max-abs {m:.2e} (floor {f_max:.2e})")
    assert s > f_snr - 1.5 and m < 2.0 * f_max


def test_same_values_in_f32_and_f16_storage_differ_like_the_reference(tmp_path):
    """Two files with IDENTICAL numbers (pre-rounded to F16), one storing them as F32 and one as F16: the reference's
    outputs differ by its activation rounding (48.5 dB).  The GPU must (a) match the reference on each
    file and (b) show the same distance between the two files -- i.e. it really switches arithmetic on the storage type."""
    from miocodec_b200 import binding, synth
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref is not built on this host")
    codes, emb = synth.synth_codes(60, 9), synth.synth_voice(5)
    out = {}
    for tag, kw in (("f32", dict(matrices="f32", convs="f32", pre_round="f16")), ("f16", dict(matrices="f16", convs="f32"))):
        path = str(tmp_path / f"{tag}.gguf")
        synth.write_codec_gguf(path, seed=1, **kw)
        R = ref.Reference(path)
        c = binding.MioCodec(path, 0)
        out[tag] = (R.decode_wave(codes, emb), c.decode_wave(codes, emb))
        R.close()
        c.close()
    d_ref = snr_db(out["f16"][0], out["f32"][0])
    d_gpu = snr_db(out["f16"][1], out["f32"][1])
    print(f"F16-stored vs F32-stored, same values: reference {d_ref:.2f} dB, GPU {d_gpu:.2f} dB; "
          f"GPU vs reference: F32 file {snr_db(out['f32'][1], out['f32'][0]):.2f} dB, F16 file {snr_db(out['f16'][1], out['f16'][0]):.2f} dB")
    assert d_ref < 52.0                       # the storage type matters to the reference ...
    assert abs(d_gpu - d_ref) < 1.5           # ... and equally to the GPU path
    assert snr_db(out["f32"][1], out["f32"][0]) > 58.0


def test_rejected_geometries(tmp_path):
    """What the reference cannot decode correctly either: (k - f) odd gives the wrong length after its trim
    (src/miocodec.cpp:689-696); BF16 ConvTranspose kernels abort in ggml."""
    from miocodec_b200 import binding, synth
    import ctypes as C
    L = binding.lib()
    for kw, msg in ((dict(hp=dict(up_kernels=(8, 9))), b"(k - f) even"),
                    (dict(hp=dict(up_kernels=(18, 9))), b"(k - f) even"),
                    (dict(hp=dict(up_kernels=(21, 9))), b"wider than 5 x stride"),
                    (dict(hp=dict(up_channels=(250, 128))), b"multiples of 64")):
        path = str(tmp_path / "bad.gguf")
        synth.write_codec_gguf(path, seed=1, **kw)
        h = C.c_void_p()
        rc = L.mio_ctx_create(path.encode(), 0, C.byref(h))
        assert rc != 0 and msg in L.mio_last_error(), (kw, rc, L.mio_last_error())
