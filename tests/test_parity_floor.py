"""The parity floor, measured instead of asserted.

The reference rounds every conv1d operand to FP16 (ggml.c:4416, ggml-cpu/ops.cpp:6232-6256).  Twelve such roundings in
series make its output discontinuous in fp32 round-off: two evaluations that differ anywhere in the last bit in front of
the first conv1d decorrelate to the quantisation noise of that chain, about -59.5 dB, whatever their own precision.
This file pins that number with the only independent evaluations available on a CPU:
  * the reference against ITSELF built for another ISA (oracle/_ref v3 = AVX2, v4 = AVX-512: other SIMD widths, other
    summation order) — the floor the GPU tests are held to (tests/test_gpu_parity.py)
  * an exact (fp64) evaluation of the same graph (oracle/port.py, dtype=float64) against the reference: no better
  * the sensitivity curve: white noise of relative size r injected in front of the first conv1d costs nothing up to
    r = 1e-6 and ~0.8 dB at 1e-5 — the measurement behind the chunked tensor-core accumulation (gemm3.cu)
north_star's ">= 60 dB / <= 1e-3" is therefore not reachable by ANY second implementation with these weights, the
reference on another CPU included; what can be asked of the GPU path is "as close to the reference as the reference's
other build"."""
import os

import numpy as np
import pytest

from conftest import snr_db

FLOOR_CASES = [("ref20", None, 0), ("rand33", (33, 11), 3), ("T100", (100, 5), 2), ("T250", (250, 3), 1)]


def _both_builds():
    from oracle import ref
    flags = ref._cpu_flags()
    here = os.path.dirname(os.path.abspath(ref.__file__))
    ok = all(os.path.exists(os.path.join(here, "_ref", f"libmio_ref_{t}.so")) for t in ("v3", "v4"))
    return ok and {"avx512f", "avx512bw", "avx512dq", "avx512vl", "avx512cd"} <= flags


def _load(tag, gguf):
    from oracle import ref
    old = os.environ.get("MIO_REF_VARIANT")
    os.environ["MIO_REF_VARIANT"] = tag
    try:
        return ref.Reference(gguf)
    finally:
        if old is None:
            os.environ.pop("MIO_REF_VARIANT", None)
        else:
            os.environ["MIO_REF_VARIANT"] = old


def _codes(case):
    from miocodec_b200 import synth
    name, gen, voice = case
    codes = synth.REFERENCE_TEST_CODES if gen is None else synth.synth_codes(*gen)
    return name, codes, synth.synth_voice(voice)


@pytest.mark.skipif(not _both_builds(), reason="needs oracle/_ref v3 + v4 and an AVX-512 host")
def test_reference_against_its_other_build_is_the_floor(assets):
    """Measured here and on the GPU box's host: 59.3-59.9 dB, max-abs 1.9e-3 .. 2.9e-3 (peak ~2)."""
    R3, R4 = _load("v3", assets[0]), _load("v4", assets[0])
    rows = []
    for case in FLOOR_CASES:
        name, codes, emb = _codes(case)
        w3, w4 = R3.decode_wave(codes, emb), R4.decode_wave(codes, emb)
        assert np.array_equal(w4, R4.decode_wave(codes, emb))            # each build alone is deterministic
        s, m = snr_db(w3, w4), float(np.abs(w3 - w4).max())
        rows.append((name, s, m))
        print(f"floor {name}: v3 vs v4 SNR {s:.2f} dB, max-abs {m:.2e}, peak {np.abs(w4).max():.2f}")
    R3.close(); R4.close()
    snrs = np.array([r[1] for r in rows])
    # the floor exists (the two builds do NOT agree to fp32 accuracy) and sits where DESIGN.md says it does
    assert 58.5 < snrs.min() and snrs.max() < 61.0, rows
    assert all(1.0e-3 < r[2] < 4.0e-3 for r in rows), rows               # max-abs of the floor itself exceeds north_star's 1e-3


def test_exact_evaluation_reaches_the_floor_and_no_more(assets, weights):
    """fp64 arithmetic everywhere (conv operands still rounded to FP16 as the graph prescribes) vs the reference."""
    import torch
    from oracle import port, ref
    if not ref.available():
        pytest.skip("oracle/_ref not built")
    hp, W = weights
    R = ref.Reference(assets[0])
    for case in FLOOR_CASES[:3]:
        name, codes, emb = _codes(case)
        w_ref = R.decode_wave(codes, emb)
        w64 = port.istft(port.decode(W, hp, codes, emb, dtype=torch.float64))
        w32 = port.istft(port.decode(W, hp, codes, emb))
        s64, s32, s6432 = snr_db(w64, w_ref), snr_db(w32, w_ref), snr_db(w32, w64)
        print(f"floor {name}: fp64 port vs reference {s64:.2f} dB, fp32 port vs reference {s32:.2f} dB, fp32 vs fp64 port {s6432:.2f} dB")
        assert 58.5 < s64 < 61.5 and 58.5 < s32 < 61.5 and 58.5 < s6432 < 61.5
    R.close()


def test_sensitivity_to_noise_in_front_of_the_first_conv(weights):
    """White noise of relative rms r added to the input of the first conv1d ('upsample' tap) of the fp64 evaluation:
    the output moves by the full floor as soon as r reaches fp32 round-off (1e-7), and the distance to an unperturbed evaluation only
    starts to grow beyond r ~ 1e-6.  A GPU path whose pre-conv stages are within ~2e-6 (114 dB) of exact is therefore at
    the floor; the unchunked tensor-core accumulation sat at 1.6e-5 (95.7 dB) and cost 0.8 dB."""
    import torch
    from miocodec_b200 import synth
    from oracle import port
    hp, W = weights
    codes, emb = synth.synth_codes(60, 5), synth.synth_voice(2)
    base = port.istft(port.decode(W, hp, codes, emb, dtype=torch.float64))
    out = {}
    for rel in (1e-7, 1e-6, 1e-4):
        g = torch.Generator().manual_seed(1)

        def pert(name, x):
            return x + rel * x.pow(2).mean().sqrt() * torch.randn(x.shape, generator=g, dtype=x.dtype)
        out[rel] = snr_db(port.istft(port.decode(W, hp, codes, emb, dtype=torch.float64, perturb=pert)), base)
        print(f"noise {rel:.0e} in front of the first conv1d: output moves to {out[rel]:.2f} dB of the unperturbed evaluation")
    assert out[1e-7] < 66.0                      # discontinuous: an fp32-ulp perturbation already costs everything above ~62 dB
    assert out[1e-6] > out[1e-4] + 1.5           # and the loss grows slowly with the size of the perturbation
    assert out[1e-6] > 58.0
