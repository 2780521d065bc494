"""Per-op parity of the tcgen05 tap-GEMM (the pattern of llama.cpp's tests/test-backend-ops.cpp):
one bare GEMM through the product kernel (mio_debug_gemm) against an fp64 product of the same split operands, over
the (N, K) shapes the decoder's linear layers use (src/miocodec.cpp:205-209) and ragged M."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _split16(x):
    hi = x.astype(np.float16)
    lo = (x - hi.astype(np.float32)).astype(np.float16)
    return hi.astype(np.float64), lo.astype(np.float64)


SHAPES = [(700, 2304, 768), (1, 768, 768), (257, 4096, 768), (513, 768, 2048), (1400, 1536, 512), (300, 3072, 512),
          (129, 512, 1536), (1000, 394, 512), (2000, 512, 128), (77, 1024, 512)]


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("npass", [3, 1])
def test_gemm_matches_fp64_product_of_the_split_operands(codec, M, N, K, npass):
    rng = np.random.default_rng(M * 7 + N * 3 + K + npass)
    A = rng.standard_normal((M, K)).astype(np.float32)
    B = (rng.standard_normal((N, K)) / np.sqrt(K)).astype(np.float32)
    out = codec.debug_gemm(A, B, npass)
    ah, al = _split16(A); bh, bl = _split16(B)
    want = ah @ bh.T if npass == 1 else (ah + al) @ (bh + bl).T
    rel = np.sqrt(np.mean((out - want) ** 2) / np.mean(want ** 2))
    print(f"M={M} N={N} K={K} npass={npass}: rel rms error {rel:.2e}, max-abs {np.abs(out - want).max():.2e}")
    assert out.shape == (M, N) and np.isfinite(out).all()
    assert rel < 3e-6          # fp32 accumulation of exact fp16 products; a dropped or doubled pass would be ~5e-4
    # an element of the output only depends on its own row of A and row of B: tile position must not matter
    if M >= 300:
        again = codec.debug_gemm(A[100:100 + 37], B, npass)
        assert np.array_equal(again, out[100:137])


def test_gemm_edge_values(codec):
    """zeros, exact small integers (exactly representable: the result must be exact) and an fp16-subnormal lo part"""
    A = np.zeros((64, 128), np.float32); B = np.zeros((64, 128), np.float32)
    assert np.array_equal(codec.debug_gemm(A, B, 3), np.zeros((64, 64), np.float32))
    rng = np.random.default_rng(1)
    A = rng.integers(-8, 9, size=(200, 256)).astype(np.float32)
    B = rng.integers(-8, 9, size=(96, 256)).astype(np.float32)
    assert np.array_equal(codec.debug_gemm(A, B, 3), (A.astype(np.float64) @ B.astype(np.float64).T).astype(np.float32))
    assert np.array_equal(codec.debug_gemm(A, B, 1), (A.astype(np.float64) @ B.astype(np.float64).T).astype(np.float32))
