// CUDA-core (SIMT) validation GEMM: the same split operands and the same fused epilogue semantics as the tcgen05
// kernel (gemm3.cu), computed with plain fp32 FMAs and libdevice transcendentals.  Selected with
// mio_ctx_set_option("gemm", "simt") / MIO_GEMM=simt.  It is a validator, not a fallback: it runs on the same B200,
// and the parity tests use it to separate tensor-core plumbing / accumulation effects from model bugs
// (profiles/r2a_parity_attribution.md).
//
// Tap-GEMM: D[m][n] = sum_{tap} sum_k A[m + tap - taps/2][k] * B[n][tap*K + k].
#include "kernels.cuh"
#include "ptx.cuh"

#include <cstdio>

namespace mio {

// ---------------------------------------------------------------------------------------------
// epilogue shared by the tcgen05 and the SIMT kernels: one row, NC consecutive columns starting at col0
// ---------------------------------------------------------------------------------------------
template <int EPI, int NC>
__device__ __forceinline__ void gemm_epilogue(const gemm_params & p, int row, int col0, float (&v)[NC]) {
    if (col0 >= p.N) return;
    const bool full = (col0 + NC <= p.N);
    const int cb = col0 % p.bias_mod;
#pragma unroll
    for (int i = 0; i < NC; i++) {
        float b = 0.f;
        if (p.bias != nullptr && (full || col0 + i < p.N)) b = __ldg(p.bias + cb + i);
        v[i] = v[i] * p.out_scale + b;
    }
    if constexpr (EPI == EPI_STORE) {
        float * o = p.out_f32 + (size_t) row * p.ld_out + col0;
        if (full) {
#pragma unroll
            for (int i = 0; i < NC; i += 4) *(float4 *) (o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        } else {
            for (int i = 0; i < NC && col0 + i < p.N; i++) o[i] = v[i];
        }
    } else if constexpr (EPI == EPI_RESID) {
        float * o = p.out_f32 + (size_t) row * p.ld_out + col0;
        const float * g = nullptr;
        if (p.gate != nullptr) {
            const int u = p.row_utt[row / p.row_div];
            if (u < 0) return;                                 // pad row: leave untouched
            g = p.gate + (size_t) p.gate_voice[u] * p.gate_ld + col0;
        }
        if (full) {
#pragma unroll
            for (int i = 0; i < NC; i += 4) {
                float4 x = *(float4 *) (o + i);
                float4 gg = g ? __ldg((const float4 *) (g + i)) : make_float4(1.f, 1.f, 1.f, 1.f);
                x.x += v[i] * gg.x; x.y += v[i + 1] * gg.y; x.z += v[i + 2] * gg.z; x.w += v[i + 3] * gg.w;
                *(float4 *) (o + i) = x;
            }
        } else {
            for (int i = 0; i < NC && col0 + i < p.N; i++) o[i] += v[i] * (g ? g[i] : 1.f);
        }
    } else if constexpr (EPI == EPI_SWIGLU) {
        // (gate, up) interleaved: h = silu(g) * u  (swiglu_ffn, src/miocodec.cpp:220-225)
        __half hi[NC / 2], lo[NC / 2];
#pragma unroll
        for (int i = 0; i < NC / 2; i++) {
            const float g = v[2 * i], u = v[2 * i + 1];
            const float h = (g / (1.0f + expf(-g))) * u;
            split_half2(h, hi[i], lo[i]);
        }
        __half * oh = p.out_hi + (size_t) row * p.ld_h + col0 / 2;
        __half * ol = p.out_lo + (size_t) row * p.ld_h + col0 / 2;
#pragma unroll
        for (int i = 0; i < NC / 2; i += 4) {
            *(uint2 *) (oh + i) = *(uint2 *) (hi + i);
            *(uint2 *) (ol + i) = *(uint2 *) (lo + i);
        }
    } else if constexpr (EPI == EPI_SPLIT || EPI == EPI_SNAKE) {
        if constexpr (EPI == EPI_SPLIT) {
            if (p.resid != nullptr) {       // resblock tail: h + r, zeros on pad rows (see gemm3.cu)
                const bool pad = p.row_utt[row / p.row_div] < 0;
                for (int i = 0; i < NC && col0 + i < p.N; i++) v[i] = pad ? 0.f : v[i] + p.resid[(size_t) row * p.ld_resid + col0 + i];
            }
        }
        if (p.snake_ea != nullptr) {
            // snake(x) = x + sin^2(e^alpha x) / e^beta  (snake_activation, src/miocodec.cpp:410-420, :717-725)
#pragma unroll
            for (int i = 0; i < NC; i++) {
                if (full || col0 + i < p.N) {
                    const float ea = __ldg(p.snake_ea + cb + i), eb = __ldg(p.snake_eb + cb + i);
                    const float s = sinf(v[i] * ea);
                    v[i] = v[i] + (s * s) / eb;
                }
            }
        }
        if constexpr (EPI == EPI_SNAKE) {
            float * o = p.out_f32 + (size_t) row * p.ld_out + col0;
            if (full) {
#pragma unroll
                for (int i = 0; i < NC; i += 4) *(float4 *) (o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
            } else {
                for (int i = 0; i < NC && col0 + i < p.N; i++) o[i] = v[i];
            }
        } else {
            __half hi[NC], lo[NC];
#pragma unroll
            for (int i = 0; i < NC; i++) split_half2(v[i], hi[i], lo[i]);
            __half * oh = p.out_hi + (size_t) row * p.ld_h + col0;
            __half * ol = p.out_lo + (size_t) row * p.ld_h + col0;
            if (full) {
#pragma unroll
                for (int i = 0; i < NC; i += 4) {
                    *(uint2 *) (oh + i) = *(uint2 *) (hi + i);
                    *(uint2 *) (ol + i) = *(uint2 *) (lo + i);
                }
            } else {
                for (int i = 0; i < NC && col0 + i < p.N; i++) { oh[i] = hi[i]; ol[i] = lo[i]; }
            }
        }
    } else if constexpr (EPI == EPI_QKV) {
        // RoPE (adjacent pairs, utterance-relative position) on q|k, q pre-scaled, split fp16 (see gemm3.cu)
        if (col0 < p.rope_cols) {
            const int u = p.row_utt[row / p.row_div];
            if (u >= 0) {
                const int pos = min(row - p.utt_start[u] * p.row_div, p.rope_max_pos - 1);
                const float qs = (col0 < p.q_cols) ? p.q_scale : 1.0f;
#pragma unroll
                for (int i = 0; i < NC; i += 2) {
                    const float2 cs = __ldg((const float2 *) (p.rope + ((size_t) pos * 32 + (((col0 + i) & 63) >> 1)) * 2));
                    const float x0 = v[i], x1 = v[i + 1];
                    v[i] = (x0 * cs.x - x1 * cs.y) * qs;
                    v[i + 1] = (x0 * cs.y + x1 * cs.x) * qs;
                }
            }
        }
        __half hi[NC], lo[NC];
#pragma unroll
        for (int i = 0; i < NC; i++) split_half2(v[i], hi[i], lo[i]);
        __half * oh = p.out_hi + (size_t) row * p.ld_h + col0;
        __half * ol = p.out_lo + (size_t) row * p.ld_h + col0;
#pragma unroll
        for (int i = 0; i < NC; i += 4) {
            *(uint2 *) (oh + i) = *(uint2 *) (hi + i);
            *(uint2 *) (ol + i) = *(uint2 *) (lo + i);
        }
    } else if constexpr (EPI == EPI_HEAD) {
        // (log-magnitude, phase) interleaved -> (re, im)  (src/miocodec.cpp:732-737)
        float * o = p.out_f32 + (size_t) row * p.ld_out + col0;
#pragma unroll
        for (int i = 0; i < NC; i += 2) {
            if (full || col0 + i < p.N) {
                const float mag = fminf(fmaxf(expf(v[i]), 0.0f), 100.0f);
                float sn, cs;
                sincosf(v[i + 1], &sn, &cs);
                *(float2 *) (o + i) = make_float2(mag * cs, mag * sn);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// SIMT validation path: same operands, same epilogues, FP32 CUDA-core FMAs ((a_hi+a_lo)*(b_hi+b_lo) in full).
// Selected with mio_ctx_set_option("gemm", "simt"); used by the parity tests to separate tensor-core plumbing
// bugs from model bugs.  64x64 tile, 256 threads, 2 rows x 8 columns per thread.
// ---------------------------------------------------------------------------------------------
// BF16-stored matrices keep their bit patterns in the __half buffers (npass 1)
__device__ __forceinline__ float bf16_bits_to_float(__half h) { return __uint_as_float((uint32_t) __half_as_ushort(h) << 16); }

template <int EPI>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const gemm_params p) {
    pdl_sync();
    __shared__ float As[16][64 + 1];
    __shared__ float Bs[16][64 + 1];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    const int tr = (tid >> 3) * 2, tc = (tid & 7) * 8;
    const int Ktot = p.taps * p.K;
    const int shift0 = -(p.taps / 2);
    float acc[2][8];
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[i][j] = 0.f;
    for (int k0 = 0; k0 < Ktot; k0 += 16) {
        for (int e = tid; e < 64 * 16; e += 256) {
            const int r = e >> 4, kk = e & 15;
            const int kg = k0 + kk;
            const int tap = kg / p.K, kc = kg - tap * p.K;
            const int ar = m0 + r + tap + shift0;
            float a = 0.f;
            if (ar >= 0 && ar < p.a_rows) {
                a = p.bf16 ? bf16_bits_to_float(p.a_hi[(size_t) ar * p.lda + kc]) : __half2float(p.a_hi[(size_t) ar * p.lda + kc]);
                if (p.npass == 3) a += __half2float(p.a_lo[(size_t) ar * p.lda + kc]);
            }
            As[kk][r] = a;
            const int bn = n0 + r;
            float b = 0.f;
            if (bn < p.N) {
                b = p.bf16 ? bf16_bits_to_float(p.b_hi[(size_t) bn * Ktot + kg]) : __half2float(p.b_hi[(size_t) bn * Ktot + kg]);
                if (p.npass == 3) b += __half2float(p.b_lo[(size_t) bn * Ktot + kg]);
            }
            Bs[kk][r] = b;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; kk++) {
            float a0 = As[kk][tr], a1 = As[kk][tr + 1];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const float b = Bs[kk][tc + j];
                acc[0][j] += a0 * b;
                acc[1][j] += a1 * b;
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const int row = m0 + tr + i;
        if (row < p.M) gemm_epilogue<EPI, 8>(p, row, n0 + tc, acc[i]);
    }
}

cudaError_t launch_gemm_simt(const gemm_params & p, cudaStream_t st) {
    if (p.M <= 0) return cudaSuccess;
    dim3 grid((p.N + 63) / 64, (p.M + 63) / 64);
    switch (p.epi) {
        case EPI_STORE: launch_k(gemm_simt_kernel<EPI_STORE>, dim3(grid), dim3(256), 0, st, p); break;
        case EPI_RESID: launch_k(gemm_simt_kernel<EPI_RESID>, dim3(grid), dim3(256), 0, st, p); break;
        case EPI_SWIGLU: launch_k(gemm_simt_kernel<EPI_SWIGLU>, dim3(grid), dim3(256), 0, st, p); break;
        case EPI_SPLIT: launch_k(gemm_simt_kernel<EPI_SPLIT>, dim3(grid), dim3(256), 0, st, p); break;
        case EPI_SNAKE: launch_k(gemm_simt_kernel<EPI_SNAKE>, dim3(grid), dim3(256), 0, st, p); break;
        case EPI_HEAD: launch_k(gemm_simt_kernel<EPI_HEAD>, dim3(grid), dim3(256), 0, st, p); break;
        case EPI_QKV: launch_k(gemm_simt_kernel<EPI_QKV>, dim3(grid), dim3(256), 0, st, p); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace mio
