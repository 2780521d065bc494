// Dense contractions of the decoder (linear layers, conv1d k=3 and ConvTranspose1d as tap-GEMMs) on the sm_100a
// tensor cores: TMA (cp.async.bulk.tensor, SWIZZLE_128B) -> shared memory -> tcgen05.mma (kind::f16, FP32
// accumulators in TMEM) -> tcgen05.ld -> fused epilogue.
//
// Numerics.  The reference computes these layers in FP32 (ggml_mul_mat on F32 weights, src/miocodec.cpp:205-209)
// except conv1d, whose operands it rounds to FP16 itself (ggml.c:4416, ops.cpp:6232-6256).  A single 16-bit pass
// costs 48 dB SNR, so FP32-class layers run as split precision: a = a_hi + a_lo, b = b_hi + b_lo
// (fp16 pairs, ~22 significant bits) and D = a_hi*b_hi + a_hi*b_lo + a_lo*b_hi with FP32 accumulation (npass = 3);
// conv1d layers run one exact-in-kind pass (npass = 1).
//
// Tap-GEMM: D[m][n] = sum_{tap} sum_k A[m + tap - (taps==3)][k] * B[n][tap*K + k].  Row shifts are plain TMA
// coordinates; out-of-range rows are zero-filled by TMA and utterances are separated by zero rows, which is the
// zero padding of conv1d / the missing taps of ConvTranspose1d.
#include "kernels.cuh"
#include "ptx.cuh"

#include <cstdio>

namespace mio {

constexpr int BM = 128, BK = 64;
constexpr int GEMM_THREADS = 192;    // warp 0: TMA producer, warp 1: MMA issuer + TMEM owner, warps 2-5: epilogue

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

template <int BN, int NPASS>
struct gemm_cfg {
    static constexpr int A_BYTES = BM * BK * 2;                      // 16 KB
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int NOPS = (NPASS == 3) ? 2 : 1;                // hi (+ lo) tiles per operand
    static constexpr int STAGE_BYTES = NOPS * (A_BYTES + B_BYTES);
    static constexpr int STAGES = (NPASS == 3) ? (BN == 256 ? 2 : 3) : (BN == 256 ? 4 : 6);
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
    static constexpr uint32_t IDESC = (1u << 4) | ((uint32_t) (BN >> 3) << 17) | ((uint32_t) (BM >> 4) << 24);
};

template <int BN, int NPASS, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA_hi, const __grid_constant__ CUtensorMap tmA_lo,
               const __grid_constant__ CUtensorMap tmB_hi, const __grid_constant__ CUtensorMap tmB_lo,
               const gemm_params p) {
    using cfg = gemm_cfg<BN, NPASS>;
    extern __shared__ uint8_t gemm_smem_raw[];
    const uint32_t raw = smem_u32(gemm_smem_raw);
    uint8_t * smem = gemm_smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    uint64_t * bars = (uint64_t *) (smem + cfg::STAGES * cfg::STAGE_BYTES);
    uint64_t * full = bars;
    uint64_t * empty = bars + cfg::STAGES;
    uint64_t * tmem_full = bars + 2 * cfg::STAGES;
    uint32_t * tmem_slot = (uint32_t *) (bars + 2 * cfg::STAGES + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM;
    const int kb_per_tap = p.K / BK;
    const int num_kb = p.taps * kb_per_tap;
    const int shift0 = (p.taps == 3) ? -1 : 0;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA_hi);
        tma_prefetch_desc(&tmB_hi);
        if (NPASS == 3) { tma_prefetch_desc(&tmA_lo); tma_prefetch_desc(&tmB_lo); }
        for (int s = 0; s < cfg::STAGES; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(tmem_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"((uint32_t) BN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            for (int it = 0; it < num_kb; it++) {
                const int s = it % cfg::STAGES;
                const uint32_t ph = (uint32_t) (it / cfg::STAGES) & 1u;
                mbar_wait(&empty[s], ph ^ 1u);
                uint8_t * st = smem + s * cfg::STAGE_BYTES;
                const int tap = it / kb_per_tap, kk = it - tap * kb_per_tap;
                const int arow = m0 + tap + shift0;
                mbar_expect_tx(&full[s], (uint32_t) cfg::STAGE_BYTES);
                tma_load_2d(st, &tmA_hi, &full[s], kk * BK, arow);
                if (NPASS == 3) tma_load_2d(st + cfg::A_BYTES, &tmA_lo, &full[s], kk * BK, arow);
                uint8_t * sb = st + cfg::NOPS * cfg::A_BYTES;
                tma_load_2d(sb, &tmB_hi, &full[s], it * BK, n0);
                if (NPASS == 3) tma_load_2d(sb + cfg::B_BYTES, &tmB_lo, &full[s], it * BK, n0);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            for (int it = 0; it < num_kb; it++) {
                const int s = it % cfg::STAGES;
                const uint32_t ph = (uint32_t) (it / cfg::STAGES) & 1u;
                mbar_wait(&full[s], ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t sa = smem_u32(smem + s * cfg::STAGE_BYTES);
                const uint32_t sb = sa + cfg::NOPS * cfg::A_BYTES;
                const uint64_t da_hi = umma_desc(sa), db_hi = umma_desc(sb);
                const uint64_t da_lo = umma_desc(sa + cfg::A_BYTES), db_lo = umma_desc(sb + cfg::B_BYTES);
#pragma unroll
                for (int k = 0; k < BK / 16; k++) {
                    const uint64_t adv = (uint64_t) (k * 2);            // 32 bytes >> 4
                    umma_f16(tmem_base, da_hi + adv, db_hi + adv, cfg::IDESC, (it > 0 || k > 0) ? 1u : 0u);
                    if (NPASS == 3) {
                        umma_f16(tmem_base, da_hi + adv, db_lo + adv, cfg::IDESC, 1u);
                        umma_f16(tmem_base, da_lo + adv, db_hi + adv, cfg::IDESC, 1u);
                    }
                }
                umma_commit(&empty[s]);          // frees the smem stage when these MMAs retire
            }
            umma_commit(tmem_full);              // accumulator complete
        }
    } else {
        // ===== epilogue warps: TMEM lane group = warp % 4 =====
        const int lg = warp & 3;
        mbar_wait(tmem_full, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int row = m0 + lg * 32 + lane;
#pragma unroll 1
        for (int c = 0; c < BN / 32; c++) {
            if (n0 + c * 32 >= p.N) break;       // warp-uniform
            float v[32];
            tmem_ld32(tmem_base + ((uint32_t) (lg * 32) << 16) + (uint32_t) (c * 32), v);
            if (row < p.M) gemm_epilogue<EPI, 32>(p, row, n0 + c * 32, v);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t) BN) : "memory");
    }
}

template <int BN, int NPASS, int EPI>
static cudaError_t launch_tc_inst(const gemm_params & p, const CUtensorMap & a_hi, const CUtensorMap & a_lo,
                                  const CUtensorMap & b_hi, const CUtensorMap & b_lo, cudaStream_t st) {
    using cfg = gemm_cfg<BN, NPASS>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<BN, NPASS, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             cfg::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    dim3 grid((p.N + BN - 1) / BN, (p.M + BM - 1) / BM);
    gemm_tc_kernel<BN, NPASS, EPI><<<grid, GEMM_THREADS, cfg::SMEM_BYTES, st>>>(a_hi, a_lo, b_hi, b_lo, p);
    return cudaGetLastError();
}

template <int BN, int NPASS>
static cudaError_t launch_tc_epi(const gemm_params & p, const CUtensorMap & a_hi, const CUtensorMap & a_lo,
                                 const CUtensorMap & b_hi, const CUtensorMap & b_lo, cudaStream_t st) {
    switch (p.epi) {
        case EPI_STORE: return launch_tc_inst<BN, NPASS, EPI_STORE>(p, a_hi, a_lo, b_hi, b_lo, st);
        case EPI_RESID: return launch_tc_inst<BN, NPASS, EPI_RESID>(p, a_hi, a_lo, b_hi, b_lo, st);
        default: break;
    }
    if constexpr (NPASS == 3) {
        switch (p.epi) {
            case EPI_SWIGLU: return launch_tc_inst<BN, 3, EPI_SWIGLU>(p, a_hi, a_lo, b_hi, b_lo, st);
            case EPI_SPLIT: return launch_tc_inst<BN, 3, EPI_SPLIT>(p, a_hi, a_lo, b_hi, b_lo, st);
            case EPI_SNAKE: return launch_tc_inst<BN, 3, EPI_SNAKE>(p, a_hi, a_lo, b_hi, b_lo, st);
            case EPI_HEAD: return launch_tc_inst<BN, 3, EPI_HEAD>(p, a_hi, a_lo, b_hi, b_lo, st);
            case EPI_QKV: return launch_tc_inst<BN, 3, EPI_QKV>(p, a_hi, a_lo, b_hi, b_lo, st);
            default: break;
        }
    }
    return cudaErrorInvalidValue;
}

cudaError_t launch_gemm_tc(const gemm_params & p, const CUtensorMap & tmA_hi, const CUtensorMap & tmA_lo,
                           const CUtensorMap & tmB_hi, const CUtensorMap & tmB_lo, int bn, cudaStream_t st) {
    if (p.M <= 0) return cudaSuccess;
    if (p.K % BK != 0 || (p.taps != 1 && p.taps != 3)) return cudaErrorInvalidValue;
    if (bn == 128) {
        if (p.npass == 3) return launch_tc_epi<128, 3>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, st);
        return launch_tc_epi<128, 1>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, st);
    }
    if (bn == 256) {
        if (p.npass == 3) return launch_tc_epi<256, 3>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, st);
        return launch_tc_epi<256, 1>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, st);
    }
    return cudaErrorInvalidValue;
}

cudaError_t gemm_tc_init() { return cudaSuccess; }

// =============================================================================================
// Generation 2: persistent, CTA-pair (cta_group::2) tap-GEMM.
//   * one cluster of 2 CTAs per SM pair, 74 clusters, static round-robin over 256 x BN output tiles
//   * each CTA loads its 128 rows of A and HALF of the B tile (BN/2 rows): 64 KB per k-block instead of 96 KB for
//     a 128x256 single-CTA tile -> fits 3 stages and cuts L2->SM operand traffic per MMA by a third
//   * the leader CTA issues tcgen05.mma.cta_group::2 (M = 256); accumulators are double buffered in TMEM
//     (2 x BN columns) so the epilogue of tile i overlaps the main loop of tile i+1
// Barriers: full[s] lives in the leader (both CTAs' TMA loads complete_tx on it), empty[s] / tmem_full[b] are
// multicast by tcgen05.commit to both CTAs, tmem_empty[b] lives in the leader and counts the 8 epilogue warps.
// =============================================================================================
template <int BN, int NPASS>
struct gemm2_cfg {
    static constexpr int A_BYTES = 128 * BK * 2;                     // this CTA's 128 rows
    static constexpr int B_BYTES = (BN / 2) * BK * 2;                // this CTA's half of the B tile
    static constexpr int NOPS = (NPASS == 3) ? 2 : 1;
    static constexpr int STAGE_BYTES = NOPS * (A_BYTES + B_BYTES);
    static constexpr int STAGES = (196608 / STAGE_BYTES) > 8 ? 8 : (196608 / STAGE_BYTES);
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
    static constexpr uint32_t IDESC = (1u << 4) | ((uint32_t) (BN >> 3) << 17) | ((uint32_t) (256 >> 4) << 24);
    static constexpr uint32_t TMEM_COLS = 2 * BN;                    // double-buffered accumulator
};

template <int BN, int NPASS, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmA_hi, const __grid_constant__ CUtensorMap tmA_lo,
                const __grid_constant__ CUtensorMap tmB_hi, const __grid_constant__ CUtensorMap tmB_lo,
                const gemm_params p) {
    using cfg = gemm2_cfg<BN, NPASS>;
    extern __shared__ uint8_t gemm_smem_raw[];
    const uint32_t raw = smem_u32(gemm_smem_raw);
    uint8_t * smem = gemm_smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    uint64_t * bars = (uint64_t *) (smem + cfg::STAGES * cfg::STAGE_BYTES);
    uint64_t * full = bars;                               // [STAGES]   (used in the leader)
    uint64_t * empty = bars + cfg::STAGES;                // [STAGES]   (both CTAs)
    uint64_t * tmem_full = bars + 2 * cfg::STAGES;        // [2]        (both CTAs)
    uint64_t * tmem_empty = bars + 2 * cfg::STAGES + 2;   // [2]        (used in the leader)
    uint32_t * tmem_slot = (uint32_t *) (bars + 2 * cfg::STAGES + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
    const int tiles_n = (p.N + BN - 1) / BN, tiles_m = (p.M + 255) / 256;
    const int num_tiles = tiles_n * tiles_m;
    const int kb_per_tap = p.K / BK;
    const int num_kb = p.taps * kb_per_tap;
    const int shift0 = (p.taps == 3) ? -1 : 0;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA_hi);
        tma_prefetch_desc(&tmB_hi);
        if (NPASS == 3) { tma_prefetch_desc(&tmA_lo); tma_prefetch_desc(&tmB_lo); }
        for (int s = 0; s < cfg::STAGES; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; b++) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(cfg::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer (both CTAs) =====
        if (lane == 0) {
            int it = 0;
            for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
                const int tm = tile / tiles_n, tn = tile - tm * tiles_n;
                const int m0 = tm * 256 + (int) rank * 128, n0 = tn * BN + (int) rank * (BN / 2);
                for (int kb = 0; kb < num_kb; kb++, it++) {
                    const int s = it % cfg::STAGES;
                    const uint32_t ph = (uint32_t) (it / cfg::STAGES) & 1u;
                    mbar_wait(&empty[s], ph ^ 1u);
                    const uint32_t lbar = mapa_u32(smem_u32(&full[s]), 0);
                    if (leader) mbar_expect_tx(&full[s], 2u * (uint32_t) cfg::STAGE_BYTES);
                    uint8_t * st = smem + s * cfg::STAGE_BYTES;
                    const int tap = kb / kb_per_tap, kk = kb - tap * kb_per_tap;
                    const int arow = m0 + tap + shift0;
                    tma_load_2d_pair(st, &tmA_hi, lbar, kk * BK, arow);
                    if (NPASS == 3) tma_load_2d_pair(st + cfg::A_BYTES, &tmA_lo, lbar, kk * BK, arow);
                    uint8_t * sb = st + cfg::NOPS * cfg::A_BYTES;
                    tma_load_2d_pair(sb, &tmB_hi, lbar, kb * BK, n0);
                    if (NPASS == 3) tma_load_2d_pair(sb + cfg::B_BYTES, &tmB_lo, lbar, kb * BK, n0);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only) =====
        if (leader && lane == 0) {
            int it = 0, tile_iter = 0;
            for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, tile_iter++) {
                const int b = tile_iter & 1;
                const uint32_t aph = (uint32_t) (tile_iter >> 1) & 1u;
                mbar_wait(&tmem_empty[b], aph ^ 1u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t d_tmem = tmem_base + (uint32_t) (b * BN);
                for (int kb = 0; kb < num_kb; kb++, it++) {
                    const int s = it % cfg::STAGES;
                    const uint32_t ph = (uint32_t) (it / cfg::STAGES) & 1u;
                    mbar_wait(&full[s], ph);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = smem_u32(smem + s * cfg::STAGE_BYTES);
                    const uint32_t sb = sa + cfg::NOPS * cfg::A_BYTES;
                    const uint64_t da_hi = umma_desc(sa), db_hi = umma_desc(sb);
                    const uint64_t da_lo = umma_desc(sa + cfg::A_BYTES), db_lo = umma_desc(sb + cfg::B_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; k++) {
                        const uint64_t adv = (uint64_t) (k * 2);
                        umma_f16_pair(d_tmem, da_hi + adv, db_hi + adv, cfg::IDESC, (kb > 0 || k > 0) ? 1u : 0u);
                        if (NPASS == 3) {
                            umma_f16_pair(d_tmem, da_hi + adv, db_lo + adv, cfg::IDESC, 1u);
                            umma_f16_pair(d_tmem, da_lo + adv, db_hi + adv, cfg::IDESC, 1u);
                        }
                    }
                    umma_commit_pair(&empty[s]);
                }
                umma_commit_pair(&tmem_full[b]);
            }
        }
    } else {
        // ===== epilogue warps (both CTAs): TMEM lane group = warp % 4 =====
        const int lg = warp & 3;
        int tile_iter = 0;
        for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, tile_iter++) {
            const int tm = tile / tiles_n, tn = tile - tm * tiles_n;
            const int b = tile_iter & 1;
            const uint32_t aph = (uint32_t) (tile_iter >> 1) & 1u;
            mbar_wait(&tmem_full[b], aph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int row = tm * 256 + (int) rank * 128 + lg * 32 + lane;
            const int n0 = tn * BN;
#pragma unroll 1
            for (int c = 0; c < BN / 32; c++) {
                if (n0 + c * 32 >= p.N) break;
                float v[32];
                tmem_ld32(tmem_base + ((uint32_t) (lg * 32) << 16) + (uint32_t) (b * BN + c * 32), v);
                if (row < p.M) gemm_epilogue<EPI, 32>(p, row, n0 + c * 32, v);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive_remote(mapa_u32(smem_u32(&tmem_empty[b]), 0));
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(cfg::TMEM_COLS) : "memory");
    }
}

static int g_num_sms = 0;

template <int BN, int NPASS, int EPI>
static cudaError_t launch_tc2_inst(const gemm_params & p, const CUtensorMap & a_hi, const CUtensorMap & a_lo,
                                   const CUtensorMap & b_hi, const CUtensorMap & b_lo, cudaStream_t st) {
    using cfg = gemm2_cfg<BN, NPASS>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(gemm_tc2_kernel<BN, NPASS, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             cfg::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    if (g_num_sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const int tiles = ((p.N + BN - 1) / BN) * ((p.M + 255) / 256);
    int clusters = g_num_sms / 2;
    if (tiles < clusters) clusters = tiles;
    gemm_tc2_kernel<BN, NPASS, EPI><<<2 * clusters, GEMM_THREADS, cfg::SMEM_BYTES, st>>>(a_hi, a_lo, b_hi, b_lo, p);
    return cudaGetLastError();
}

template <int BN, int NPASS>
static cudaError_t launch_tc2_epi(const gemm_params & p, const CUtensorMap & a_hi, const CUtensorMap & a_lo,
                                  const CUtensorMap & b_hi, const CUtensorMap & b_lo, cudaStream_t st) {
    switch (p.epi) {
        case EPI_STORE: return launch_tc2_inst<BN, NPASS, EPI_STORE>(p, a_hi, a_lo, b_hi, b_lo, st);
        case EPI_RESID: return launch_tc2_inst<BN, NPASS, EPI_RESID>(p, a_hi, a_lo, b_hi, b_lo, st);
        default: break;
    }
    if constexpr (NPASS == 3) {
        switch (p.epi) {
            case EPI_SWIGLU: return launch_tc2_inst<BN, 3, EPI_SWIGLU>(p, a_hi, a_lo, b_hi, b_lo, st);
            case EPI_SPLIT: return launch_tc2_inst<BN, 3, EPI_SPLIT>(p, a_hi, a_lo, b_hi, b_lo, st);
            case EPI_SNAKE: return launch_tc2_inst<BN, 3, EPI_SNAKE>(p, a_hi, a_lo, b_hi, b_lo, st);
            case EPI_HEAD: return launch_tc2_inst<BN, 3, EPI_HEAD>(p, a_hi, a_lo, b_hi, b_lo, st);
            case EPI_QKV: return launch_tc2_inst<BN, 3, EPI_QKV>(p, a_hi, a_lo, b_hi, b_lo, st);
            default: break;
        }
    }
    return cudaErrorInvalidValue;
}

// tmB_*: box {64, bn/2}
cudaError_t launch_gemm_tc2(const gemm_params & p, const CUtensorMap & tmA_hi, const CUtensorMap & tmA_lo,
                            const CUtensorMap & tmB_hi, const CUtensorMap & tmB_lo, int bn, cudaStream_t st) {
    if (p.M <= 0) return cudaSuccess;
    if (p.K % BK != 0 || (p.taps != 1 && p.taps != 3)) return cudaErrorInvalidValue;
    if (bn == 256) {
        if (p.npass == 3) return launch_tc2_epi<256, 3>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, st);
        return launch_tc2_epi<256, 1>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, st);
    }
    if (bn == 128) {
        if (p.npass == 3) return launch_tc2_epi<128, 3>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, st);
        return launch_tc2_epi<128, 1>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, st);
    }
    return cudaErrorInvalidValue;
}


// ---------------------------------------------------------------------------------------------
// SIMT validation path: same operands, same epilogues, FP32 CUDA-core FMAs ((a_hi+a_lo)*(b_hi+b_lo) in full).
// Selected with mio_ctx_set_option("gemm", "simt"); used by the parity tests to separate tensor-core plumbing
// bugs from model bugs.  64x64 tile, 256 threads, 2 rows x 8 columns per thread.
// ---------------------------------------------------------------------------------------------
template <int EPI>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const gemm_params p) {
    __shared__ float As[16][64 + 1];
    __shared__ float Bs[16][64 + 1];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    const int tr = (tid >> 3) * 2, tc = (tid & 7) * 8;
    const int Ktot = p.taps * p.K;
    const int shift0 = (p.taps == 3) ? -1 : 0;
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
                a = __half2float(p.a_hi[(size_t) ar * p.lda + kc]);
                if (p.npass == 3) a += __half2float(p.a_lo[(size_t) ar * p.lda + kc]);
            }
            As[kk][r] = a;
            const int bn = n0 + r;
            float b = 0.f;
            if (bn < p.N) {
                b = __half2float(p.b_hi[(size_t) bn * Ktot + kg]);
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
        case EPI_STORE: gemm_simt_kernel<EPI_STORE><<<grid, 256, 0, st>>>(p); break;
        case EPI_RESID: gemm_simt_kernel<EPI_RESID><<<grid, 256, 0, st>>>(p); break;
        case EPI_SWIGLU: gemm_simt_kernel<EPI_SWIGLU><<<grid, 256, 0, st>>>(p); break;
        case EPI_SPLIT: gemm_simt_kernel<EPI_SPLIT><<<grid, 256, 0, st>>>(p); break;
        case EPI_SNAKE: gemm_simt_kernel<EPI_SNAKE><<<grid, 256, 0, st>>>(p); break;
        case EPI_HEAD: gemm_simt_kernel<EPI_HEAD><<<grid, 256, 0, st>>>(p); break;
        case EPI_QKV: gemm_simt_kernel<EPI_QKV><<<grid, 256, 0, st>>>(p); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

}  // namespace mio
