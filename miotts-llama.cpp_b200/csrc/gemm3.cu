// Dense contractions of the decoder (linear layers, conv1d k=3 and ConvTranspose1d as tap-GEMMs) on the sm_100a
// tensor cores: TMA (cp.async.bulk.tensor, SWIZZLE_128B) -> shared memory ring -> tcgen05.mma.cta_group::2.kind::f16
// (FP32 accumulators in TMEM, double buffered) -> tcgen05.ld -> fused epilogue -> TMA store.
//
// Numerics.  The reference computes these layers in FP32 (ggml_mul_mat on F32 weights, src/miocodec.cpp:205-209)
// except conv1d, whose operands it rounds to FP16 itself (ggml.c:4416, ops.cpp:6232-6256).  A single 16-bit pass
// costs 48 dB SNR, so FP32-class layers run as split precision: a = a_hi + a_lo, b = b_hi + b_lo
// (fp16 pairs, ~22 significant bits) and D = a_hi*b_hi + a_hi*b_lo + a_lo*b_hi with FP32 accumulation (npass = 3);
// conv1d layers run one exact-in-kind pass (npass = 1).
//
// Tap-GEMM: D[m][n] = sum_{tap} sum_k A[m + tap - taps/2][k] * B[n][tap*K + k], taps in {1, 3, 5}.  Row shifts are plain TMA
// coordinates; out-of-range rows are zero-filled by TMA and utterances are separated by zero rows, which is the
// zero padding of conv1d / the missing taps of ConvTranspose1d.
//
// Layout of the kernel: one cluster of 2 CTAs per SM pair, static round-robin over 256 x BN output tiles; each CTA
// loads its 128 rows of A and HALF of the B tile; the leader CTA issues the MMAs (M = 256).  The layers have short K
// (128..2048), so a 256 x 256 tile spends only ~6 us in MMAs and the epilogue sets the tile time, hence:
//   * 8 epilogue warps per CTA (two per TMEM lane quadrant, each owning half of the tile's columns)
//   * each warp moves 32 rows x 32 accumulator columns at a time: tcgen05.ld -> registers -> fused math -> swizzled
//     shared-memory staging tile (4 KB per warp, bank-conflict free) -> one TMA store (full 128-byte lines, clipped
//     by the tensor map at the M / N edges, asynchronous)
//   * residual layers (x += gate * y) use the TMA reduce-add (cp.reduce.async.bulk.tensor .add.f32): no load of x
//   * new EPI_QKV: RoPE + 1/sqrt(d) + fp16 hi/lo split fused behind the QKV projection
#include "kernels.cuh"
#include "ptx.cuh"

namespace mio {

constexpr int G3_BK = 64;
constexpr int G3_THREADS = 320;     // warp 0: TMA producer, warp 1: MMA issuer + TMEM owner, warps 2-9: epilogue
constexpr int G3_EPI_WARPS = 8;
// Chunked accumulation (template parameter CHUNK > 0, see the kernel): 384 threads = 3 warpgroups so that setmaxnreg
// can move registers from the producer / MMA warpgroup (warps 0-3, two of them idle) to the epilogue warpgroups
// (warps 4-11), which keep a 256 x BN tile's fp32 running sum in registers: 40 / 232 registers per thread.
constexpr int G3_THREADS_CHUNKED = 384;
constexpr int G3_STAGE_TILE = 4096; // staging bytes per epilogue warp

template <int BN, int NPASS>
struct gemm3_cfg {
    static constexpr int A_BYTES = 128 * G3_BK * 2;                  // this CTA's 128 rows
    static constexpr int B_BYTES = (BN / 2) * G3_BK * 2;             // this CTA's half of the B tile
    static constexpr int NOPS = (NPASS == 3) ? 2 : 1;
    static constexpr int STAGE_BYTES = NOPS * (A_BYTES + B_BYTES);
    static constexpr int STAGES = (196608 / STAGE_BYTES) > 8 ? 8 : (196608 / STAGE_BYTES);
    static constexpr int EPI_BYTES = G3_EPI_WARPS * G3_STAGE_TILE;   // 32 KB
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 + 256;
    static constexpr uint32_t IDESC = (1u << 4) | ((uint32_t) (BN >> 3) << 17) | ((uint32_t) (256 >> 4) << 24);
    static constexpr uint32_t TMEM_COLS = 2 * BN;                    // double-buffered accumulator
};

__device__ __forceinline__ uint32_t pack_split_hi_lo(float a, float b, uint32_t & lo) {
    // same roundings as split_half2 on each value, with the packed converts (one F2FP per pair)
    const __half2 h = __floats2half2_rn(a, b);
    const float2 hf = __half22float2(h);
    const __half2 l = __floats2half2_rn(a - hf.x, b - hf.y);
    lo = *(const uint32_t *) &l;
    return *(const uint32_t *) &h;
}

// fp32 staging tile: [32 rows][128 B], SWIZZLE_128B (16-byte chunk index ^= row & 7)
__device__ __forceinline__ void stage_f32_row(uint8_t * stage, int lane, const float (&v)[32]) {
    float4 * rowp = (float4 *) (stage + lane * 128);
#pragma unroll
    for (int j = 0; j < 8; j++) rowp[j ^ (lane & 7)] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}
// fp16 staging tiles: hi at +0, lo at +2048, each [32 rows][64 B], SWIZZLE_64B (chunk index ^= (row >> 1) & 3).
// Writes outputs [first, first + 2*NP) of the 32-wide row, NP pairs given as packed half2.
template <int NP>
__device__ __forceinline__ void stage_f16_row(uint8_t * stage, int lane, int first_chunk, const uint32_t (&hi)[NP],
                                              const uint32_t (&lo)[NP]) {
    uint4 * rh = (uint4 *) (stage + lane * 64);
    uint4 * rl = (uint4 *) (stage + 2048 + lane * 64);
    const int sw = (lane >> 1) & 3;
#pragma unroll
    for (int j = 0; j < NP / 4; j++) {
        const int c = (first_chunk + j) ^ sw;
        rh[c] = make_uint4(hi[4 * j], hi[4 * j + 1], hi[4 * j + 2], hi[4 * j + 3]);
        rl[c] = make_uint4(lo[4 * j], lo[4 * j + 1], lo[4 * j + 2], lo[4 * j + 3]);
    }
}

// One row (this thread's), 32 accumulator columns starting at col0: fused math + write into the staging tile.
// `part` is the half of an fp16 output row this chunk fills (EPI_SWIGLU only: 32 accumulators -> 16 outputs).
// `aux` is the per-row value the tile loop resolved once, while waiting for the accumulator: the RoPE position of the
// row (EPI_QKV) or the voice whose gate vector applies (gated EPI_RESID); -1 on pad rows / rows past M.
template <int EPI>
__device__ __forceinline__ void epi3_chunk(const gemm_params & p, int row, int col0, float (&v)[32], uint8_t * stage,
                                           int lane, int part, int aux) {
    // Per-column parameter vectors (bias, snake) are readable up to the end of the 32-column chunk (the engine pads
    // them), so there are no per-element bounds checks: columns >= N are computed and then clipped by the TMA store.
    const int cb = col0 % p.bias_mod;
    if (p.bias != nullptr) {
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
            const float4 b = __ldg((const float4 *) (p.bias + cb + i));
            v[i] = v[i] * p.out_scale + b.x; v[i + 1] = v[i + 1] * p.out_scale + b.y;
            v[i + 2] = v[i + 2] * p.out_scale + b.z; v[i + 3] = v[i + 3] * p.out_scale + b.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 32; i++) v[i] *= p.out_scale;
    }
    if constexpr (EPI == EPI_STORE) {
        stage_f32_row(stage, lane, v);
    } else if constexpr (EPI == EPI_RESID) {
        // x += gate[utt] * y: the add itself is done by the TMA reduce; pad rows of gated layers contribute 0
        if (p.gate != nullptr) {
            if (aux < 0) {
#pragma unroll
                for (int i = 0; i < 32; i++) v[i] = 0.f;
            } else {
                const float * g = p.gate + (size_t) aux * p.gate_ld + col0;
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    const float4 gg = __ldg((const float4 *) (g + i));
                    v[i] *= gg.x; v[i + 1] *= gg.y; v[i + 2] *= gg.z; v[i + 3] *= gg.w;
                }
            }
        }
        stage_f32_row(stage, lane, v);
    } else if constexpr (EPI == EPI_SWIGLU) {
        // (gate, up) interleaved: h = silu(g) * u  (swiglu_ffn, src/miocodec.cpp:220-225)
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const float g0 = v[4 * i], u0 = v[4 * i + 1], g1 = v[4 * i + 2], u1 = v[4 * i + 3];
            // libdevice expf; the quotient by 1 + e >= 1 through the SFU reciprocal (<= 2 ulp, below the 2^-22 of the
            // hi/lo split that follows) instead of the IEEE divide sequence: the epilogue has to hide behind the MMAs
            const float h0 = __fdividef(g0, 1.0f + expf(-g0)) * u0;
            const float h1 = __fdividef(g1, 1.0f + expf(-g1)) * u1;
            hi[i] = pack_split_hi_lo(h0, h1, lo[i]);
        }
        stage_f16_row<8>(stage, lane, 2 * part, hi, lo);
    } else if constexpr (EPI == EPI_SPLIT || EPI == EPI_SNAKE) {
        if constexpr (EPI == EPI_SPLIT) {
            if (p.resid != nullptr) {       // resblock tail: h + r (resnet_block, src/miocodec.cpp:403-404); the same fp32 add the TMA reduce did
                if (aux < 0) {
#pragma unroll
                    for (int i = 0; i < 32; i++) v[i] = 0.f;
                } else {
                    const float * r = p.resid + (size_t) row * p.ld_resid + col0;
#pragma unroll
                    for (int i = 0; i < 32; i += 4) {
                        const float4 x = __ldg((const float4 *) (r + i));
                        v[i] += x.x; v[i + 1] += x.y; v[i + 2] += x.z; v[i + 3] += x.w;
                    }
                }
            }
        }
        if (p.snake_ea != nullptr) {
            // snake(x) = x + sin^2(e^alpha x) / e^beta  (snake_activation, src/miocodec.cpp:410-420, :717-725)
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
                const float4 ea = __ldg((const float4 *) (p.snake_ea + cb + i)), eb = __ldg((const float4 *) (p.snake_eb + cb + i));
                float s;
                s = fast_sinf(v[i] * ea.x);         v[i] += __fdividef(s * s, eb.x);
                s = fast_sinf(v[i + 1] * ea.y);     v[i + 1] += __fdividef(s * s, eb.y);
                s = fast_sinf(v[i + 2] * ea.z);     v[i + 2] += __fdividef(s * s, eb.z);
                s = fast_sinf(v[i + 3] * ea.w);     v[i + 3] += __fdividef(s * s, eb.w);
            }
        }
        if constexpr (EPI == EPI_SNAKE) {
            stage_f32_row(stage, lane, v);
        } else {
            uint32_t hi[16], lo[16];
#pragma unroll
            for (int i = 0; i < 16; i++) hi[i] = pack_split_hi_lo(v[2 * i], v[2 * i + 1], lo[i]);
            stage_f16_row<16>(stage, lane, 0, hi, lo);
        }
    } else if constexpr (EPI == EPI_HEAD) {
        // (log-magnitude, phase) interleaved -> (re, im)  (src/miocodec.cpp:732-737)
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
            const float mag = fminf(fmaxf(expf(v[i]), 0.0f), 100.0f);
            float sn, cs;
            fast_sincosf(v[i + 1], &sn, &cs);
            v[i] = mag * cs;
            v[i + 1] = mag * sn;
        }
        stage_f32_row(stage, lane, v);
    } else if constexpr (EPI == EPI_QKV) {
        // RoPE mode 0 (adjacent pairs) on the q and k columns with the utterance-relative position
        // (mha_rope, src/miocodec.cpp:260-263; ggml rope cache ops.cpp:5612-5626); q pre-scaled by 1/sqrt(head_dim)
        if (col0 < p.rope_cols) {
            if (aux >= 0) {
                const int pos = aux;
                // position-minor table: lanes (consecutive rows = consecutive positions) read contiguous float4s
                const float4 * cs = (const float4 *) p.rope_t + (size_t) ((col0 & 63) >> 5) * 8 * p.rope_max_pos + pos;
                const float qs = (col0 < p.q_cols) ? p.q_scale : 1.0f;
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const float4 t = __ldg(cs + (size_t) i * p.rope_max_pos);   // (cos, sin) of two consecutive pairs
                    const float x0 = v[4 * i], x1 = v[4 * i + 1], x2 = v[4 * i + 2], x3 = v[4 * i + 3];
                    v[4 * i] = (x0 * t.x - x1 * t.y) * qs;
                    v[4 * i + 1] = (x0 * t.y + x1 * t.x) * qs;
                    v[4 * i + 2] = (x2 * t.z - x3 * t.w) * qs;
                    v[4 * i + 3] = (x2 * t.w + x3 * t.z) * qs;
                }
            }
        }
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int i = 0; i < 16; i++) hi[i] = pack_split_hi_lo(v[2 * i], v[2 * i + 1], lo[i]);
        stage_f16_row<16>(stage, lane, 0, hi, lo);
    }
}

template <int EPI> struct epi3_out {
    static constexpr bool F16 = (EPI == EPI_SWIGLU || EPI == EPI_SPLIT || EPI == EPI_QKV);
    static constexpr bool REDUCE = (EPI == EPI_RESID);
    static constexpr int ACC_PER_OUT = (EPI == EPI_SWIGLU) ? 2 : 1;   // accumulator columns per output column
};

// CHUNK = 0: the whole K extent accumulates in one TMEM accumulator (double buffered across tiles).
// CHUNK = 1: "chunked accumulation" (chunk length p.chunk k-blocks, a run-time value).  The tensor core adds into its fp32 accumulator with TRUNCATION toward zero, once
// per MMA instruction (profiles/r2a_parity_attribution.md: probes + a signed bias of -2e-6 of the output rms at K = 768,
// -5e-6 at K = 2048, against 5e-7 unbiased for fp32 FMAs).  That bias is what kept the decoder ~0.8 dB under the
// reference's own cross-build parity floor: the layers in front of the first conv1d need fp32-grade sums.  Here every
// p.chunk k-blocks (p.chunk * 64 of K) go to a FRESH TMEM accumulator (the two buffers alternate per chunk instead of per
// tile), the small cross terms (hi*lo, lo*hi) of a k-block are issued before its hi*hi, and the epilogue warps add the
// finished chunk to a running sum in registers with round-to-nearest.  The truncations then act on partial sums of
// short runs whose signs are independent, so the error stops growing with K: ~4e-7 of the output rms for any K.
template <int BN, int NPASS, int EPI, int CHUNK>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(CHUNK > 0 ? G3_THREADS_CHUNKED : G3_THREADS, 1)
gemm_tc3_kernel(const __grid_constant__ CUtensorMap tmA_hi, const __grid_constant__ CUtensorMap tmA_lo,
                const __grid_constant__ CUtensorMap tmB_hi, const __grid_constant__ CUtensorMap tmB_lo,
                const __grid_constant__ CUtensorMap tmO0, const __grid_constant__ CUtensorMap tmO1,
                const gemm_params p) {
    using cfg = gemm3_cfg<BN, NPASS>;
    using eo = epi3_out<EPI>;
    extern __shared__ uint8_t gemm3_smem_raw[];
    const uint32_t raw = smem_u32(gemm3_smem_raw);
    uint8_t * smem = gemm3_smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    uint8_t * epi_smem = smem + cfg::STAGES * cfg::STAGE_BYTES;            // 8 x 4 KB, 1024-aligned
    uint64_t * bars = (uint64_t *) (epi_smem + cfg::EPI_BYTES);
    uint64_t * full = bars;                               // [STAGES]   (used in the leader)
    uint64_t * empty = bars + cfg::STAGES;                // [STAGES]   (both CTAs)
    uint64_t * tmem_full = bars + 2 * cfg::STAGES;        // [2]        (both CTAs)
    uint64_t * tmem_empty = bars + 2 * cfg::STAGES + 2;   // [2]        (used in the leader)
    uint32_t * tmem_slot = (uint32_t *) (bars + 2 * cfg::STAGES + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int EPI_WARP0 = (CHUNK > 0) ? 4 : 2;            // first epilogue warp
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
    const int tiles_n = (p.N + BN - 1) / BN, tiles_m = (p.M + 255) / 256;
    const int num_tiles = tiles_n * tiles_m;
    const int kb_per_tap = p.K / G3_BK;
    const int num_kb = p.taps * kb_per_tap;
    const int shift0 = -(p.taps / 2);                      // taps 1 / 3 / 5: rows m, m-1..m+1, m-2..m+2

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA_hi);
        tma_prefetch_desc(&tmB_hi);
        if (NPASS == 3) { tma_prefetch_desc(&tmA_lo); tma_prefetch_desc(&tmB_lo); }
        tma_prefetch_desc(&tmO0);
        if (eo::F16) tma_prefetch_desc(&tmO1);
        for (int s = 0; s < cfg::STAGES; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; b++) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], 2 * G3_EPI_WARPS); }
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
    // everything above (barriers, TMEM allocation, descriptor prefetch, cluster rendezvous) ran under the predecessor's tail
    pdl_sync();

    if (warp < EPI_WARP0) {
    // chunked variant: this warpgroup (producer, MMA issuer, two idle warps) hands its registers to the epilogue warpgroups
    if constexpr (CHUNK > 0) asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    if (warp == 0) {
        // ===== TMA producer (both CTAs) =====
        if (lane == 0) {
            int it = 0;
            for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
                const int t2 = p.reverse ? num_tiles - 1 - tile : tile;       // serpentine row order between kernels
                const int tm = t2 / tiles_n, tn = t2 - tm * tiles_n;
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
                    tma_load_2d_pair(st, &tmA_hi, lbar, kk * G3_BK, arow);
                    if (NPASS == 3) tma_load_2d_pair(st + cfg::A_BYTES, &tmA_lo, lbar, kk * G3_BK, arow);
                    uint8_t * sb = st + cfg::NOPS * cfg::A_BYTES;
                    tma_load_2d_pair(sb, &tmB_hi, lbar, kb * G3_BK, n0);
                    if (NPASS == 3) tma_load_2d_pair(sb + cfg::B_BYTES, &tmB_lo, lbar, kb * G3_BK, n0);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only) =====
        if (leader && lane == 0) {
            // operand format of kind::f16: bits 7-9 (A) and 10-12 (B), 0 = F16, 1 = BF16 (BF16-stored matrices, npass 1)
            const uint32_t idesc = cfg::IDESC | (p.bf16 ? ((1u << 7) | (1u << 10)) : 0u);
            if constexpr (CHUNK > 0) {
                int it = 0, ci = 0;                               // ci: running chunk counter = accumulator hand-offs
                for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
                    for (int kb0 = 0; kb0 < num_kb; kb0 += p.chunk, ci++) {
                        const int b = ci & 1;
                        const uint32_t aph = (uint32_t) (ci >> 1) & 1u;
                        mbar_wait(&tmem_empty[b], aph ^ 1u);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        const uint32_t d_tmem = tmem_base + (uint32_t) (b * BN);
                        const int nk = (num_kb - kb0 < p.chunk) ? num_kb - kb0 : p.chunk;
                        for (int j = 0; j < nk; j++, it++) {
                            const int s = it % cfg::STAGES;
                            const uint32_t ph = (uint32_t) (it / cfg::STAGES) & 1u;
                            mbar_wait(&full[s], ph);
                            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                            const uint32_t sa = smem_u32(smem + s * cfg::STAGE_BYTES);
                            const uint32_t sb = sa + cfg::NOPS * cfg::A_BYTES;
                            const uint64_t da_hi = umma_desc(sa), db_hi = umma_desc(sb);
                            if constexpr (NPASS == 3) {
                                const uint64_t da_lo = umma_desc(sa + cfg::A_BYTES), db_lo = umma_desc(sb + cfg::B_BYTES);
                                // cross terms first: they are 2^-11 of the hi*hi terms, so their truncations are free
#pragma unroll
                                for (int k = 0; k < G3_BK / 16; k++) umma_f16_pair(d_tmem, da_hi + (uint64_t) (k * 2), db_lo + (uint64_t) (k * 2), idesc, (j > 0 || k > 0) ? 1u : 0u);
#pragma unroll
                                for (int k = 0; k < G3_BK / 16; k++) umma_f16_pair(d_tmem, da_lo + (uint64_t) (k * 2), db_hi + (uint64_t) (k * 2), idesc, 1u);
#pragma unroll
                                for (int k = 0; k < G3_BK / 16; k++) umma_f16_pair(d_tmem, da_hi + (uint64_t) (k * 2), db_hi + (uint64_t) (k * 2), idesc, 1u);
                            } else {
#pragma unroll
                                for (int k = 0; k < G3_BK / 16; k++) umma_f16_pair(d_tmem, da_hi + (uint64_t) (k * 2), db_hi + (uint64_t) (k * 2), idesc, (j > 0 || k > 0) ? 1u : 0u);
                            }
                            umma_commit_pair(&empty[s]);
                        }
                        umma_commit_pair(&tmem_full[b]);
                    }
                }
            } else {
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
                    if (NPASS == 3 && p.mma_order == 2) {          // pass-major inside the k-block: +1.3 % (profiles/r1n_summary.md)
#pragma unroll
                        for (int k = 0; k < G3_BK / 16; k++) umma_f16_pair(d_tmem, da_hi + (uint64_t) (k * 2), db_hi + (uint64_t) (k * 2), idesc, (kb > 0 || k > 0) ? 1u : 0u);
#pragma unroll
                        for (int k = 0; k < G3_BK / 16; k++) umma_f16_pair(d_tmem, da_hi + (uint64_t) (k * 2), db_lo + (uint64_t) (k * 2), idesc, 1u);
#pragma unroll
                        for (int k = 0; k < G3_BK / 16; k++) umma_f16_pair(d_tmem, da_lo + (uint64_t) (k * 2), db_hi + (uint64_t) (k * 2), idesc, 1u);
                    } else {
#pragma unroll
                    for (int k = 0; k < G3_BK / 16; k++) {
                        const uint64_t adv = (uint64_t) (k * 2);
                        umma_f16_pair(d_tmem, da_hi + adv, db_hi + adv, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                        if (NPASS == 3) {
                            if (p.mma_order == 1) {
                                umma_f16_pair(d_tmem, da_lo + adv, db_hi + adv, idesc, 1u);
                                umma_f16_pair(d_tmem, da_hi + adv, db_lo + adv, idesc, 1u);
                            } else {
                                umma_f16_pair(d_tmem, da_hi + adv, db_lo + adv, idesc, 1u);
                                umma_f16_pair(d_tmem, da_lo + adv, db_hi + adv, idesc, 1u);
                            }
                        }
                    }
                    }
                    umma_commit_pair(&empty[s]);
                }
                umma_commit_pair(&tmem_full[b]);
            }
            }
        }
    }
    } else {
        // ===== epilogue warps (both CTAs): TMEM lane quadrant = warp % 4, column half = (warp - EPI_WARP0) / 4 =====
        if constexpr (CHUNK > 0) asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
        const int lg = warp & 3;
        const int half = (warp - EPI_WARP0) >> 2;
        constexpr int CH = BN / 2 / 32;                          // 32-column chunks per warp per tile
        uint8_t * stage = epi_smem + (warp - EPI_WARP0) * G3_STAGE_TILE;
        const uint32_t tmem_empty_leader0 = mapa_u32(smem_u32(&tmem_empty[0]), 0);
        const uint32_t tmem_empty_leader1 = mapa_u32(smem_u32(&tmem_empty[1]), 0);
        int tile_iter = 0;
        [[maybe_unused]] int ci = 0;                             // CHUNK > 0: running chunk counter (accumulator hand-offs)
        for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, tile_iter++) {
            const int t2 = p.reverse ? num_tiles - 1 - tile : tile;
            const int tm = t2 / tiles_n, tn = t2 - tm * tiles_n;
            const int b = tile_iter & 1;
            const uint32_t aph = (uint32_t) (tile_iter >> 1) & 1u;
            const int row0 = tm * 256 + (int) rank * 128 + lg * 32;      // first row of this warp's 32-row slab
            const int row = row0 + lane;
            const int colw = tn * BN + half * (BN / 2);                  // first accumulator column of this warp
            int nch = (p.N - colw + 31) / 32;                            // valid chunks (warp-uniform)
            nch = nch < 0 ? 0 : (nch > CH ? CH : nch);
            if (row0 >= p.M) nch = 0;
            // per-row metadata of this tile, resolved while the MMAs of the tile are still running
            int aux = -1;
            if constexpr (EPI == EPI_QKV) {
                const int u = (row < p.M) ? p.row_utt[row / p.row_div] : -1;
                if (u >= 0) aux = min(row - p.utt_start[u] * p.row_div, p.rope_max_pos - 1);
            } else if constexpr (EPI == EPI_RESID) {
                if (p.gate != nullptr) {
                    const int u = (row < p.M) ? p.row_utt[row / p.row_div] : -1;
                    if (u >= 0) aux = p.gate_voice[u];
                }
            } else if constexpr (EPI == EPI_SPLIT) {
                if (p.resid != nullptr) aux = (row < p.M) ? p.row_utt[row / p.row_div] : -1;
            }
            if constexpr (CHUNK > 0) {
                // running sum of this thread's row: CH pieces of 32 columns, in registers
                float acc[CH][32];
#pragma unroll
                for (int j = 0; j < CH; j++) {
#pragma unroll
                    for (int i = 0; i < 32; i++) acc[j][i] = 0.f;
                }
                const int nchunks = (num_kb + p.chunk - 1) / p.chunk;
                for (int cc = 0; cc < nchunks; cc++, ci++) {
                    const int bb = ci & 1;
                    mbar_wait(&tmem_full[bb], (uint32_t) (ci >> 1) & 1u);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t t_addr = tmem_base + ((uint32_t) (lg * 32) << 16) + (uint32_t) (bb * BN + half * (BN / 2));
#pragma unroll
                    for (int j = 0; j < CH; j++) {
                        if (j < nch) {                                   // warp-uniform
                            float t[32];
                            tmem_ld32(t_addr + (uint32_t) (j * 32), t);
#pragma unroll
                            for (int i = 0; i < 32; i++) acc[j][i] += t[i];       // round-to-nearest, unlike the MMA's adder
                        }
                    }
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive_remote(bb ? tmem_empty_leader1 : tmem_empty_leader0);
                }
#pragma unroll 1
                for (int c = 0; c < nch; c++) {
                    float v[32];
#pragma unroll
                    for (int i = 0; i < 32; i++) v[i] = acc[0][i];
#pragma unroll
                    for (int j = 0; j + 1 < CH; j++) {                   // rotate: the piece index stays static for the compiler
#pragma unroll
                        for (int i = 0; i < 32; i++) acc[j][i] = acc[j + 1][i];
                    }
                    const int part = (eo::ACC_PER_OUT == 2) ? (c & 1) : 0;
                    if (part == 0) {
                        if (lane == 0) tma_store_wait_read0();
                        __syncwarp();
                    }
                    epi3_chunk<EPI>(p, row, colw + c * 32, v, stage, lane, part, aux);
                    const bool flush = (eo::ACC_PER_OUT == 1) || part == 1 || c == nch - 1;
                    if (flush) {
                        fence_async_smem();
                        __syncwarp();
                        if (lane == 0) {
                            const int oc = (eo::ACC_PER_OUT == 2) ? (colw + (c & ~1) * 32) / 2 : colw + c * 32;
                            if (eo::REDUCE) tma_reduce_add_2d(&tmO0, stage, oc, row0);
                            else tma_store_2d(&tmO0, stage, oc, row0);
                            if (eo::F16) tma_store_2d(&tmO1, stage + 2048, oc, row0);
                            tma_store_commit();
                        }
                    }
                }
                continue;
            }
            mbar_wait(&tmem_full[b], aph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t t_addr = tmem_base + ((uint32_t) (lg * 32) << 16) + (uint32_t) (b * BN + half * (BN / 2));
#pragma unroll 1
            for (int c = 0; c < nch; c++) {
                float v[32];
                tmem_ld32(t_addr + (uint32_t) (c * 32), v);
                if (c == nch - 1) {                                      // accumulator drained: hand it back early
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive_remote(b ? tmem_empty_leader1 : tmem_empty_leader0);
                }
                const int part = (eo::ACC_PER_OUT == 2) ? (c & 1) : 0;
                if (part == 0) {                                         // staging tile free again?
                    if (lane == 0) tma_store_wait_read0();
                    __syncwarp();
                }
                epi3_chunk<EPI>(p, row, colw + c * 32, v, stage, lane, part, aux);
                const bool flush = (eo::ACC_PER_OUT == 1) || part == 1 || c == nch - 1;
                if (flush) {
                    fence_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        const int oc = (eo::ACC_PER_OUT == 2) ? (colw + (c & ~1) * 32) / 2 : colw + c * 32;
                        if (eo::REDUCE) tma_reduce_add_2d(&tmO0, stage, oc, row0);
                        else tma_store_2d(&tmO0, stage, oc, row0);
                        if (eo::F16) tma_store_2d(&tmO1, stage + 2048, oc, row0);
                        tma_store_commit();
                    }
                }
            }
            if (nch == 0) {
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive_remote(b ? tmem_empty_leader1 : tmem_empty_leader0);
            }
        }
        if (lane == 0) tma_store_wait_all0();
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(cfg::TMEM_COLS) : "memory");
    }
}

constexpr int G3_CHUNK = 1;   // template flag of the chunked kernels; the chunk length (k-blocks) is the run-time p.chunk

template <int BN, int NPASS, int EPI, int CHUNK>
static cudaError_t launch_tc3_inst(const gemm_params & p, const CUtensorMap & a_hi, const CUtensorMap & a_lo,
                                   const CUtensorMap & b_hi, const CUtensorMap & b_lo, const CUtensorMap & o0,
                                   const CUtensorMap & o1, int num_sms, cudaStream_t st) {
    using cfg = gemm3_cfg<BN, NPASS>;
    static_assert(cfg::SMEM_BYTES <= 232448, "shared memory budget");
    const int tiles = ((p.N + BN - 1) / BN) * ((p.M + 255) / 256);
    int clusters = num_sms / 2;
    if (tiles < clusters) clusters = tiles;
    launch_k(gemm_tc3_kernel<BN, NPASS, EPI, CHUNK>, dim3(2 * clusters), dim3(CHUNK > 0 ? G3_THREADS_CHUNKED : G3_THREADS), cfg::SMEM_BYTES, st, 
        a_hi, a_lo, b_hi, b_lo, o0, o1, p);
    return cudaGetLastError();
}

// chunked accumulation exists for the fp32-class layers in front of the first conv1d (engine::run_chunk sets p.chunk there)
template <int EPI> struct epi3_chunkable {
    static constexpr bool value = (EPI == EPI_STORE || EPI == EPI_RESID || EPI == EPI_SWIGLU || EPI == EPI_SPLIT || EPI == EPI_QKV);
};

template <int BN, int NPASS, int EPI>
static cudaError_t launch_tc3_pick(const gemm_params & p, const CUtensorMap & a_hi, const CUtensorMap & a_lo,
                                   const CUtensorMap & b_hi, const CUtensorMap & b_lo, const CUtensorMap & o0,
                                   const CUtensorMap & o1, int num_sms, cudaStream_t st) {
    if constexpr (NPASS == 3 && epi3_chunkable<EPI>::value) {
        if (p.chunk > 0) return launch_tc3_inst<BN, NPASS, EPI, G3_CHUNK>(p, a_hi, a_lo, b_hi, b_lo, o0, o1, num_sms, st);
    }
    return launch_tc3_inst<BN, NPASS, EPI, 0>(p, a_hi, a_lo, b_hi, b_lo, o0, o1, num_sms, st);
}

template <int BN, int NPASS, int EPI>
static cudaError_t set_attr_inst() {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc3_kernel<BN, NPASS, EPI, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         gemm3_cfg<BN, NPASS>::SMEM_BYTES);
    if constexpr (NPASS == 3 && epi3_chunkable<EPI>::value) {
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(gemm_tc3_kernel<BN, NPASS, EPI, G3_CHUNK>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     gemm3_cfg<BN, NPASS>::SMEM_BYTES);
    }
    return e;
}
template <int BN, int NPASS>
static cudaError_t set_attr_np() {
    cudaError_t e;
    if ((e = set_attr_inst<BN, NPASS, EPI_STORE>()) != cudaSuccess) return e;
    if ((e = set_attr_inst<BN, NPASS, EPI_RESID>()) != cudaSuccess) return e;
    if constexpr (BN >= 128) { if ((e = set_attr_inst<BN, NPASS, EPI_SWIGLU>()) != cudaSuccess) return e; }
    if ((e = set_attr_inst<BN, NPASS, EPI_SPLIT>()) != cudaSuccess) return e;
    if ((e = set_attr_inst<BN, NPASS, EPI_SNAKE>()) != cudaSuccess) return e;
    if ((e = set_attr_inst<BN, NPASS, EPI_HEAD>()) != cudaSuccess) return e;
    return set_attr_inst<BN, NPASS, EPI_QKV>();
}
template <int BN>
static cudaError_t set_attr_bn() {
    cudaError_t e;
    if ((e = set_attr_np<BN, 1>()) != cudaSuccess) return e;
    return set_attr_np<BN, 3>();
}
// The opt-in to > 48 KB of dynamic shared memory is stored per device, so every context sets it for ITS device
// (a process-wide "done" flag would leave the second GPU of a multi-device process without it).
cudaError_t gemm_tc3_init_device() {
    cudaError_t e;
    if ((e = set_attr_bn<256>()) != cudaSuccess) return e;
    if ((e = set_attr_bn<128>()) != cudaSuccess) return e;
    return set_attr_bn<64>();
}

template <int BN, int NPASS>
static cudaError_t launch_tc3_epi(const gemm_params & p, const CUtensorMap & a_hi, const CUtensorMap & a_lo,
                                  const CUtensorMap & b_hi, const CUtensorMap & b_lo, const CUtensorMap & o0,
                                  const CUtensorMap & o1, int num_sms, cudaStream_t st) {
    switch (p.epi) {
        case EPI_STORE: return launch_tc3_pick<BN, NPASS, EPI_STORE>(p, a_hi, a_lo, b_hi, b_lo, o0, o1, num_sms, st);
        case EPI_RESID: return launch_tc3_pick<BN, NPASS, EPI_RESID>(p, a_hi, a_lo, b_hi, b_lo, o0, o1, num_sms, st);
        default: break;
    }
    // NPASS 1 with these epilogues: F16/BF16-stored matrices (the reference rounds the activations to that type and
    // multiplies once, ggml-cpu.c:1291-1326)
    switch (p.epi) {
        case EPI_SWIGLU:
            if (p.N % 64 != 0) return cudaErrorInvalidValue;
            if constexpr (BN >= 128) return launch_tc3_pick<BN, NPASS, EPI_SWIGLU>(p, a_hi, a_lo, b_hi, b_lo, o0, o1, num_sms, st);
            else return cudaErrorInvalidValue;   // a (gate, up) pair row needs 64 accumulator columns per warp
        case EPI_SPLIT: return launch_tc3_pick<BN, NPASS, EPI_SPLIT>(p, a_hi, a_lo, b_hi, b_lo, o0, o1, num_sms, st);
        case EPI_SNAKE: return launch_tc3_pick<BN, NPASS, EPI_SNAKE>(p, a_hi, a_lo, b_hi, b_lo, o0, o1, num_sms, st);
        case EPI_HEAD: return launch_tc3_pick<BN, NPASS, EPI_HEAD>(p, a_hi, a_lo, b_hi, b_lo, o0, o1, num_sms, st);
        case EPI_QKV: return launch_tc3_pick<BN, NPASS, EPI_QKV>(p, a_hi, a_lo, b_hi, b_lo, o0, o1, num_sms, st);
        default: break;
    }
    return cudaErrorInvalidValue;
}

cudaError_t launch_gemm_tc3(const gemm_params & p, const CUtensorMap & tmA_hi, const CUtensorMap & tmA_lo,
                            const CUtensorMap & tmB_hi, const CUtensorMap & tmB_lo, const CUtensorMap & tmO0,
                            const CUtensorMap & tmO1, int bn, int num_sms, cudaStream_t st) {
    if (p.M <= 0) return cudaSuccess;
    if (p.K % G3_BK != 0 || (p.taps != 1 && p.taps != 3 && p.taps != 5) || (p.bf16 && p.npass != 1)) return cudaErrorInvalidValue;
    if (bn == 256) {
        if (p.npass == 3) return launch_tc3_epi<256, 3>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, tmO0, tmO1, num_sms, st);
        return launch_tc3_epi<256, 1>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, tmO0, tmO1, num_sms, st);
    }
    if (bn == 128) {
        if (p.npass == 3) return launch_tc3_epi<128, 3>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, tmO0, tmO1, num_sms, st);
        return launch_tc3_epi<128, 1>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, tmO0, tmO1, num_sms, st);
    }
    if (bn == 64) {   // latency mode: small M, more (narrower) tiles
        if (p.npass == 3) return launch_tc3_epi<64, 3>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, tmO0, tmO1, num_sms, st);
        return launch_tc3_epi<64, 1>(p, tmA_hi, tmA_lo, tmB_hi, tmB_lo, tmO0, tmO1, num_sms, st);
    }
    return cudaErrorInvalidValue;
}

}  // namespace mio
