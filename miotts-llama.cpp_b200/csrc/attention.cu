// Banded local attention (|i-j| <= 32, head_dim 64, per-utterance boundaries) on operands that the QKV projection's
// epilogue (EPI_QKV, gemm3.cu) already rotated (RoPE), pre-scaled (q / sqrt(64)) and split into fp16 hi/lo pairs.
// Replaces the score/softmax/PV part of mha_rope (src/miocodec.cpp:264-279 of the reference; the reference
// materialises dense S x S scores with a -inf mask, :233-242 — masked terms are exactly 0 after softmax, so the band is
// mathematically identical).
//
// This stage is HBM-bound (8*D bytes read + 4*D written per row against ~1 kFLOP/byte of tensor work), so the kernel
// is built around data movement:
//   * one block = 64 queries x 1 head; 8 KB TMA box loads (q, k as hi and lo; keys [q0-32, q0+96)) land in
//     SWIZZLE_128B shared memory behind one mbarrier; rows outside [0, M) are zero-filled by the TMA unit.  V is
//     loaded into K's tiles as soon as every warp has finished S = Q K^T, so the load hides behind the softmax and a
//     block needs 48 KB instead of 80 KB: four blocks (16 warps) per SM instead of two
//   * fragments come from ldmatrix (ldmatrix.trans for V), bank-conflict free under the 128-byte swizzle
//   * S = Q K^T and O = P V are warp-level m16n8k16 MMAs in split precision (hi*hi + hi*lo + lo*hi, fp32 accumulate);
//     each warp owns 16 queries and the 80 keys covering their bands; softmax runs on the accumulator fragments
//     (valid key range per query precomputed by attn_bounds_kernel; exp via ex2.approx on log2e-scaled scores)
//   * the result is split to fp16 hi/lo in the warp's own (dead) Q rows and written with two TMA stores
#include "kernels.cuh"
#include "ptx.cuh"

namespace mio {

constexpr int AT_QB = 64;                       // queries per block
constexpr int AT_KR = 128;                      // key rows staged
constexpr int AT_TILE = 64 * 128;               // bytes of one 64-row x 64-half box
constexpr int AT_SMEM = 6 * AT_TILE + 1024 + 64;

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// (x0, x1) -> packed half2 hi and lo with x = hi + lo (same roundings as split_half2, packed converts)
__device__ __forceinline__ void split_pack2(float x0, float x1, uint32_t & hi, uint32_t & lo) {
    const __half2 h = __floats2half2_rn(x0, x1);
    const float2 hf = __half22float2(h);
    const __half2 l = __floats2half2_rn(x0 - hf.x, x1 - hf.y);
    hi = *(const uint32_t *) &h;
    lo = *(const uint32_t *) &l;
}
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// byte offset of (row, 16-byte chunk) inside a [rows][128 B] SWIZZLE_128B tile whose base is 1024-byte aligned
__device__ __forceinline__ uint32_t sw128(int row, int chunk) { return (uint32_t) (row * 128 + ((chunk ^ (row & 7)) << 4)); }

// Valid key range of every query row: keys j with |i - j| <= half_w inside the row's own utterance
// (fill_local_attn_mask, src/miocodec.cpp:233-242, plus the utterance boundaries of the ragged batch).
// bounds[i] = (first key, number of keys); pad rows get (0, 0).  Computed once per chunk and resolution, reused by
// every layer.
__global__ void attn_bounds_kernel(const int32_t * __restrict__ row_utt, const int32_t * __restrict__ start,
                                   const int32_t * __restrict__ len, int rows, int row_div, int half_w, int2 * __restrict__ bounds) {
    pdl_sync();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows) return;
    const int u = row_utt[i / row_div];
    int2 b = make_int2(0, 0);
    if (u >= 0) {
        const int us = start[u] * row_div, ue = us + len[u] * row_div;
        const int lo = max(us, i - half_w), hi = min(ue - 1, i + half_w);
        b = make_int2(lo, hi - lo + 1);
    }
    bounds[i] = b;
}

cudaError_t launch_attn_bounds(const row_map & rm, int rows, int row_div, int window, int2 * bounds, cudaStream_t st) {
    if (rows <= 0) return cudaSuccess;
    launch_k(attn_bounds_kernel, dim3((rows + 255) / 256), dim3(256), 0, st, rm.row_utt, rm.start, rm.len, rows, row_div, window / 2, bounds);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(128, 4) attention_tc_kernel(
        const __grid_constant__ CUtensorMap tm_in_hi, const __grid_constant__ CUtensorMap tm_in_lo,
        const __grid_constant__ CUtensorMap tm_out_hi, const __grid_constant__ CUtensorMap tm_out_lo,
        int rows, int D, const int2 * __restrict__ bounds, int rev) {
    pdl_sync();
    extern __shared__ uint8_t at_smem_raw[];
    const uint32_t raw = smem_u32(at_smem_raw);
    uint8_t * smem = at_smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    // tiles: 0 Qh, 1 Ql, 2-3 Kh then Vh, 4-5 Kl then Vl
    uint64_t * bar = (uint64_t *) (smem + 6 * AT_TILE);     // [0]: q + k landed, [1]: v landed
    const int q0 = (rev ? gridDim.x - 1 - blockIdx.x : blockIdx.x) * AT_QB;
    const int h = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k_lo = q0 - 32;

    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(&bar[0], 6u * AT_TILE);
        const int cq = h * 64, ck = D + h * 64;
        tma_load_2d(smem + 0 * AT_TILE, &tm_in_hi, &bar[0], cq, q0);
        tma_load_2d(smem + 1 * AT_TILE, &tm_in_lo, &bar[0], cq, q0);
        tma_load_2d(smem + 2 * AT_TILE, &tm_in_hi, &bar[0], ck, k_lo);
        tma_load_2d(smem + 3 * AT_TILE, &tm_in_hi, &bar[0], ck, k_lo + 64);
        tma_load_2d(smem + 4 * AT_TILE, &tm_in_lo, &bar[0], ck, k_lo);
        tma_load_2d(smem + 5 * AT_TILE, &tm_in_lo, &bar[0], ck, k_lo + 64);
    }
    // per-thread row metadata while the loads fly: rows i0 = g, i1 = g + 8 of the warp's 16 queries
    const int g = lane >> 2, t = lane & 3;
    const int qr = warp * 16;
    const int kw = warp * 16;                       // first key row (staged index) of this warp's 80-key window
    const int i0 = q0 + qr + g, i1 = i0 + 8;
    const int2 b0 = (i0 < rows) ? __ldg(bounds + i0) : make_int2(0, 0);
    const int2 b1 = (i1 < rows) ? __ldg(bounds + i1) : make_int2(0, 0);
    __syncthreads();                                // barriers initialised
    mbar_wait(&bar[0], 0);

    const uint32_t sQh = smem_u32(smem), sQl = sQh + AT_TILE, sKh = sQh + 2 * AT_TILE, sKl = sQh + 4 * AT_TILE;
    // ---- S = Q K^T over 10 key tiles of 8
    float s[10][4];
#pragma unroll
    for (int nt = 0; nt < 10; nt++) { s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f; }
#pragma unroll
    for (int ks = 0; ks < 4; ks++) {
        uint32_t ah[4], al[4];
        {
            const int r = qr + (lane & 7) + ((lane >> 3) & 1) * 8, c = 2 * ks + (lane >> 4);
            ldsm_x4(ah, sQh + sw128(r, c));
            ldsm_x4(al, sQl + sw128(r, c));
        }
        // all ten key tiles of this k-step first, then three passes over ten independent accumulators (the three
        // MMAs that feed one accumulator are ten instructions apart, so the tensor pipe never waits on a dependency)
        uint32_t bh[5][4], bl[5][4];
#pragma unroll
        for (int np = 0; np < 5; np++) {            // pairs of key tiles
            const int r = kw + np * 16 + (lane & 7) + ((lane >> 4) & 1) * 8, c = 2 * ks + ((lane >> 3) & 1);
            ldsm_x4(bh[np], sKh + sw128(r, c));
            ldsm_x4(bl[np], sKl + sw128(r, c));
        }
#pragma unroll
        for (int np = 0; np < 5; np++) { mma16816(s[2 * np], al, bh[np][0], bh[np][1]); mma16816(s[2 * np + 1], al, bh[np][2], bh[np][3]); }
#pragma unroll
        for (int np = 0; np < 5; np++) { mma16816(s[2 * np], ah, bl[np][0], bl[np][1]); mma16816(s[2 * np + 1], ah, bl[np][2], bl[np][3]); }
#pragma unroll
        for (int np = 0; np < 5; np++) { mma16816(s[2 * np], ah, bh[np][0], bh[np][1]); mma16816(s[2 * np + 1], ah, bh[np][2], bh[np][3]); }
    }
    // ---- every warp is done with K: V replaces it (the load overlaps the softmax below)
    __syncthreads();
    if (threadIdx.x == 0) {
        fence_async_smem();
        mbar_expect_tx(&bar[1], 4u * AT_TILE);
        const int cv = 2 * D + h * 64;
        tma_load_2d(smem + 2 * AT_TILE, &tm_in_hi, &bar[1], cv, k_lo);
        tma_load_2d(smem + 3 * AT_TILE, &tm_in_hi, &bar[1], cv, k_lo + 64);
        tma_load_2d(smem + 4 * AT_TILE, &tm_in_lo, &bar[1], cv, k_lo);
        tma_load_2d(smem + 5 * AT_TILE, &tm_in_lo, &bar[1], cv, k_lo + 64);
    }
    // ---- masked softmax on the fragments (softmax over keys: subtract row max, exp, normalise; ops.cpp:5286-5323).
    // Key j of element (nt, e) is j = k_lo + kw + nt*8 + 2t + e; it is valid iff (unsigned)(j - first) < count.
    const float L2E = 1.4426950408889634f;
    const int rel0 = k_lo + kw + 2 * t - b0.x, rel1 = k_lo + kw + 2 * t - b1.x;
    float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 10; nt++) {
#pragma unroll
        for (int e = 0; e < 2; e++) {
            const bool ok0 = (unsigned) (rel0 + nt * 8 + e) < (unsigned) b0.y;
            const bool ok1 = (unsigned) (rel1 + nt * 8 + e) < (unsigned) b1.y;
            s[nt][e] = ok0 ? s[nt][e] : -INFINITY;
            s[nt][2 + e] = ok1 ? s[nt][2 + e] : -INFINITY;
            m0 = fmaxf(m0, s[nt][e]);
            m1 = fmaxf(m1, s[nt][2 + e]);
        }
    }
    m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
    m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
    // pad rows have no valid key (max = -inf): use 0 so that every term is ex2(-inf) = 0 instead of NaN
    const float c0 = (b0.y > 0) ? -m0 * L2E : 0.f, c1 = (b1.y > 0) ? -m1 * L2E : 0.f;
    float l0 = 0.f, l1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 10; nt++) {
#pragma unroll
        for (int e = 0; e < 2; e++) {
            const float p0 = ex2_approx(fmaf(s[nt][e], L2E, c0));
            const float p1 = ex2_approx(fmaf(s[nt][2 + e], L2E, c1));
            s[nt][e] = p0; s[nt][2 + e] = p1;
            l0 += p0; l1 += p1;
        }
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float inv0 = (b0.y > 0) ? 1.0f / l0 : 0.f, inv1 = (b1.y > 0) ? 1.0f / l1 : 0.f;
    // ---- O = P V over 5 key steps of 16, 8 d-tiles of 8
    const uint32_t sVh = sKh, sVl = sKl;
    float o[8][4];
#pragma unroll
    for (int nd = 0; nd < 8; nd++) { o[nd][0] = o[nd][1] = o[nd][2] = o[nd][3] = 0.f; }
    uint32_t ph[5][4], pl[5][4];
#pragma unroll
    for (int ks = 0; ks < 5; ks++) {
        split_pack2(s[2 * ks][0] * inv0, s[2 * ks][1] * inv0, ph[ks][0], pl[ks][0]);
        split_pack2(s[2 * ks][2] * inv1, s[2 * ks][3] * inv1, ph[ks][1], pl[ks][1]);
        split_pack2(s[2 * ks + 1][0] * inv0, s[2 * ks + 1][1] * inv0, ph[ks][2], pl[ks][2]);
        split_pack2(s[2 * ks + 1][2] * inv1, s[2 * ks + 1][3] * inv1, ph[ks][3], pl[ks][3]);
    }
    mbar_wait(&bar[1], 0);
#pragma unroll
    for (int ks = 0; ks < 5; ks++) {
        uint32_t bh[4][4], bl[4][4];
#pragma unroll
        for (int dp = 0; dp < 4; dp++) {            // pairs of d tiles
            const int r = kw + ks * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, c = 2 * dp + (lane >> 4);
            ldsm_x4_t(bh[dp], sVh + sw128(r, c));
            ldsm_x4_t(bl[dp], sVl + sw128(r, c));
        }
#pragma unroll
        for (int dp = 0; dp < 4; dp++) { mma16816(o[2 * dp], pl[ks], bh[dp][0], bh[dp][1]); mma16816(o[2 * dp + 1], pl[ks], bh[dp][2], bh[dp][3]); }
#pragma unroll
        for (int dp = 0; dp < 4; dp++) { mma16816(o[2 * dp], ph[ks], bl[dp][0], bl[dp][1]); mma16816(o[2 * dp + 1], ph[ks], bl[dp][2], bl[dp][3]); }
#pragma unroll
        for (int dp = 0; dp < 4; dp++) { mma16816(o[2 * dp], ph[ks], bh[dp][0], bh[dp][1]); mma16816(o[2 * dp + 1], ph[ks], bh[dp][2], bh[dp][3]); }
    }
    // ---- split to fp16 hi/lo into the warp's own Q rows (dead since the S loop), then two TMA stores
    __syncwarp();
#pragma unroll
    for (int nd = 0; nd < 8; nd++) {
        uint32_t hi, lo;
        split_pack2(o[nd][0], o[nd][1], hi, lo);
        *(uint32_t *) (smem + sw128(qr + g, nd) + 4 * t) = hi;
        *(uint32_t *) (smem + AT_TILE + sw128(qr + g, nd) + 4 * t) = lo;
        split_pack2(o[nd][2], o[nd][3], hi, lo);
        *(uint32_t *) (smem + sw128(qr + g + 8, nd) + 4 * t) = hi;
        *(uint32_t *) (smem + AT_TILE + sw128(qr + g + 8, nd) + 4 * t) = lo;
    }
    fence_async_smem();
    __syncwarp();
    if (lane == 0 && q0 + qr < rows) {
        tma_store_2d(&tm_out_hi, smem + qr * 128, h * 64, q0 + qr);
        tma_store_2d(&tm_out_lo, smem + AT_TILE + qr * 128, h * 64, q0 + qr);
        tma_store_commit();
        tma_store_wait_read0();
    }
}

cudaError_t attention_tc_init() {
    return cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM);
}

// tm_in_*: maps over the fp16 [rows][3*D] q|k|v buffers, box {64, 64}, SWIZZLE_128B, row extent == rows;
// tm_out_*: maps over the fp16 [rows][D] outputs, box {64, 16}, SWIZZLE_128B; bounds: launch_attn_bounds(window).
cudaError_t launch_attention_tc(const CUtensorMap & tm_in_hi, const CUtensorMap & tm_in_lo, const CUtensorMap & tm_out_hi,
                                const CUtensorMap & tm_out_lo, int rows, int D, int n_head, int window, const int2 * bounds,
                                cudaStream_t st, bool rev) {
    if (rows <= 0) return cudaSuccess;
    if (D != n_head * 64 || window / 2 != 32) return cudaErrorInvalidValue;
    dim3 grid((rows + AT_QB - 1) / AT_QB, n_head);
    launch_k(attention_tc_kernel, dim3(grid), dim3(128), AT_SMEM, st, tm_in_hi, tm_in_lo, tm_out_hi, tm_out_lo, rows, D, bounds, rev ? 1 : 0);
    return cudaGetLastError();
}

}  // namespace mio
