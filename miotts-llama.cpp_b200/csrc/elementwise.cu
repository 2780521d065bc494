// Memory-bound stages of the decoder: code-table gather, LayerNorm / AdaLN, banded RoPE attention,
// GroupNorm statistics + apply, operand splitting, AdaLN conditioning GEMV.
// Semantics follow the reference graph (src/miocodec.cpp:205-420) as executed by the ggml CPU backend
// (formulae: SURVEY.md Appendix A).
#include "kernels.cuh"
#include "ptx.cuh"

#include <cuda_bf16.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace mio {

bool pdl_enabled() {
    static const bool on = [] { const char * e = getenv("MIO_PDL"); return !(e && e[0] == '0'); }();
    return on;
}


__device__ __forceinline__ void split_half(float x, __half & hi, __half & lo) {
    hi = __float2half_rn(x);
    lo = __float2half_rn(x - __half2float(hi));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---------------------------------------------------------------------------------------------
// gather: x[row][:] = table[code[row]][:]   (ggml_get_rows, src/miocodec.cpp:599-600)
// one block per row, 128-bit loads/stores; pad rows (code < 0) are zero-filled; out-of-range codes raise the
// device error flag (the reference aborts in GGML_ASSERT, ggml-cpu/ops.cpp:4811).
__global__ void gather_embed_kernel(const float4 * __restrict__ table, const int32_t * __restrict__ codes,
                                    float4 * __restrict__ x, int D4, int vocab, int * __restrict__ err) {
    pdl_sync();
    const int row = blockIdx.x;
    int code = codes[row];
    const bool pad = code < 0;
    if (!pad && code >= vocab) {
        if (threadIdx.x == 0) atomicExch(err, 1 + row);
        code = 0;
    }
    const float4 * src = table + (size_t) (pad ? 0 : code) * D4;
    float4 * dst = x + (size_t) row * D4;
    for (int i = threadIdx.x; i < D4; i += blockDim.x) {
        float4 v = pad ? make_float4(0.f, 0.f, 0.f, 0.f) : __ldg(src + i);
        dst[i] = v;
    }
}

cudaError_t launch_gather(const float * table, const int32_t * codes_padded, float * x, int rows, int D, int vocab,
                          int * err_flag, cudaStream_t st) {
    if (rows <= 0) return cudaSuccess;
    launch_k(gather_embed_kernel, dim3(rows), dim3(192), 0, st, (const float4 *) table, codes_padded, (float4 *) x, D / 4, vocab, err_flag);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// LayerNorm: one warp per row, NV float4 per lane (D = 128*NV).
//   prenet  : y = (x-mean)*rstd*w + b                       (layer_norm, src/miocodec.cpp:212-217)
//   decoder : y = (x-mean)*rstd*(1+scale[utt]) + shift[utt] (adaln_norm, src/miocodec.cpp:322-330)
// biased variance, rstd = 1/sqrtf(var+eps) (ggml_compute_forward_norm_f32, ggml-cpu/ops.cpp:3642-3687).
template <int NV>
__global__ void __launch_bounds__(256) layernorm_kernel(
        const float * __restrict__ x, int rows, float eps, const float * __restrict__ w, const float * __restrict__ b,
        const float * __restrict__ mod, int mod_ld, int shift_off, int scale_off,
        const int32_t * __restrict__ row_utt, const int32_t * __restrict__ voice, int row_div,
        __half * __restrict__ out_hi, __half * __restrict__ out_lo, float * __restrict__ out_f32, int rev) {
    pdl_sync();
    constexpr int D = NV * 128;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = (rev ? gridDim.x - 1 - blockIdx.x : blockIdx.x) * 8 + warp;
    if (row >= rows) return;
    const int utt = row_utt[row / row_div];
    const size_t base = (size_t) row * D;
    if (utt < 0) {   // pad row: zeros
#pragma unroll
        for (int i = 0; i < NV; i++) {
            const int c = (i * 32 + lane) * 4;
            if (out_hi) {
                *(uint2 *) (out_hi + base + c) = make_uint2(0u, 0u);
                if (out_lo) *(uint2 *) (out_lo + base + c) = make_uint2(0u, 0u);
            }
            if (out_f32) *(float4 *) (out_f32 + base + c) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        return;
    }
    float4 v[NV];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; i++) {
        v[i] = *(const float4 *) (x + base + (i * 32 + lane) * 4);
        s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
    const float mean = warp_sum(s) / (float) D;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < NV; i++) {
        v[i].x -= mean; v[i].y -= mean; v[i].z -= mean; v[i].w -= mean;
        q += (v[i].x * v[i].x + v[i].y * v[i].y) + (v[i].z * v[i].z + v[i].w * v[i].w);
    }
    const float var = warp_sum(q) / (float) D;
    const float rstd = 1.0f / sqrtf(var + eps);
    const float * shift = nullptr, * scale = nullptr;
    if (mod) {
        const float * m = mod + (size_t) voice[utt] * mod_ld;
        shift = m + shift_off;
        scale = m + scale_off;
    }
#pragma unroll
    for (int i = 0; i < NV; i++) {
        const int c = (i * 32 + lane) * 4;
        float y[4] = {v[i].x * rstd, v[i].y * rstd, v[i].z * rstd, v[i].w * rstd};
        if (mod) {
            const float4 sc = *(const float4 *) (scale + c), sh = *(const float4 *) (shift + c);
            y[0] = y[0] * (1.0f + sc.x) + sh.x; y[1] = y[1] * (1.0f + sc.y) + sh.y;
            y[2] = y[2] * (1.0f + sc.z) + sh.z; y[3] = y[3] * (1.0f + sc.w) + sh.w;
        } else {
            const float4 ww = *(const float4 *) (w + c), bb = *(const float4 *) (b + c);
            y[0] = y[0] * ww.x + bb.x; y[1] = y[1] * ww.y + bb.y;
            y[2] = y[2] * ww.z + bb.z; y[3] = y[3] * ww.w + bb.w;
        }
        if (out_hi) {
            __half h[4], l[4];
#pragma unroll
            for (int j = 0; j < 4; j++) split_half(y[j], h[j], l[j]);
            *(uint2 *) (out_hi + base + c) = *(uint2 *) h;
            if (out_lo) *(uint2 *) (out_lo + base + c) = *(uint2 *) l;
        }
        if (out_f32) *(float4 *) (out_f32 + base + c) = make_float4(y[0], y[1], y[2], y[3]);
    }
}

cudaError_t launch_layernorm(const float * x, int rows, int D, float eps, const float * w, const float * b,
                             const float * mod, int mod_ld, int shift_off, int scale_off, const row_map & rm, int row_div,
                             __half * out_hi, __half * out_lo, float * out_f32, cudaStream_t st, bool rev) {
    if (rows <= 0) return cudaSuccess;
    const int grid = (rows + 7) / 8;
    if (D == 768) {
        launch_k(layernorm_kernel<6>, dim3(grid), dim3(256), 0, st, x, rows, eps, w, b, mod, mod_ld, shift_off, scale_off, rm.row_utt,
                                                  rm.voice, row_div, out_hi, out_lo, out_f32, rev ? 1 : 0);
    } else if (D == 512) {
        launch_k(layernorm_kernel<4>, dim3(grid), dim3(256), 0, st, x, rows, eps, w, b, mod, mod_ld, shift_off, scale_off, rm.row_utt,
                                                  rm.voice, row_div, out_hi, out_lo, out_f32, rev ? 1 : 0);
    } else {
        return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Banded local attention, CUDA-core validation kernel (fp32 dot products): head_dim 64, |i-j| <= window/2,
// per-utterance boundaries; q, k arrive rotated/scaled as fp16 hi/lo pairs from the QKV epilogue (EPI_QKV).
// (mha_rope, src/miocodec.cpp:245-286; the reference materialises dense scores with a -inf mask, :233-242 — masked
// terms are exactly 0 after softmax so the band is identical.)
// Block: 4 warps, ATT_QB query rows x 1 head; K and V of rows [q0-32, q0+QB+32) staged in smem.
constexpr int ATT_QB = 32;
constexpr int ATT_HALF = 32;
constexpr int ATT_KROWS = ATT_QB + 2 * ATT_HALF;   // 96
constexpr int ATT_KLD = 65;                        // padded: lane j reads K[j][d] conflict-free

__device__ __forceinline__ float2 ld_hl2(const __half * hi, const __half * lo, size_t idx) {
    const float2 a = __half22float2(*(const __half2 *) (hi + idx)), b = __half22float2(*(const __half2 *) (lo + idx));
    return make_float2(a.x + b.x, a.y + b.y);
}

__global__ void __launch_bounds__(128) attention_kernel(
        const __half * __restrict__ qkv_hi, const __half * __restrict__ qkv_lo, int rows, int D, int half_w,
        const int32_t * __restrict__ row_utt, const int32_t * __restrict__ start, const int32_t * __restrict__ len,
        int row_div, __half * __restrict__ out_hi, __half * __restrict__ out_lo) {
    pdl_sync();
    extern __shared__ float att_smem[];
    float * Ks = att_smem;                         // [KROWS][65]
    float * Vs = Ks + ATT_KROWS * ATT_KLD;         // [KROWS][64]
    float * Qs = Vs + ATT_KROWS * 64;              // [4][64]
    const int q0 = blockIdx.x * ATT_QB;
    const int h = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ld = 3 * D;
    const int k_lo = q0 - ATT_HALF;

    // stage K (with RoPE) and V: each warp takes rows warp, warp+4, ...; lane owns the pair (2*lane, 2*lane+1)
    for (int r = warp; r < ATT_KROWS; r += 4) {
        const int j = k_lo + r;
        float2 kk = make_float2(0.f, 0.f), vv = make_float2(0.f, 0.f);
        if (j >= 0 && j < rows) {
            const int u = row_utt[j / row_div];
            if (u >= 0) {
                kk = ld_hl2(qkv_hi, qkv_lo, (size_t) j * ld + D + h * 64 + 2 * lane);
                vv = ld_hl2(qkv_hi, qkv_lo, (size_t) j * ld + 2 * D + h * 64 + 2 * lane);
            }
        }
        Ks[r * ATT_KLD + 2 * lane] = kk.x;
        Ks[r * ATT_KLD + 2 * lane + 1] = kk.y;
        *(float2 *) (Vs + r * 64 + 2 * lane) = vv;
    }
    __syncthreads();

    float * qs = Qs + warp * 64;
    for (int qi = warp; qi < ATT_QB; qi += 4) {
        const int i = q0 + qi;
        if (i >= rows) break;                       // warp-uniform
        const int u = row_utt[i / row_div];
        const size_t obase = (size_t) i * D + h * 64;
        if (u < 0) {                                // pad row
            out_hi[obase + lane] = __float2half_rn(0.f); out_hi[obase + 32 + lane] = __float2half_rn(0.f);
            out_lo[obase + lane] = __float2half_rn(0.f); out_lo[obase + 32 + lane] = __float2half_rn(0.f);
            continue;
        }
        const int ustart = start[u] * row_div, uend = ustart + len[u] * row_div;
        {
            const float2 qin = ld_hl2(qkv_hi, qkv_lo, (size_t) i * ld + h * 64 + 2 * lane);
            __syncwarp();
            qs[2 * lane] = qin.x;
            qs[2 * lane + 1] = qin.y;
            __syncwarp();
        }
        // scores for keys i-32+jj, jj = lane, lane+32, lane+64
        float sc[3];
#pragma unroll
        for (int t = 0; t < 3; t++) {
            const int jj = lane + 32 * t;
            const int j = i - half_w + jj;
            const bool valid = (jj <= 2 * half_w) && j >= ustart && j < uend;
            float acc = 0.f;
            if (valid) {
                const float * kr = Ks + (j - k_lo) * ATT_KLD;
#pragma unroll 16
                for (int d = 0; d < 64; d++) acc += qs[d] * kr[d];   // q carries the 1/sqrt(64) factor (EPI_QKV)
            }
            sc[t] = valid ? acc : -INFINITY;
        }
        const float mx = warp_max(fmaxf(sc[0], fmaxf(sc[1], sc[2])));
        float p[3], ps = 0.f;
#pragma unroll
        for (int t = 0; t < 3; t++) { p[t] = (sc[t] == -INFINITY) ? 0.f : expf(sc[t] - mx); ps += p[t]; }
        const float inv = 1.0f / warp_sum(ps);
#pragma unroll
        for (int t = 0; t < 3; t++) p[t] *= inv;
        // out[d] = sum_j p_j V[j][d], lane owns d = lane and lane+32
        float o0 = 0.f, o1 = 0.f;
#pragma unroll
        for (int t = 0; t < 3; t++) {
            const int nj = (t < 2) ? 32 : (2 * half_w + 1 - 64);
            for (int l = 0; l < nj; l++) {
                const float pj = __shfl_sync(0xffffffffu, p[t], l);
                const int j = i - half_w + l + 32 * t;
                if (pj != 0.f) {                    // warp-uniform (pj broadcast)
                    const float * vr = Vs + (j - k_lo) * 64;
                    o0 += pj * vr[lane];
                    o1 += pj * vr[lane + 32];
                }
            }
        }
        __half h0, l0, h1, l1;
        split_half(o0, h0, l0);
        split_half(o1, h1, l1);
        out_hi[obase + lane] = h0; out_hi[obase + 32 + lane] = h1;
        out_lo[obase + lane] = l0; out_lo[obase + 32 + lane] = l1;
    }
}


static constexpr size_t ATT_SMEM = (size_t) (ATT_KROWS * ATT_KLD + ATT_KROWS * 64 + 4 * 64) * sizeof(float);

cudaError_t attention_init() {
    cudaError_t ce = cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) ATT_SMEM);
    if (ce != cudaSuccess) return ce;
    return attention_tc_init();
}

// CUDA-core validation kernel (MIO_ATTN=simt); the product path is launch_attention_tc (attention.cu)
cudaError_t launch_attention_simt(const __half * qkv_hi, const __half * qkv_lo, int rows, int D, int n_head, int window,
                                  const row_map & rm, int row_div, __half * out_hi, __half * out_lo, cudaStream_t st) {
    if (rows <= 0) return cudaSuccess;
    if (D != n_head * 64 || window / 2 != ATT_HALF) return cudaErrorInvalidValue;
    dim3 grid((rows + ATT_QB - 1) / ATT_QB, n_head);
    launch_k(attention_kernel, dim3(grid), dim3(128), ATT_SMEM, st, qkv_hi, qkv_lo, rows, D, window / 2, rm.row_utt, rm.start, rm.len, row_div,
                                                  out_hi, out_lo);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// GroupNorm statistics over (all rows of an utterance) x (C/32 channels)
// (conv_group_norm, src/miocodec.cpp:358-371; ggml_compute_forward_group_norm_f32, ggml-cpu/ops.cpp:3955-4028).
//
// Deterministic and batch-invariant, like the reference: a fixed-order two-stage reduction whose
// partition depends only on the utterance.  Stage 1: one block per (utterance, segment of GN_SEG_CODES codes counted
// from the utterance's first row); thread (c4, rl) sums one float4 column over the rows rl, rl + RL, ... of the segment
// in ascending order, the partial sums of a group are combined by an xor-shuffle tree and then over rl in ascending
// order -> part[segment][group] = (sum, sum of squares) in fp64.  Stage 2 (gn_finalize_kernel): one warp per
// (utterance, group) adds the segments in a fixed order.  No atomics: the same utterance gives the same bits alone,
// in any batch, at any position and under any chunking.
__global__ void __launch_bounds__(256) gn_stats_kernel(const float * __restrict__ x, int C, const int32_t * __restrict__ start,
                                                       const int32_t * __restrict__ len, const int32_t * __restrict__ seg_cu,
                                                       int n_utts, int row_div, double2 * __restrict__ part, int rev) {
    pdl_sync();
    __shared__ float2 red[8][32];
    const int sb = rev ? gridDim.x - 1 - blockIdx.x : blockIdx.x;
    int lo = 0, hi = n_utts;                       // largest u with seg_cu[u] <= sb
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (seg_cu[mid] <= sb) lo = mid; else hi = mid;
    }
    const int u = lo, seg = sb - seg_cu[u];
    const int C4 = C >> 2;                         // threads per row: 128, 64 or 32
    const int RL = 256 / C4;                       // row lanes: 2, 4 or 8
    const int c4 = threadIdx.x % C4, rl = threadIdx.x / C4;
    const int cpg = C / 32, tpg = cpg / 4;         // channels / threads per group: (16,4), (8,2), (4,1)
    const int r0 = (start[u] + seg * GN_SEG_CODES) * row_div;
    const int nrows = min(GN_SEG_CODES, len[u] - seg * GN_SEG_CODES) * row_div;
    const float * xp = x + (size_t) r0 * C + c4 * 4;
    float s = 0.f, q = 0.f;
    for (int rb = rl; rb < nrows; rb += 8 * RL) {
        float4 v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int r = rb + k * RL;
            v[k] = (r < nrows) ? __ldg((const float4 *) (xp + (size_t) r * C)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int k = 0; k < 8; k++) {
            s += (v[k].x + v[k].y) + (v[k].z + v[k].w);
            q += (v[k].x * v[k].x + v[k].y * v[k].y) + (v[k].z * v[k].z + v[k].w * v[k].w);
        }
    }
    for (int o = tpg >> 1; o > 0; o >>= 1) {       // the tpg threads of a group are adjacent lanes of one warp
        s += __shfl_xor_sync(0xffffffffu, s, o);
        q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    if ((c4 % tpg) == 0) red[rl][c4 / tpg] = make_float2(s, q);
    __syncthreads();
    if (threadIdx.x < 32) {
        double ds = 0.0, dq = 0.0;
        for (int k = 0; k < RL; k++) { ds += (double) red[k][threadIdx.x].x; dq += (double) red[k][threadIdx.x].y; }
        part[(size_t) sb * 32 + threadIdx.x] = make_double2(ds, dq);
    }
}

// Any channel count that is a multiple of 32 (upsampler widths other than 512 / 256 / 128): warp w of the block owns
// groups w, w + 8, ...; its lanes walk the (row, channel-in-group) cells of the segment in a fixed strided order and a
// shuffle tree adds the 32 partial sums, so the result is as run-to-run and batch invariant as the fast kernel's.
__global__ void __launch_bounds__(256) gn_stats_generic_kernel(const float * __restrict__ x, int C, const int32_t * __restrict__ start,
                                                               const int32_t * __restrict__ len, const int32_t * __restrict__ seg_cu,
                                                               int n_utts, int row_div, double2 * __restrict__ part) {
    pdl_sync();
    const int sb = blockIdx.x;
    int lo = 0, hi = n_utts;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (seg_cu[mid] <= sb) lo = mid; else hi = mid;
    }
    const int u = lo, seg = sb - seg_cu[u];
    const int cpg = C / 32;
    const int r0 = (start[u] + seg * GN_SEG_CODES) * row_div;
    const int nrows = min(GN_SEG_CODES, len[u] - seg * GN_SEG_CODES) * row_div;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int g = warp; g < 32; g += 8) {
        float s = 0.f, q = 0.f;
        const float * xp = x + (size_t) r0 * C + g * cpg;
        for (int i = lane; i < nrows * cpg; i += 32) {
            const int r = i / cpg, c = i - r * cpg;
            const float v = __ldg(xp + (size_t) r * C + c);
            s += v; q += v * v;
        }
        for (int o = 16; o > 0; o >>= 1) {
            s += __shfl_xor_sync(0xffffffffu, s, o);
            q += __shfl_xor_sync(0xffffffffu, q, o);
        }
        if (lane == 0) part[(size_t) sb * 32 + g] = make_double2((double) s, (double) q);
    }
}

cudaError_t launch_gn_stats(const float * x, int C, const row_map & rm, int row_div, const int32_t * seg_cu, int total_segs,
                            double2 * part, cudaStream_t st, bool rev) {
    if (total_segs <= 0) return cudaSuccess;
    if (C <= 0 || C % 32 != 0) return cudaErrorInvalidValue;
    if (C != 512 && C != 256 && C != 128) {
        launch_k(gn_stats_generic_kernel, dim3(total_segs), dim3(256), 0, st, x, C, rm.start, rm.len, seg_cu, rm.n_utts, row_div, part);
        return cudaGetLastError();
    }
    launch_k(gn_stats_kernel, dim3(total_segs), dim3(256), 0, st, x, C, rm.start, rm.len, seg_cu, rm.n_utts, row_div, part, rev ? 1 : 0);
    return cudaGetLastError();
}

__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + expf(-x)); }

// stage 2: segments of an utterance added in ascending order (fp64), then (sum, sum of squares) -> (mean,
// 1/sqrt(var + eps)) in fp32, once per (utterance, group): keeps the fp64 arithmetic (slow on this part) out of the
// per-element kernel.  var = sum((x-mean)^2)/n with the float-rounded mean, like the reference's second pass
// (ops.cpp:3990-4010).
// One WARP per (utterance, group): lane l adds segments l, l + 32, ... in ascending order, a fixed xor-shuffle tree adds
// the 32 lane sums -- an order that depends on the utterance's own segment count only, so still batch invariant, and
// 44 segments (a 700-code utterance) cost two dependent loads per lane instead of 44 in a row.
__global__ void __launch_bounds__(256) gn_finalize_kernel(const double2 * __restrict__ part, const int32_t * __restrict__ seg_cu,
                                                          const int32_t * __restrict__ len, int n_utts, int row_div,
                                                          int cpg, float eps, float2 * __restrict__ mr) {
    pdl_sync();
    const int idx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (idx >= n_utts * 32) return;                // whole warps leave together
    const int u = idx >> 5, g = idx & 31;
    double sum = 0.0, sq = 0.0;
    for (int sg = seg_cu[u] + lane; sg < seg_cu[u + 1]; sg += 32) {
        const double2 p = part[(size_t) sg * 32 + g];
        sum += p.x; sq += p.y;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        sq += __shfl_xor_sync(0xffffffffu, sq, o);
    }
    if (lane != 0) return;
    const double n = (double) len[u] * row_div * cpg;
    const float mean = (float) (sum / n);
    const double md = (double) mean;
    const float var = (float) ((sq - 2.0 * md * sum) / n + md * md);
    mr[idx] = make_float2(mean, 1.0f / sqrtf(var + eps));
}

// GroupNorm apply + SiLU + round to fp16 (the conv operand: ggml_conv_1d's im2col rounds activations to F16,
// ggml-cpu/ops.cpp:6232-6256, so this rounding is the reference's own, not an approximation).
// One thread owns 8 fixed channels (affine parameters live in registers) and walks the block's rows four at a time, so
// four independent 32-byte loads are in flight per thread; SiLU uses the SFU exp/reciprocal (relative error ~2^-20,
// far below the fp16 rounding that follows).
constexpr int GNA_ROWS = 64;                     // rows per block

__device__ __forceinline__ float silu_fast(float x) { return __fdividef(x, 1.0f + __expf(-x)); }
// PRECISE (MIO_GN_PRECISE=1, accuracy experiments): libdevice expf + IEEE divide
template <bool PRECISE> __device__ __forceinline__ float silu_pick(float x) { return PRECISE ? silu_f(x) : silu_fast(x); }

template <bool PRECISE>
__global__ void __launch_bounds__(256) gn_apply_silu_kernel(
        const float * __restrict__ x, int rows, int C, const int32_t * __restrict__ row_utt, int row_div,
        const float2 * __restrict__ mr, const float * __restrict__ w, const float * __restrict__ b, __half * __restrict__ out,
        int rows_per_block, int rev) {
    pdl_sync();
    const int C8 = C >> 3;                       // threads per row: 64, 32 or 16
    const int rpp = 256 / C8;                    // rows per pass of the block
    const int c = (threadIdx.x % C8) * 8;
    const int rl = threadIdx.x / C8;
    const int cpg = C / 32;
    const int g0 = c / cpg, g1 = (c + 4) / cpg;
    const float4 w0 = __ldg((const float4 *) (w + c)), w1 = __ldg((const float4 *) (w + c + 4));
    const float4 b0 = __ldg((const float4 *) (b + c)), b1 = __ldg((const float4 *) (b + c + 4));
    const int r_begin = (rev ? gridDim.x - 1 - blockIdx.x : blockIdx.x) * rows_per_block, r_end = min(rows, r_begin + rows_per_block);
    for (int rb = r_begin + rl; rb < r_end; rb += 4 * rpp) {
        int u[4];
        float4 v0[4], v1[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int r = rb + k * rpp;
            u[k] = (r < r_end) ? __ldg(row_utt + r / row_div) : -2;
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int r = rb + k * rpp;
            if (u[k] >= 0) {
                v0[k] = __ldg((const float4 *) (x + (size_t) r * C + c));
                v1[k] = __ldg((const float4 *) (x + (size_t) r * C + c + 4));
            }
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int r = rb + k * rpp;
            if (u[k] == -2) continue;
            uint4 o = make_uint4(0u, 0u, 0u, 0u);
            if (u[k] >= 0) {
                const float2 m0 = __ldg(mr + (size_t) u[k] * 32 + g0), m1 = __ldg(mr + (size_t) u[k] * 32 + g1);
                const __half2 h0 = __floats2half2_rn(silu_pick<PRECISE>((v0[k].x - m0.x) * m0.y * w0.x + b0.x), silu_pick<PRECISE>((v0[k].y - m0.x) * m0.y * w0.y + b0.y));
                const __half2 h1 = __floats2half2_rn(silu_pick<PRECISE>((v0[k].z - m0.x) * m0.y * w0.z + b0.z), silu_pick<PRECISE>((v0[k].w - m0.x) * m0.y * w0.w + b0.w));
                const __half2 h2 = __floats2half2_rn(silu_pick<PRECISE>((v1[k].x - m1.x) * m1.y * w1.x + b1.x), silu_pick<PRECISE>((v1[k].y - m1.x) * m1.y * w1.y + b1.y));
                const __half2 h3 = __floats2half2_rn(silu_pick<PRECISE>((v1[k].z - m1.x) * m1.y * w1.z + b1.z), silu_pick<PRECISE>((v1[k].w - m1.x) * m1.y * w1.w + b1.w));
                o = make_uint4(*(const uint32_t *) &h0, *(const uint32_t *) &h1, *(const uint32_t *) &h2, *(const uint32_t *) &h3);
            }
            *(uint4 *) (out + (size_t) r * C + c) = o;
        }
    }
}

// Any even channel count per group (C a multiple of 64): one thread per channel pair, grid-stride over (row, pair).
template <bool PRECISE>
__global__ void __launch_bounds__(256) gn_apply_silu_generic_kernel(
        const float * __restrict__ x, int rows, int C, const int32_t * __restrict__ row_utt, int row_div,
        const float2 * __restrict__ mr, const float * __restrict__ w, const float * __restrict__ b, __half * __restrict__ out) {
    pdl_sync();
    const int C2 = C >> 1, cpg = C / 32;
    const size_t total = (size_t) rows * C2;
    for (size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t) gridDim.x * blockDim.x) {
        const int r = (int) (idx / C2), c = (int) (idx - (size_t) r * C2) * 2;
        const int u = __ldg(row_utt + r / row_div);
        __half2 h = __floats2half2_rn(0.f, 0.f);
        if (u >= 0) {
            const float2 v = *(const float2 *) (x + (size_t) r * C + c);
            const float2 m = __ldg(mr + (size_t) u * 32 + c / cpg);
            h = __floats2half2_rn(silu_pick<PRECISE>((v.x - m.x) * m.y * __ldg(w + c) + __ldg(b + c)),
                                  silu_pick<PRECISE>((v.y - m.x) * m.y * __ldg(w + c + 1) + __ldg(b + c + 1)));
        }
        *(__half2 *) (out + (size_t) r * C + c) = h;
    }
}

cudaError_t launch_gn_apply_silu(const float * x, int rows, int C, const row_map & rm, int row_div, const double2 * part,
                                 const int32_t * seg_cu, float eps, const float * w, const float * b, float2 * mr_scratch,
                                 __half * out, cudaStream_t st, bool rev) {
    if (rows <= 0) return cudaSuccess;
    if (C <= 0 || C % 64 != 0) return cudaErrorInvalidValue;
    launch_k(gn_finalize_kernel, dim3((rm.n_utts * 32 + 7) / 8), dim3(256), 0, st, part, seg_cu, rm.len, rm.n_utts, row_div, C / 32, eps, mr_scratch);
    static int precise_set = -1;
    if (precise_set < 0) {
        const char * ev = getenv("MIO_GN_PRECISE");
        precise_set = (ev && ev[0] == '1') ? 1 : 0;
    }
    if (C != 128 && C != 256 && C != 512 && C != 1024 && C != 2048) {     // the fast kernel needs C / 8 to divide 256
        const int grid = (int) min((size_t) 148 * 16, ((size_t) rows * (C / 2) + 255) / 256);
        if (precise_set) launch_k(gn_apply_silu_generic_kernel<true>, dim3(grid), dim3(256), 0, st, x, rows, C, rm.row_utt, row_div, mr_scratch, w, b, out);
        else launch_k(gn_apply_silu_generic_kernel<false>, dim3(grid), dim3(256), 0, st, x, rows, C, rm.row_utt, row_div, mr_scratch, w, b, out);
        return cudaGetLastError();
    }
    const int rpb = (rows >= GNA_ROWS * 148 * 2) ? GNA_ROWS : 16;      // few rows: more, shorter blocks (latency mode)
    if (precise_set) launch_k(gn_apply_silu_kernel<true>, dim3((rows + rpb - 1) / rpb), dim3(256), 0, st, x, rows, C, rm.row_utt, row_div, mr_scratch, w, b, out, rpb, rev ? 1 : 0);
    else launch_k(gn_apply_silu_kernel<false>, dim3((rows + rpb - 1) / rpb), dim3(256), 0, st, x, rows, C, rm.row_utt, row_div, mr_scratch, w, b, out, rpb, rev ? 1 : 0);
    return cudaGetLastError();
}

// fp32 -> (hi, lo) fp16 operand pair, zeros on pad rows (inputs of the ConvTranspose GEMMs, which read row+-1)
__global__ void split_kernel(const float * __restrict__ x, int rows, int C, const int32_t * __restrict__ row_utt,
                             int row_div, __half * __restrict__ hi, __half * __restrict__ lo) {
    pdl_sync();
    const int C4 = C / 4;
    const size_t total = (size_t) rows * C4;
    for (size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t) gridDim.x * blockDim.x) {
        const int r = (int) (idx / C4);
        const int u = row_utt[r / row_div];
        __half h[4], l[4];
        if (u < 0) {
#pragma unroll
            for (int j = 0; j < 4; j++) h[j] = l[j] = __float2half_rn(0.f);
        } else {
            const float4 v = *(const float4 *) (x + idx * 4);
            split_half(v.x, h[0], l[0]); split_half(v.y, h[1], l[1]);
            split_half(v.z, h[2], l[2]); split_half(v.w, h[3], l[3]);
        }
        *(uint2 *) (hi + idx * 4) = *(uint2 *) h;
        *(uint2 *) (lo + idx * 4) = *(uint2 *) l;
    }
}

cudaError_t launch_split(const float * x, int rows, int C, const row_map & rm, int row_div, __half * hi, __half * lo,
                         cudaStream_t st) {
    if (rows <= 0) return cudaSuccess;
    const size_t total = (size_t) rows * (C / 4);
    const int grid = (int) min((size_t) 148 * 16, (total + 255) / 256);
    launch_k(split_kernel, dim3(grid), dim3(256), 0, st, x, rows, C, rm.row_utt, row_div, hi, lo);
    return cudaGetLastError();
}

// BF16-stored matrices: the reference rounds the activations to BF16 (round to nearest even, ggml_compute_fp32_to_bf16)
// and multiplies once.  The (hi, lo) operand pair becomes one BF16 operand IN PLACE (bit patterns in the hi buffer);
// hi + lo carries 22 bits of the fp32 value, far beyond the 8 that survive.
__global__ void pair_to_bf16_kernel(__half * __restrict__ hi, const __half * __restrict__ lo, size_t n2) {
    pdl_sync();
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (size_t) gridDim.x * blockDim.x) {
        const float2 h = __half22float2(((const __half2 *) hi)[i]);
        float2 v = h;
        if (lo) { const float2 l = __half22float2(((const __half2 *) lo)[i]); v.x += l.x; v.y += l.y; }
        const __nv_bfloat162 o = __floats2bfloat162_rn(v.x, v.y);
        ((uint32_t *) hi)[i] = *(const uint32_t *) &o;
    }
}

cudaError_t launch_pair_to_bf16(__half * hi, const __half * lo, size_t n, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    if (n % 2) return cudaErrorInvalidValue;
    const int grid = (int) min((size_t) 148 * 16, (n / 2 + 255) / 256);
    launch_k(pair_to_bf16_kernel, dim3(grid), dim3(256), 0, st, hi, lo, n / 2);
    return cudaGetLastError();
}

// the activation rounding of a matrix product whose matrix is stored as F16 (1) / BF16 (2); 0 = F32, none
__device__ __forceinline__ float round_act(float x, int store) {
    if (store == 1) return __half2float(__float2half_rn(x));
    if (store == 2) return __bfloat162float(__float2bfloat16_rn(x));
    return x;
}

// AdaLN conditioning: mod[v][j] = b[j] + sum_k W[j][k] * silu(emb[v][k])
// (compute_adaln3, src/miocodec.cpp:310-319 and :654-655). One warp per output row j, all voices.
__global__ void cond_silu_kernel(const float * __restrict__ voice_table, int n, float * __restrict__ out) {
    pdl_sync();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = silu_f(voice_table[i]);
}

// one warp per output row j (its 128 weights live in registers), looping over the voices of this blockIdx.y
__global__ void cond_kernel(const float * __restrict__ voice_silu, int n_voices, int adaln_dim,
                            const float * __restrict__ w, const float * __restrict__ b, int cond_rows,
                            const uint8_t * __restrict__ row_store, float * __restrict__ mod) {
    pdl_sync();
    const int j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (j >= cond_rows) return;
    const float bj = b[j];
    const int store = row_store ? row_store[j] : 0;      // storage type of the matrix this row came from
    if (adaln_dim == 128) {
        const float4 wj = __ldg((const float4 *) (w + (size_t) j * 128) + lane);
        for (int v = blockIdx.y; v < n_voices; v += gridDim.y) {
            float4 s = __ldg((const float4 *) (voice_silu + (size_t) v * 128) + lane);
            if (store) s = make_float4(round_act(s.x, store), round_act(s.y, store), round_act(s.z, store), round_act(s.w, store));
            // same summation order per lane as the strided loop below would not matter: fp32 dot of 128 terms,
            // reduced over the warp; the reference's ggml_mul_mat sums in SIMD lanes of its own order
            float acc = wj.x * s.x + wj.y * s.y + wj.z * s.z + wj.w * s.w;
            acc = warp_sum(acc);
            if (lane == 0) mod[(size_t) v * cond_rows + j] = acc + bj;
        }
        return;
    }
    for (int v = blockIdx.y; v < n_voices; v += gridDim.y) {
        float acc = 0.f;
        for (int k = lane; k < adaln_dim; k += 32) acc += w[(size_t) j * adaln_dim + k] * round_act(voice_silu[(size_t) v * adaln_dim + k], store);
        acc = warp_sum(acc);
        if (lane == 0) mod[(size_t) v * cond_rows + j] = acc + bj;
    }
}

cudaError_t launch_cond(const float * voice_table, int n_voices, int adaln_dim, const float * w, const float * b,
                        int cond_rows, const uint8_t * row_store, float * silu_scratch, float * mod, cudaStream_t st) {
    if (n_voices <= 0) return cudaSuccess;
    const int n = n_voices * adaln_dim;
    launch_k(cond_silu_kernel, dim3((n + 255) / 256), dim3(256), 0, st, voice_table, n, silu_scratch);
    dim3 grid((cond_rows + 7) / 8, min(n_voices, 8));
    launch_k(cond_kernel, dim3(grid), dim3(256), 0, st, silu_scratch, n_voices, adaln_dim, w, b, cond_rows, row_store, mod);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// per-utterance peak normalisation: gain = 0.95/max|s| if peak > 1e-8 (src/test-to-speech.cpp:289-300)
__global__ void peak_max_kernel(const float * __restrict__ wave, const int32_t * __restrict__ len,
                                const int64_t * __restrict__ out_off, int spc, float * __restrict__ peak) {
    pdl_sync();
    const int u = blockIdx.y;
    const int64_t n = (int64_t) len[u] * spc;
    const float * w = wave + out_off[u];
    float m = 0.f;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) m = fmaxf(m, fabsf(w[i]));
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0) atomicMax((int *) (peak + u), __float_as_int(m));   // m >= 0: int order == float order
}
__global__ void peak_scale_kernel(float * __restrict__ wave, const int32_t * __restrict__ len,
                                  const int64_t * __restrict__ out_off, int spc, const float * __restrict__ peak) {
    pdl_sync();
    const int u = blockIdx.y;
    const float pk = peak[u];
    if (!(pk > 1e-8f)) return;
    const float gain = 0.95f / pk;
    const int64_t n = (int64_t) len[u] * spc;
    float * w = wave + out_off[u];
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) w[i] *= gain;
}

cudaError_t launch_peak_normalize(float * wave, const int32_t * len, const int64_t * out_off, int n_utts,
                                  int samples_per_code, float * peak_scratch, cudaStream_t st) {
    if (n_utts <= 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(peak_scratch, 0, sizeof(float) * n_utts, st);
    if (e != cudaSuccess) return e;
    dim3 grid(32, n_utts);
    launch_k(peak_max_kernel, dim3(grid), dim3(256), 0, st, wave, len, out_off, samples_per_code, peak_scratch);
    launch_k(peak_scale_kernel, dim3(grid), dim3(256), 0, st, wave, len, out_off, samples_per_code, peak_scratch);
    return cudaGetLastError();
}

// float -> PCM16 exactly as the reference's WAV writer does on the host (src/wav-writer.cpp:39-42):
// int16(clamp(s * 32767.0f, -32768.0f, 32767.0f)), i.e. truncation toward zero after the clamp.  Eight samples per
// thread (two 16-byte loads, one 16-byte store); the tail is handled per element.  NaN maps to -32768.
__device__ __forceinline__ int pcm16_of(float s) { return __float2int_rz(fminf(fmaxf(s * 32767.0f, -32768.0f), 32767.0f)); }

__global__ void __launch_bounds__(256) pcm16_kernel(const float * __restrict__ wave, int16_t * __restrict__ pcm,
                                                    int64_t begin, int64_t end) {
    pdl_sync();
    // absolute sample indices: [va, vb) is the 8-sample-aligned body (bases are 256-byte aligned allocations)
    int64_t va = (begin + 7) & ~(int64_t) 7, vb = end & ~(int64_t) 7;
    if (va > vb) { va = end; vb = end; }
    const int64_t i0 = va >> 3, i1 = vb >> 3;
    for (int64_t i = i0 + (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < i1; i += (int64_t) gridDim.x * blockDim.x) {
        const float4 a = __ldg((const float4 *) wave + 2 * i), b = __ldg((const float4 *) wave + 2 * i + 1);
        uint4 o;
        o.x = (uint32_t) (pcm16_of(a.x) & 0xffff) | ((uint32_t) pcm16_of(a.y) << 16);
        o.y = (uint32_t) (pcm16_of(a.z) & 0xffff) | ((uint32_t) pcm16_of(a.w) << 16);
        o.z = (uint32_t) (pcm16_of(b.x) & 0xffff) | ((uint32_t) pcm16_of(b.y) << 16);
        o.w = (uint32_t) (pcm16_of(b.z) & 0xffff) | ((uint32_t) pcm16_of(b.w) << 16);
        ((uint4 *) pcm)[i] = o;
    }
    if (blockIdx.x == 0) {                                  // unaligned head and tail: fewer than 8 samples each
        const int64_t h = begin + threadIdx.x, t = vb + threadIdx.x;
        if (threadIdx.x < 8 && h < va) pcm[h] = (int16_t) pcm16_of(wave[h]);
        if (threadIdx.x < 8 && t >= va && t < end) pcm[t] = (int16_t) pcm16_of(wave[t]);
    }
}

// converts samples [begin, end) of wave into the same positions of pcm (both base pointers of cudaMalloc allocations)
cudaError_t launch_pcm16(const float * wave, int16_t * pcm, int64_t begin, int64_t end, cudaStream_t st) {
    if (end <= begin) return cudaSuccess;
    const int64_t n8 = (end - begin) >> 3;
    const int grid = (int) std::min<int64_t>((int64_t) 148 * 8, std::max<int64_t>(1, (n8 + 255) / 256));
    launch_k(pcm16_kernel, dim3(grid), dim3(256), 0, st, wave, pcm, begin, end);
    return cudaGetLastError();
}

}  // namespace mio
