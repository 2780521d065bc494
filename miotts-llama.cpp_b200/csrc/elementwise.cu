// Memory-bound stages of the decoder: code-table gather, LayerNorm / AdaLN, banded RoPE attention,
// GroupNorm statistics + apply, operand splitting, AdaLN conditioning GEMV.
// Semantics follow the reference graph (src/miocodec.cpp:205-420) as executed by the ggml CPU backend
// (formulae: SURVEY.md Appendix A).
#include "kernels.cuh"

#include <cstdlib>
#include <cstring>

namespace mio {

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
    gather_embed_kernel<<<rows, 192, 0, st>>>((const float4 *) table, codes_padded, (float4 *) x, D / 4, vocab, err_flag);
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
        __half * __restrict__ out_hi, __half * __restrict__ out_lo, float * __restrict__ out_f32) {
    constexpr int D = NV * 128;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = blockIdx.x * 8 + warp;
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
                             __half * out_hi, __half * out_lo, float * out_f32, cudaStream_t st) {
    if (rows <= 0) return cudaSuccess;
    const int grid = (rows + 7) / 8;
    if (D == 768) {
        layernorm_kernel<6><<<grid, 256, 0, st>>>(x, rows, eps, w, b, mod, mod_ld, shift_off, scale_off, rm.row_utt,
                                                  rm.voice, row_div, out_hi, out_lo, out_f32);
    } else if (D == 512) {
        layernorm_kernel<4><<<grid, 256, 0, st>>>(x, rows, eps, w, b, mod, mod_ld, shift_off, scale_off, rm.row_utt,
                                                  rm.voice, row_div, out_hi, out_lo, out_f32);
    } else {
        return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Banded local attention, head_dim 64, |i-j| <= window/2, per-utterance boundaries, RoPE mode 0 (adjacent pairs)
// on q and k with per-utterance positions (mha_rope, src/miocodec.cpp:245-286; the reference materialises dense
// scores with a -inf mask, :233-242 — masked terms are exactly 0 after softmax so the band is identical).
// Block: 4 warps, ATT_QB query rows x 1 head; K (RoPE applied) and V of rows [q0-32, q0+QB+32) staged in smem.
constexpr int ATT_QB = 32;
constexpr int ATT_HALF = 32;
constexpr int ATT_KROWS = ATT_QB + 2 * ATT_HALF;   // 96
constexpr int ATT_KLD = 65;                        // padded: lane j reads K[j][d] conflict-free

__global__ void __launch_bounds__(128) attention_kernel(
        const float * __restrict__ qkv, int rows, int D, int half_w, const float * __restrict__ rope, int rope_max_pos,
        const int32_t * __restrict__ row_utt, const int32_t * __restrict__ start, const int32_t * __restrict__ len,
        int row_div, __half * __restrict__ out_hi, __half * __restrict__ out_lo) {
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
                int pos = j - start[u] * row_div;
                pos = min(pos, rope_max_pos - 1);
                const float2 cs = *(const float2 *) (rope + ((size_t) pos * 32 + lane) * 2);
                const float2 kin = *(const float2 *) (qkv + (size_t) j * ld + D + h * 64 + 2 * lane);
                kk.x = kin.x * cs.x - kin.y * cs.y;
                kk.y = kin.x * cs.y + kin.y * cs.x;
                vv = *(const float2 *) (qkv + (size_t) j * ld + 2 * D + h * 64 + 2 * lane);
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
            int pos = min(i - ustart, rope_max_pos - 1);
            const float2 cs = *(const float2 *) (rope + ((size_t) pos * 32 + lane) * 2);
            const float2 qin = *(const float2 *) (qkv + (size_t) i * ld + h * 64 + 2 * lane);
            __syncwarp();
            qs[2 * lane] = qin.x * cs.x - qin.y * cs.y;
            qs[2 * lane + 1] = qin.x * cs.y + qin.y * cs.x;
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
                for (int d = 0; d < 64; d++) acc += qs[d] * kr[d];
                acc *= 0.125f;                      // 1/sqrt(64)
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


// ---------------------------------------------------------------------------------------------
// Tensor-core version of the same banded attention (default).  FA2-style: one warp owns 16 query rows and the
// 80 keys [r0-32, r0+48) that cover their bands; S = Q.K^T and O = P.V run as warp-level m16n8k16 MMAs on
// fragments with split-precision operands (x = x_hi + x_lo in fp16, three MMAs per product, fp32 accumulate), so
// the result keeps fp32-class accuracy; softmax is done on the accumulator fragments.
// Block = 4 warps = 64 queries x 1 head; Q, K (RoPE applied) and V^T are staged once per block as fp16 pairs.
constexpr int AM_QB = 64;
constexpr int AM_KR = AM_QB + 64;          // 128 key rows staged
constexpr int AM_LDQ = 72;                 // padded row strides (halves): fragment loads are bank-conflict free
constexpr int AM_LDV = AM_KR + 8;          // 136
constexpr size_t AM_SMEM = (size_t) (2 * AM_QB * AM_LDQ + 2 * AM_KR * AM_LDQ + 2 * 64 * AM_LDV) * sizeof(__half);

__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(__half a, __half b) {
    return (uint32_t) __half_as_ushort(a) | ((uint32_t) __half_as_ushort(b) << 16);
}
__device__ __forceinline__ void split_pack(float x0, float x1, uint32_t & hi, uint32_t & lo) {
    __half h0, l0, h1, l1;
    split_half(x0, h0, l0);
    split_half(x1, h1, l1);
    hi = pack_h2(h0, h1);
    lo = pack_h2(l0, l1);
}

__global__ void __launch_bounds__(128, 2) attention_mma_kernel(
        const float * __restrict__ qkv, int rows, int D, int half_w, const float * __restrict__ rope, int rope_max_pos,
        const int32_t * __restrict__ row_utt, const int32_t * __restrict__ start, const int32_t * __restrict__ len,
        int row_div, __half * __restrict__ out_hi, __half * __restrict__ out_lo) {
    extern __shared__ __align__(16) unsigned char am_smem_raw[];
    __half * Qh = (__half *) am_smem_raw;           // [64][72]
    __half * Ql = Qh + AM_QB * AM_LDQ;
    __half * Kh = Ql + AM_QB * AM_LDQ;              // [128][72]
    __half * Kl = Kh + AM_KR * AM_LDQ;
    __half * Vh = Kl + AM_KR * AM_LDQ;              // V^T [64][136]
    __half * Vl = Vh + 64 * AM_LDV;
    const int q0 = blockIdx.x * AM_QB;
    const int h = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ld = 3 * D;
    const int k_lo = q0 - 32;

    // ---- staging: lane owns the RoPE pair (2*lane, 2*lane+1) of a row
    for (int r = warp; r < AM_KR; r += 4) {
        const int j = k_lo + r;
        float2 kk = make_float2(0.f, 0.f), vv = make_float2(0.f, 0.f), qq = make_float2(0.f, 0.f);
        const bool is_q = (r >= 32 && r < 32 + AM_QB);
        if (j >= 0 && j < rows) {
            const int u = row_utt[j / row_div];
            if (u >= 0) {
                const int pos = min(j - start[u] * row_div, rope_max_pos - 1);
                const float2 cs = *(const float2 *) (rope + ((size_t) pos * 32 + lane) * 2);
                const float * base = qkv + (size_t) j * ld + h * 64 + 2 * lane;
                const float2 kin = *(const float2 *) (base + D);
                kk.x = kin.x * cs.x - kin.y * cs.y;
                kk.y = kin.x * cs.y + kin.y * cs.x;
                vv = *(const float2 *) (base + 2 * D);
                if (is_q) {
                    const float2 qin = *(const float2 *) base;
                    qq.x = qin.x * cs.x - qin.y * cs.y;
                    qq.y = qin.x * cs.y + qin.y * cs.x;
                }
            }
        }
        uint32_t hi, lo;
        split_pack(kk.x, kk.y, hi, lo);
        *(uint32_t *) (Kh + r * AM_LDQ + 2 * lane) = hi;
        *(uint32_t *) (Kl + r * AM_LDQ + 2 * lane) = lo;
        if (is_q) {
            split_pack(qq.x, qq.y, hi, lo);
            *(uint32_t *) (Qh + (r - 32) * AM_LDQ + 2 * lane) = hi;
            *(uint32_t *) (Ql + (r - 32) * AM_LDQ + 2 * lane) = lo;
        }
        __half h0, l0, h1, l1;
        split_half(vv.x, h0, l0);
        split_half(vv.y, h1, l1);
        Vh[(2 * lane) * AM_LDV + r] = h0; Vh[(2 * lane + 1) * AM_LDV + r] = h1;
        Vl[(2 * lane) * AM_LDV + r] = l0; Vl[(2 * lane + 1) * AM_LDV + r] = l1;
    }
    __syncthreads();

    const int g = lane >> 2, t = lane & 3;
    const int qr = warp * 16;                       // warp's first query row inside the block
    const int kw = warp * 16;                       // warp's first key row inside the staged window (query - 32)
    // ---- S = Q K^T over 10 key tiles of 8
    float s[10][4];
#pragma unroll
    for (int nt = 0; nt < 10; nt++) { s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f; }
#pragma unroll
    for (int ks = 0; ks < 4; ks++) {
        uint32_t ah[4], al[4];
        const int c = ks * 16 + 2 * t;
        ah[0] = *(const uint32_t *) (Qh + (qr + g) * AM_LDQ + c);      al[0] = *(const uint32_t *) (Ql + (qr + g) * AM_LDQ + c);
        ah[1] = *(const uint32_t *) (Qh + (qr + g + 8) * AM_LDQ + c);  al[1] = *(const uint32_t *) (Ql + (qr + g + 8) * AM_LDQ + c);
        ah[2] = *(const uint32_t *) (Qh + (qr + g) * AM_LDQ + c + 8);  al[2] = *(const uint32_t *) (Ql + (qr + g) * AM_LDQ + c + 8);
        ah[3] = *(const uint32_t *) (Qh + (qr + g + 8) * AM_LDQ + c + 8); al[3] = *(const uint32_t *) (Ql + (qr + g + 8) * AM_LDQ + c + 8);
#pragma unroll
        for (int nt = 0; nt < 10; nt++) {
            const int kr = kw + nt * 8 + g;
            const uint32_t bh0 = *(const uint32_t *) (Kh + kr * AM_LDQ + c), bh1 = *(const uint32_t *) (Kh + kr * AM_LDQ + c + 8);
            const uint32_t bl0 = *(const uint32_t *) (Kl + kr * AM_LDQ + c), bl1 = *(const uint32_t *) (Kl + kr * AM_LDQ + c + 8);
            mma_16816(s[nt], al, bh0, bh1);
            mma_16816(s[nt], ah, bl0, bl1);
            mma_16816(s[nt], ah, bh0, bh1);
        }
    }
    // ---- masked softmax on the fragments: thread holds rows i0 = g, i1 = g+8; key columns nt*8 + 2t, +1
    const int i0 = q0 + qr + g, i1 = i0 + 8;
    int us0 = 0, ue0 = 0, us1 = 0, ue1 = 0;
    bool v0 = false, v1 = false;
    if (i0 < rows) { const int u = row_utt[i0 / row_div]; if (u >= 0) { v0 = true; us0 = start[u] * row_div; ue0 = us0 + len[u] * row_div; } }
    if (i1 < rows) { const int u = row_utt[i1 / row_div]; if (u >= 0) { v1 = true; us1 = start[u] * row_div; ue1 = us1 + len[u] * row_div; } }
    float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 10; nt++) {
#pragma unroll
        for (int e = 0; e < 2; e++) {
            const int j = k_lo + kw + nt * 8 + 2 * t + e;
            const bool ok0 = v0 && j >= us0 && j < ue0 && j >= i0 - half_w && j <= i0 + half_w;
            const bool ok1 = v1 && j >= us1 && j < ue1 && j >= i1 - half_w && j <= i1 + half_w;
            s[nt][e] = ok0 ? s[nt][e] * 0.125f : -INFINITY;
            s[nt][2 + e] = ok1 ? s[nt][2 + e] * 0.125f : -INFINITY;
            m0 = fmaxf(m0, s[nt][e]);
            m1 = fmaxf(m1, s[nt][2 + e]);
        }
    }
    m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
    m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
    float l0 = 0.f, l1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 10; nt++) {
#pragma unroll
        for (int e = 0; e < 2; e++) {
            const float p0 = (s[nt][e] == -INFINITY) ? 0.f : expf(s[nt][e] - m0);
            const float p1 = (s[nt][2 + e] == -INFINITY) ? 0.f : expf(s[nt][2 + e] - m1);
            s[nt][e] = p0; s[nt][2 + e] = p1;
            l0 += p0; l1 += p1;
        }
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float inv0 = v0 ? 1.0f / l0 : 0.f, inv1 = v1 ? 1.0f / l1 : 0.f;
    // ---- O = P V over 5 key steps of 16, 8 d-tiles of 8
    float o[8][4];
#pragma unroll
    for (int nd = 0; nd < 8; nd++) { o[nd][0] = o[nd][1] = o[nd][2] = o[nd][3] = 0.f; }
#pragma unroll
    for (int ks = 0; ks < 5; ks++) {
        uint32_t ph[4], pl[4];
        split_pack(s[2 * ks][0] * inv0, s[2 * ks][1] * inv0, ph[0], pl[0]);
        split_pack(s[2 * ks][2] * inv1, s[2 * ks][3] * inv1, ph[1], pl[1]);
        split_pack(s[2 * ks + 1][0] * inv0, s[2 * ks + 1][1] * inv0, ph[2], pl[2]);
        split_pack(s[2 * ks + 1][2] * inv1, s[2 * ks + 1][3] * inv1, ph[3], pl[3]);
        const int kc = kw + ks * 16 + 2 * t;
#pragma unroll
        for (int nd = 0; nd < 8; nd++) {
            const int dr = nd * 8 + g;
            const uint32_t bh0 = *(const uint32_t *) (Vh + dr * AM_LDV + kc), bh1 = *(const uint32_t *) (Vh + dr * AM_LDV + kc + 8);
            const uint32_t bl0 = *(const uint32_t *) (Vl + dr * AM_LDV + kc), bl1 = *(const uint32_t *) (Vl + dr * AM_LDV + kc + 8);
            mma_16816(o[nd], pl, bh0, bh1);
            mma_16816(o[nd], ph, bl0, bl1);
            mma_16816(o[nd], ph, bh0, bh1);
        }
    }
    // ---- store split fp16 (pad rows get zeros: inv = 0)
#pragma unroll
    for (int nd = 0; nd < 8; nd++) {
        const int c = h * 64 + nd * 8 + 2 * t;
        uint32_t hi, lo;
        if (i0 < rows) {
            split_pack(o[nd][0], o[nd][1], hi, lo);
            *(uint32_t *) (out_hi + (size_t) i0 * D + c) = hi;
            *(uint32_t *) (out_lo + (size_t) i0 * D + c) = lo;
        }
        if (i1 < rows) {
            split_pack(o[nd][2], o[nd][3], hi, lo);
            *(uint32_t *) (out_hi + (size_t) i1 * D + c) = hi;
            *(uint32_t *) (out_lo + (size_t) i1 * D + c) = lo;
        }
    }
}

static constexpr size_t ATT_SMEM = (size_t) (ATT_KROWS * ATT_KLD + ATT_KROWS * 64 + 4 * 64) * sizeof(float);

static bool g_attn_simt = false;   // MIO_ATTN=simt selects the CUDA-core validation kernel

cudaError_t attention_init() {
    const char * e = getenv("MIO_ATTN");
    g_attn_simt = e && !strcmp(e, "simt");
    cudaError_t ce = cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) ATT_SMEM);
    if (ce != cudaSuccess) return ce;
    return cudaFuncSetAttribute(attention_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) AM_SMEM);
}

cudaError_t launch_attention(const float * qkv, int rows, int D, int n_head, int window, const float * rope,
                             int rope_max_pos, const row_map & rm, int row_div, __half * out_hi, __half * out_lo,
                             cudaStream_t st) {
    if (rows <= 0) return cudaSuccess;
    if (D != n_head * 64 || window / 2 != ATT_HALF) return cudaErrorInvalidValue;
    if (!g_attn_simt) {
        dim3 gm((rows + AM_QB - 1) / AM_QB, n_head);
        attention_mma_kernel<<<gm, 128, AM_SMEM, st>>>(qkv, rows, D, window / 2, rope, rope_max_pos, rm.row_utt, rm.start,
                                                       rm.len, row_div, out_hi, out_lo);
        return cudaGetLastError();
    }
    dim3 grid((rows + ATT_QB - 1) / ATT_QB, n_head);
    attention_kernel<<<grid, 128, ATT_SMEM, st>>>(qkv, rows, D, window / 2, rope, rope_max_pos, rm.row_utt, rm.start,
                                                  rm.len, row_div, out_hi, out_lo);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// GroupNorm statistics over (all rows of an utterance) x (C/32 channels)
// (conv_group_norm, src/miocodec.cpp:358-371; ggml_compute_forward_group_norm_f32, ggml-cpu/ops.cpp:3955-4028).
// Block = C/4 threads (one float4 of channels each) x GN_ROWS rows; partial sums are flushed with fp64 atomics
// whenever the utterance changes inside the chunk.
constexpr int GN_ROWS = 64;

__global__ void gn_stats_kernel(const float * __restrict__ x, int rows, int C, const int32_t * __restrict__ row_utt,
                                int row_div, double * __restrict__ stats) {
    const int c4 = threadIdx.x;                    // float4 index within the row
    const int cpg = C / 32;                        // channels per group: 16, 8 or 4
    const int tpg = cpg / 4;                       // threads per group: 4, 2 or 1
    const int group = (c4 * 4) / cpg;
    const int r0 = blockIdx.x * GN_ROWS;
    const int r1 = min(rows, r0 + GN_ROWS);
    float s = 0.f, q = 0.f;
    int cur = -1;
    auto flush = [&](int utt) {
        float ss = s, qq = q;
        for (int o = tpg >> 1; o > 0; o >>= 1) {
            ss += __shfl_xor_sync(0xffffffffu, ss, o);
            qq += __shfl_xor_sync(0xffffffffu, qq, o);
        }
        if (utt >= 0 && (c4 % tpg) == 0) {
            atomicAdd(stats + ((size_t) utt * 32 + group) * 2, (double) ss);
            atomicAdd(stats + ((size_t) utt * 32 + group) * 2 + 1, (double) qq);
        }
        s = 0.f; q = 0.f;
    };
    for (int r = r0; r < r1; r++) {
        const int u = row_utt[r / row_div];        // block-uniform
        if (u != cur) {
            flush(cur);
            cur = u;
        }
        if (u >= 0) {
            const float4 v = *(const float4 *) (x + (size_t) r * C + c4 * 4);
            s += (v.x + v.y) + (v.z + v.w);
            q += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
        }
    }
    flush(cur);
}

cudaError_t launch_gn_stats(const float * x, int rows, int C, const row_map & rm, int row_div, double * stats,
                            cudaStream_t st) {
    if (rows <= 0) return cudaSuccess;
    if (C % 128 != 0) return cudaErrorInvalidValue;
    gn_stats_kernel<<<(rows + GN_ROWS - 1) / GN_ROWS, C / 4, 0, st>>>(x, rows, C, rm.row_utt, row_div, stats);
    return cudaGetLastError();
}

__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + expf(-x)); }

// GroupNorm apply + SiLU + round to fp16 (the conv operand: ggml_conv_1d's im2col rounds activations to F16,
// ggml-cpu/ops.cpp:6232-6256, so this rounding is the reference's own, not an approximation).
__global__ void gn_apply_silu_kernel(const float * __restrict__ x, int rows, int C, const int32_t * __restrict__ row_utt,
                                     const int32_t * __restrict__ len, int row_div, const double * __restrict__ stats,
                                     float eps, const float * __restrict__ w, const float * __restrict__ b,
                                     __half * __restrict__ out) {
    const int C4 = C / 4;
    const size_t total = (size_t) rows * C4;
    const int cpg = C / 32;
    for (size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t) gridDim.x * blockDim.x) {
        const int r = (int) (idx / C4);
        const int c = (int) (idx % C4) * 4;
        const int u = row_utt[r / row_div];
        __half h[4];
        if (u < 0) {
            h[0] = h[1] = h[2] = h[3] = __float2half_rn(0.f);
        } else {
            const int g = c / cpg;
            const double n = (double) len[u] * row_div * cpg;
            const double sum = stats[((size_t) u * 32 + g) * 2], sq = stats[((size_t) u * 32 + g) * 2 + 1];
            const float mean = (float) (sum / n);
            // sum((x-mean)^2)/n with the float-rounded mean, like the reference's second pass
            const double md = (double) mean;
            const float var = (float) ((sq - 2.0 * md * sum) / n + md * md);
            const float rstd = 1.0f / sqrtf(var + eps);
            const float4 v = *(const float4 *) (x + (size_t) r * C + c);
            const float4 ww = *(const float4 *) (w + c), bb = *(const float4 *) (b + c);
            h[0] = __float2half_rn(silu_f((v.x - mean) * rstd * ww.x + bb.x));
            h[1] = __float2half_rn(silu_f((v.y - mean) * rstd * ww.y + bb.y));
            h[2] = __float2half_rn(silu_f((v.z - mean) * rstd * ww.z + bb.z));
            h[3] = __float2half_rn(silu_f((v.w - mean) * rstd * ww.w + bb.w));
        }
        *(uint2 *) (out + (size_t) r * C + c) = *(uint2 *) h;
    }
}

cudaError_t launch_gn_apply_silu(const float * x, int rows, int C, const row_map & rm, int row_div, const double * stats,
                                 float eps, const float * w, const float * b, __half * out, cudaStream_t st) {
    if (rows <= 0) return cudaSuccess;
    const size_t total = (size_t) rows * (C / 4);
    const int grid = (int) min((size_t) 148 * 16, (total + 255) / 256);
    gn_apply_silu_kernel<<<grid, 256, 0, st>>>(x, rows, C, rm.row_utt, rm.len, row_div, stats, eps, w, b, out);
    return cudaGetLastError();
}

// fp32 -> (hi, lo) fp16 operand pair, zeros on pad rows (inputs of the ConvTranspose GEMMs, which read row+-1)
__global__ void split_kernel(const float * __restrict__ x, int rows, int C, const int32_t * __restrict__ row_utt,
                             int row_div, __half * __restrict__ hi, __half * __restrict__ lo) {
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
    split_kernel<<<grid, 256, 0, st>>>(x, rows, C, rm.row_utt, row_div, hi, lo);
    return cudaGetLastError();
}

// AdaLN conditioning: mod[v][j] = b[j] + sum_k W[j][k] * silu(emb[v][k])
// (compute_adaln3, src/miocodec.cpp:310-319 and :654-655). One warp per output row j, all voices.
__global__ void cond_kernel(const float * __restrict__ voice_table, int n_voices, int adaln_dim,
                            const float * __restrict__ w, const float * __restrict__ b, int cond_rows,
                            float * __restrict__ mod) {
    const int j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (j >= cond_rows) return;
    for (int v = blockIdx.y; v < n_voices; v += gridDim.y) {
        float acc = 0.f;
        for (int k = lane; k < adaln_dim; k += 32) acc += w[(size_t) j * adaln_dim + k] * silu_f(voice_table[(size_t) v * adaln_dim + k]);
        acc = warp_sum(acc);
        if (lane == 0) mod[(size_t) v * cond_rows + j] = acc + b[j];
    }
}

cudaError_t launch_cond(const float * voice_table, int n_voices, int adaln_dim, const float * w, const float * b,
                        int cond_rows, float * mod, cudaStream_t st) {
    if (n_voices <= 0) return cudaSuccess;
    dim3 grid((cond_rows + 7) / 8, min(n_voices, 64));
    cond_kernel<<<grid, 256, 0, st>>>(voice_table, n_voices, adaln_dim, w, b, cond_rows, mod);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// per-utterance peak normalisation: gain = 0.95/max|s| if peak > 1e-8 (src/test-to-speech.cpp:289-300)
__global__ void peak_max_kernel(const float * __restrict__ wave, const int32_t * __restrict__ len,
                                const int64_t * __restrict__ out_off, int spc, float * __restrict__ peak) {
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
    peak_max_kernel<<<grid, 256, 0, st>>>(wave, len, out_off, samples_per_code, peak_scratch);
    peak_scale_kernel<<<grid, 256, 0, st>>>(wave, len, out_off, samples_per_code, peak_scratch);
    return cudaGetLastError();
}

}  // namespace mio
