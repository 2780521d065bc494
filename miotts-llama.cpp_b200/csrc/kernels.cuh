// Launchers of the sm_100a kernels.  Everything here works on the padded row space described in DESIGN.md:
// activations are row-major [rows][C] (channels contiguous), rows of all utterances of a batch are
// concatenated, each utterance in a slot of roundup(len + 1, 16) rows (zero rows behind it), times r at resolution r in
// {1, 2, 2*f0, 2*f0*f1} (the upsampler factors of the file; 1 / 2 / 6 / 18 for [3, 3]).
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>

namespace mio {

// ------------------------------------------------------------------------------------------ launches
// Programmatic dependent launch (PDL): every kernel of the decode sequence is launched with
// cudaLaunchAttributeProgrammaticStreamSerialization and executes `griddepcontrol.wait` (pdl_wait(), ptx.cuh) before it
// touches anything its predecessor wrote, then `griddepcontrol.launch_dependents`.  The NEXT kernel's CTAs are therefore
// scheduled while this kernel's last CTAs drain, and their launch latency and prologue (barrier init, TMEM allocation,
// tensor-map prefetch) run under the predecessor's tail instead of after it: ~160 launches per decode, which is most of
// what a single short utterance (BASELINE config C2) spends its time on.  MIO_PDL=0 launches plain (A/B switch).
bool pdl_enabled();
template <class... P, class... A>
inline cudaError_t launch_k(void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A &&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, P(args)...);
}

// ------------------------------------------------------------------------------------------ GEMM
enum epi_kind : int {
    EPI_STORE = 0,   // out_f32[row][col] = acc*s + bias
    EPI_RESID = 1,   // out_f32[row][col] += (acc*s + bias) * gate[utt][col]
    EPI_SWIGLU = 2,  // columns interleaved (gate, up): h = silu(g)*u -> split fp16 at col/2
    EPI_SPLIT = 3,   // y = acc*s + bias (-> snake) -> split fp16
    EPI_SNAKE = 4,   // y = snake(acc*s + bias) -> out_f32
    EPI_HEAD = 5,    // columns interleaved (log-mag, phase): (re, im) -> out_f32
    EPI_QKV = 6      // stacked q|k|v: RoPE (adjacent pairs) on q and k, q pre-scaled by 1/sqrt(64) -> split fp16
};

struct gemm_params {
    int M = 0, N = 0, K = 0, taps = 1;     // D[M][N] = sum_tap A[row + tap - taps/2][K] * B[N][tap*K + k], taps 1 / 3 / 5
    int npass = 3;                         // 3: (a_hi + a_lo)(b_hi + b_lo) without lo*lo; 1: a_hi * b_hi
    int bf16 = 0;                          // npass 1 only: both operands hold BF16 bit patterns (BF16-stored matrices)
    int mma_order = 2;                     // MIO_MMA_ORDER: 2 (default) = pass-major inside a 64-wide k-block (4 x hi*hi, 4 x
                                           // hi*lo, 4 x lo*hi), 0 / 1 = the three passes interleaved per 16-wide k step
    int chunk = 0;                         // > 0: chunked accumulation (gemm3.cu): a fresh tensor-core accumulator every chunk * 64 of K,
                                           // chunks added in fp32 registers with round-to-nearest; the layers in front of
                                           // the first conv1d use it (npass 3 only)
    int reverse = 0;                       // walk the M tiles from the last row down (generation 3): consecutive kernels
                                           // alternate, so each starts on the rows its predecessor left in L2
    float out_scale = 1.f;
    int epi = EPI_STORE;
    float * out_f32 = nullptr; int ld_out = 0;
    __half * out_hi = nullptr; __half * out_lo = nullptr; int ld_h = 0;
    const float * bias = nullptr; int bias_mod = 1;
    const float * gate = nullptr; int gate_ld = 0;        // gate[gate_voice[utt]*gate_ld + col]
    const int32_t * gate_voice = nullptr;
    const int32_t * row_utt = nullptr; int row_div = 1;   // utt = row_utt[row / row_div] (<0: pad row)
    const float * snake_ea = nullptr; const float * snake_eb = nullptr;   // exp(alpha), exp(beta) [bias_mod]
    // EPI_SPLIT only: y += resid[row][col] before the split, pad rows (row_utt < 0) written as zeros -- the last conv of a
    // resblock whose result only feeds the next product: replaces the residual reduce-add into x AND the split pass over x
    const float * resid = nullptr; int ld_resid = 0;
    // EPI_QKV: rope[pos][32][2] (cos, sin), pos = row - utt_start[utt] * row_div; columns < rope_cols are rotated,
    // columns < q_cols additionally multiplied by q_scale
    const float * rope = nullptr; int rope_max_pos = 0; const int32_t * utt_start = nullptr;
    const float * rope_t = nullptr;        // same values as [2][8][rope_max_pos][4] (position-minor), generation 3
    int rope_cols = 0, q_cols = 0; float q_scale = 1.f;
    // raw operand pointers (SIMT validation path; the tcgen05 path reads through the tensor maps)
    const __half * a_hi = nullptr; const __half * a_lo = nullptr; int lda = 0; int a_rows = 0;
    const __half * b_hi = nullptr; const __half * b_lo = nullptr;
};

// tcgen05 path (gemm3.cu): persistent CTA-pair kernel (cta_group::2, 256 x bn tiles, double-buffered TMEM accumulators),
// 8 epilogue warps that stage results in swizzled shared memory and write them with TMA stores (reduce-add for
// EPI_RESID).  tmA_*: 2-D maps over [a_rows][K] fp16, box {64, 128}; tmB_*: over [N][taps*K], box {64, bn/2};
// tmO0: fp32 output or fp16 hi, tmO1: fp16 lo; box {32, 32}, SWIZZLE_128B (fp32) / SWIZZLE_64B (fp16), row extent == M
// so the TMA unit clips the last tile.  num_sms: multiProcessorCount of the device the stream belongs to.
cudaError_t launch_gemm_tc3(const gemm_params & p, const CUtensorMap & tmA_hi, const CUtensorMap & tmA_lo,
                            const CUtensorMap & tmB_hi, const CUtensorMap & tmB_lo, const CUtensorMap & tmO0,
                            const CUtensorMap & tmO1, int bn, int num_sms, cudaStream_t st);
// CUDA-core validation kernel (gemm_simt.cu): fp32 FMAs on the same operands, libdevice transcendentals
cudaError_t launch_gemm_simt(const gemm_params & p, cudaStream_t st);
// > 48 KB dynamic shared memory is a per-DEVICE function attribute: every context calls these once for its device
cudaError_t gemm_tc3_init_device();

// ------------------------------------------------------------------------------------------ elementwise
struct row_map {                 // device pointers describing the padded row space of the current batch
    const int32_t * row_utt;     // [rows1] utterance id or -1 (code resolution)
    const int32_t * start;       // [n_utts] first row (code resolution)
    const int32_t * len;         // [n_utts] codes
    const int32_t * voice;       // [n_utts]
    int rows1;
    int n_utts;
};

cudaError_t launch_gather(const float * table, const int32_t * codes_padded, float * x, int rows, int D, int vocab,
                          int * err_flag, cudaStream_t st);
// LayerNorm over D (768 or 512). affine (w,b) for the prenet; AdaLN modulation (mod) for the decoder.
// Outputs: split fp16 (out_hi/out_lo) and/or fp32 (out_f32). Pad rows produce zeros.
cudaError_t launch_layernorm(const float * x, int rows, int D, float eps, const float * w, const float * b,
                             const float * mod, int mod_ld, int shift_off, int scale_off, const row_map & rm, int row_div,
                             __half * out_hi, __half * out_lo, float * out_f32, cudaStream_t st, bool rev = false);
// banded local attention on the rotated / scaled / split q|k|v written by EPI_QKV.  qkv_*: [rows][3*D] fp16;
// out: split fp16 [rows][D].  _tc: TMA + ldmatrix + warp MMA (attention.cu, product path); _simt: fp32 validation.
// bounds[i] = (first valid key, number of valid keys) of query row i, written once per chunk by launch_attn_bounds.
cudaError_t launch_attn_bounds(const row_map & rm, int rows, int row_div, int window, int2 * bounds, cudaStream_t st);
cudaError_t launch_attention_tc(const CUtensorMap & tm_in_hi, const CUtensorMap & tm_in_lo, const CUtensorMap & tm_out_hi,
                                const CUtensorMap & tm_out_lo, int rows, int D, int n_head, int window, const int2 * bounds,
                                cudaStream_t st, bool rev = false);
cudaError_t launch_attention_simt(const __half * qkv_hi, const __half * qkv_lo, int rows, int D, int n_head, int window,
                                  const row_map & rm, int row_div, __half * out_hi, __half * out_lo, cudaStream_t st);
cudaError_t attention_init();      // per device (dynamic shared memory attributes), called by every context
cudaError_t attention_tc_init();
cudaError_t istft_init_device();   // per device
// GroupNorm statistics, deterministic two-stage reduction: part[segment][32] = (sum, sum of squares) of one segment of
// GN_SEG_CODES codes of one utterance; seg_cu[n_utts+1] = prefix of the per-utterance segment counts
constexpr int GN_SEG_CODES = 16;        // codes per statistics segment (counted from the utterance start)
constexpr int ROW_ALIGN = 16;           // utterances start on multiples of 16 rows (code resolution), see engine::upload
cudaError_t launch_gn_stats(const float * x, int C, const row_map & rm, int row_div, const int32_t * seg_cu, int total_segs,
                            double2 * part, cudaStream_t st, bool rev = false);
// y = fp16( silu( (x-mean)*rstd*w + b ) ), zeros on pad rows
// (two launches: segments -> per-(utterance, group) mean / rstd into mr_scratch[n_utts*32], then the per-element pass)
cudaError_t launch_gn_apply_silu(const float * x, int rows, int C, const row_map & rm, int row_div, const double2 * part,
                                 const int32_t * seg_cu, float eps, const float * w, const float * b, float2 * mr_scratch,
                                 __half * out, cudaStream_t st, bool rev = false);
cudaError_t launch_split(const float * x, int rows, int C, const row_map & rm, int row_div, __half * hi, __half * lo,
                         cudaStream_t st);
// mod[voice][cond_rows] = W[cond_rows][128] * silu(emb[voice]) + b
// row_store[cond_rows] (or nullptr): 1 / 2 where the row's matrix is stored as F16 / BF16 -> silu(emb) rounded to that type
cudaError_t launch_cond(const float * voice_table, int n_voices, int adaln_dim, const float * w, const float * b,
                        int cond_rows, const uint8_t * row_store, float * silu_scratch /*[n_voices*adaln_dim]*/, float * mod,
                        cudaStream_t st);
// (hi, lo) fp16 operand pair -> BF16 bit patterns in the hi buffer, in place (BF16-stored matrices); n even
cudaError_t launch_pair_to_bf16(__half * hi, const __half * lo, size_t n, cudaStream_t st);

// ------------------------------------------------------------------------------------------ ISTFT
struct istft_params {
    const float * spec;          // [rows18][spec_ld], 394 floats used per row
    int spec_ld;                 // row pitch in floats (even)
    float * wave;                // output samples
    const int32_t * start;       // [n_utts] first row at code resolution (frames start at fpc*start)
    const int32_t * len;         // [n_utts] codes
    const int64_t * out_off;     // [n_utts] sample offset in wave
    const int32_t * chunk_cu;    // [n_utts+1] prefix sum of per-utterance chunk counts
    const float * hann;          // [392]
    const float * tw;            // [784] twiddles (istft_twiddles), device
    int n_utts;
    int frames_per_code;         // 18
    int total_chunks;
};
constexpr int ISTFT_SEG_PER_BLOCK = 44;   // output hops (98 samples) per block; +4 halo frames
void istft_twiddles(float * host784);
cudaError_t launch_istft(const istft_params & p, cudaStream_t st);
// per-utterance peak normalisation 0.95/max|s| (src/test-to-speech.cpp:289-300)
cudaError_t launch_peak_normalize(float * wave, const int32_t * len, const int64_t * out_off, int n_utts,
                                  int samples_per_code, float * peak_scratch, cudaStream_t st);
// float -> PCM16 with the reference WAV writer's clamp + truncation (src/wav-writer.cpp:39-42)
cudaError_t launch_pcm16(const float * wave, int16_t * pcm, int64_t begin, int64_t end, cudaStream_t st);

}  // namespace mio
