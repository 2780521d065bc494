// Host-side model description: hyper-parameters, device-resident weights, workspace layout.
// Replaces miocodec_context (src/miocodec.cpp:50-90 of the reference) — but weights live in HBM, pre-split
// into fp16 hi/lo pairs for the tcgen05 GEMMs, instead of a host byte cache that is re-copied per call.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "gguf_reader.h"

namespace mio {

struct hparams {                      // defaults: src/miocodec.cpp:59-85
    int sample_rate = 44100, n_fft = 392, hop = 98, n_freq = 197, samples_per_token = 1764, head_out = 394;
    int prenet_layers = 6, prenet_dim = 768, prenet_heads = 12, prenet_ff = 2048, prenet_window = 65;
    int dec_layers = 8, dec_dim = 512, dec_heads = 8, dec_ff = 1536, dec_window = 65, adaln_dim = 128;
    int resnet_blocks = 2, resnet_groups = 32, up_stages = 2;
    float rope_theta = 10000.f, norm_eps = 1e-5f, gn_eps = 1e-6f;
    int vocab = 12800;
    int up_factor[4] = {3, 3, 1, 1}, up_kernel[4] = {9, 9, 1, 1}, up_ch[4] = {256, 128, 0, 0};
    int up_res[4] = {6, 18, 0, 0};   // rows per code after stage s: 2 * f_0 * ... * f_s
    int out_proj_dim = 512;
    int frames_per_code = 18;
};

// One GEMM weight: B[N][Ktot] K-major fp16, hi (+ lo when the layer needs fp32-class accuracy).
struct gemm_weight {
    int N = 0, K = 0, taps = 1;       // Ktot = taps*K
    int npass = 3;                    // 3 = split precision (hi*hi + hi*lo + lo*hi), 1 = fp16 operands (conv1d)
    int store = 0;                    // how the file stores the matrix: 0 F32, 1 F16, 2 BF16.  1 / 2: the reference rounds the
                                      // activations to that type and multiplies once (ggml-cpu.c:1291-1326) -> npass 1, and
                                      // for 2 both operands hold BF16 bit patterns
    float out_scale = 1.f;            // epilogue multiplier (inverse of the power-of-two weight pre-scale)
    __half * hi = nullptr;
    __half * lo = nullptr;
    CUtensorMap tm2_hi, tm2_lo;       // box {64, bn2 / 2}     (CTA pair per 256 x bn2 tile: each CTA loads half of B)
    int bn2 = 0;
    // narrower tiles for small M (a single short utterance has too few 256 x bn2 tiles to fill 74 CTA pairs):
    // box {64, 64} and {64, 32} for 256 x 128 and 256 x 64 tiles
    CUtensorMap tm128_hi, tm128_lo, tm64_hi, tm64_lo;
};

struct resblock_w {
    float *n1w, *n1b, *c1b, *n2w, *n2b, *c2b;
    gemm_weight c1, c2;
    int C;
};

struct prenet_layer_w {
    float *an_w, *an_b, *fn_w, *fn_b;
    gemm_weight qkv, wo, gu, down;
};

struct dec_layer_w {
    gemm_weight qkv, wo, gu, down;
};

struct weights {
    float * token_embd = nullptr;                 // [vocab][768] fp32
    std::vector<prenet_layer_w> prenet;
    float *pre_norm_w, *pre_norm_b, *pre_out_b;
    gemm_weight pre_out;                          // 768 -> 512
    gemm_weight up0; float * up0_b;               // ConvT k2 s2 as GEMM N=2*512
    std::vector<resblock_w> prior, post, up_res;
    std::vector<dec_layer_w> dec;
    // all AdaLN conditioning linears stacked: rows = 16*1536 + 1024, K = 128 (fp32; tiny GEMV per utterance)
    float * cond_w = nullptr; float * cond_b = nullptr; int cond_rows = 0;
    uint8_t * cond_store = nullptr;               // [cond_rows] storage type of the matrix each row came from (gemm_weight::store)
    std::vector<gemm_weight> upT; std::vector<float *> upT_b, snake_ea, snake_eb;   // per stage
    gemm_weight out_proj; float *out_proj_b, *out_snake_ea, *out_snake_eb;
    gemm_weight head; float * head_b;             // columns interleaved (logmag_f, phase_f)
    float * rope_pre = nullptr;                   // [max_pos][32][2] cos,sin
    float * rope_t = nullptr;         // the same table position-minor: [2][8][rope_max_pos][4] (generation-3 QKV epilogue)
    int rope_max_pos = 0;
    float * hann = nullptr;                       // [n_fft]
    float * istft_tw = nullptr;                   // [784] ISTFT twiddles (istft_twiddles)
};

// Padded row space (DESIGN.md "row space"): utterance u owns rows [start[u], start[u]+len[u]) at code
// resolution followed by ONE zero row; at resolution r everything is multiplied by r.
struct batch_layout {
    int n_utts = 0;
    int rows1 = 0;                 // sum(len) + n_utts
    int64_t total_codes = 0;
    std::vector<int32_t> start, len, voice, row_utt, codes_padded;
    std::vector<int64_t> out_off;  // sample offset of each utterance in the output wave
};

}  // namespace mio
