#include "engine.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

namespace mio {

#define CK(expr)                                                                                       \
    do {                                                                                               \
        cudaError_t _e = (expr);                                                                       \
        if (_e != cudaSuccess) return std::string(#expr) + ": " + cudaGetErrorString(_e);              \
    } while (0)

constexpr int SPEC_LD = 400;   // spectrogram row pitch in floats: 394 padded to a 16-byte multiple (TMA store pitch rule)

// ---------------------------------------------------------------------------------------------
// TMA descriptors (driver entry point fetched at run time: the library has no link-time libcuda dependency, so it
// loads — and its symbol table can be checked — on machines without a driver)
// ---------------------------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static encode_tiled_fn g_encode = nullptr;

static std::string get_encode() {
    if (g_encode) return {};
    void * fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    if (!fn || q != cudaDriverEntryPointSuccess) return "cuTensorMapEncodeTiled not available";
    g_encode = (encode_tiled_fn) fn;
    return {};
}

// 2-D fp16 [rows][cols] row-major, box {64, box_rows}, SWIZZLE_128B, OOB -> zeros
static std::string make_map(CUtensorMap * tm, const __half * base, int rows, int cols, int box_rows) {
    std::string e = get_encode();
    if (!e.empty()) return e;
    cuuint64_t dims[2] = {(cuuint64_t) cols, (cuuint64_t) rows};
    cuuint64_t strides[1] = {(cuuint64_t) cols * sizeof(__half)};
    cuuint32_t box[2] = {64u, (cuuint32_t) box_rows};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = g_encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (void *) base, dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return "cuTensorMapEncodeTiled failed (" + std::to_string((int) r) + ")";
    return {};
}

// output tile map of the generation-3 GEMM epilogue: [rows][cols] elements of `esize` bytes with row pitch ld_bytes,
// box {32, 32}; fp32 -> SWIZZLE_128B (128-byte box rows), fp16 -> SWIZZLE_64B.  rows == M: the TMA unit clips.
static std::string make_out_map(CUtensorMap * tm, const void * base, int esize, int rows, int cols, size_t ld_bytes,
                                int box_c, int box_r) {
    std::string e = get_encode();
    if (!e.empty()) return e;
    if (ld_bytes % 16 != 0 || ((uintptr_t) base & 15) != 0) return "internal: GEMM output not 16-byte aligned for TMA";
    cuuint64_t dims[2] = {(cuuint64_t) cols, (cuuint64_t) rows};
    cuuint64_t strides[1] = {(cuuint64_t) ld_bytes};
    cuuint32_t box[2] = {(cuuint32_t) box_c, (cuuint32_t) box_r};
    cuuint32_t estr[2] = {1u, 1u};
    if (box_c * esize != 128 && box_c * esize != 64) return "internal: unsupported TMA box width";
    CUresult r = g_encode(tm, esize == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (void *) base,
                          dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          box_c * esize == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                          CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return "cuTensorMapEncodeTiled (output) failed (" + std::to_string((int) r) + ")";
    return {};
}

// Returns the descriptor BY VALUE (128 bytes): the cache may evict at any call without invalidating what earlier
// calls of the same launch handed out.  Keys include the row count M, so a streaming session (a new M nearly every
// round) keeps inserting; the cache is simply emptied when it reaches 4096 entries.
bool engine::out_map(CUtensorMap * out, const void * base, int esize, int rows, int cols, size_t ld_bytes, int box_c, int box_r) {
    out_map_key k{base, esize, rows, cols, ld_bytes, box_c, box_r};
    auto it = out_maps.find(k);
    if (it != out_maps.end()) { *out = it->second; return true; }
    std::string e = make_out_map(out, base, esize, rows, cols, ld_bytes, box_c, box_r);
    if (!e.empty()) { err = e; return false; }
    if (out_maps.size() >= 4096) out_maps.clear();
    out_maps[k] = *out;
    return true;
}

// ---------------------------------------------------------------------------------------------
// small device helpers used at load time
// ---------------------------------------------------------------------------------------------
__global__ void split_weights_kernel(const float * __restrict__ w, size_t n, float scale, __half * __restrict__ hi,
                                     __half * __restrict__ lo) {
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        const float x = w[i] * scale;
        const __half h = __float2half_rn(x);
        hi[i] = h;
        if (lo) lo[i] = __float2half_rn(x - __half2float(h));
    }
}

// BF16-stored matrix: its values are BF16 numbers, kept as bit patterns in the fp16-typed operand buffer
__global__ void bf16_weights_kernel(const float * __restrict__ w, size_t n, __half * __restrict__ out) {
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x)
        out[i] = __ushort_as_half((unsigned short) (__float_as_uint(w[i]) >> 16));
}

template <class T>
static std::string dev_alloc(T ** p, size_t n) {
    CK(cudaMalloc((void **) p, std::max<size_t>(n, 1) * sizeof(T)));
    return {};
}

static std::string upload_f32(const std::vector<float> & h, float ** d) {
    std::string e = dev_alloc(d, h.size());
    if (!e.empty()) return e;
    CK(cudaMemcpy(*d, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
    return {};
}

engine::~engine() {
    // process-lifetime allocations are released with the context; individual frees omitted for weights created
    // through cudaMalloc are handled here for the big buffers
    cudaSetDevice(device);
    if (stream && own_stream) cudaStreamSynchronize(stream);
    auto F = [](void * p) { if (p) cudaFree(p); };
    F(x768); F(qkv_hi); F(qkv_lo); F(x512); F(y512); F(xu0); F(yu0); F(xu1); F(yu1); F(spec); F(mod); F(peak);
    F(aA_hi); F(aA_lo); F(aFF_hi); F(aFF_lo); F(c16); F(hH_hi); F(hH_lo); F(gn_part); F(gn_mr); F(voice_silu); F(attn_b1); F(attn_b2); F(err_flag);
    F(d_codes); F(d_row_utt); F(d_start); F(d_len); F(d_voice); F(d_chunk_cu); F(d_seg_cu); F(d_out_off); F(d_voice_table); F(d_wave);
    for (void * p : w_allocs) F(p);
    F(d_emit); F(d_emit_jobs); F(d_pcm); F(w.rope_pre); F(w.rope_t); F(istft_spec);
    for (emit_block * b : emit_pool) { if (b->host) cudaFreeHost(b->host); delete b; }
    if (h_stage) cudaFreeHost(h_stage);
    if (stream && own_stream) cudaStreamDestroy(stream);
    if (chunk_done) cudaEventDestroy(chunk_done);
    for (cudaEvent_t ev : tune_ev) if (ev) cudaEventDestroy(ev);
    if (copy_stream) cudaStreamDestroy(copy_stream);
}

// A free pinned block of at least `samples` floats.  *pooled = false: the pool is exhausted (every block still has unread
// slices); the caller gets the scratch block (last entry, never referenced) and must copy the samples out of it.
engine::emit_block * engine::acquire_emit_block(int64_t samples, bool * pooled) {
    emit_block * best = nullptr;
    int in_use = 0;
    for (emit_block * b : emit_pool) {
        if (b->refs > 0) { in_use++; continue; }
        if (b->cap >= samples && (!best || b->cap < best->cap)) best = b;
    }
    *pooled = in_use < EMIT_POOL_MAX;
    if (best) return best;
    // grow: reuse (reallocate) a free block that is too small before adding a new one
    emit_block * blk = nullptr;
    for (emit_block * b : emit_pool) if (b->refs == 0) { blk = b; break; }
    if (!blk) { blk = new emit_block(); emit_pool.push_back(blk); }
    if (blk->host) { cudaFreeHost(blk->host); blk->host = nullptr; blk->cap = 0; }
    const int64_t want = samples + samples / 4 + 4096;
    if (cudaMallocHost((void **) &blk->host, (size_t) want * sizeof(float)) != cudaSuccess) { blk->host = nullptr; return nullptr; }
    blk->cap = want;
    return blk;
}

// ---------------------------------------------------------------------------------------------
// load: GGUF -> HBM.  KV keys and defaults follow miocodec_load (src/miocodec.cpp:447-481).
// ---------------------------------------------------------------------------------------------
static std::string make_view(a_view & v, __half * hi, __half * lo, int rows, int K);

namespace {
struct loader {
    engine & E;
    gguf_file & g;
    std::string err;
    loader(engine & e, gguf_file & gg) : E(e), g(gg) {}

    bool get(const std::string & name, std::vector<float> & out, std::vector<int64_t> * ne = nullptr) {
        const gguf_tensor * t = g.find(name);
        if (!t) { err = "missing tensor: " + name; return false; }
        if (!g.read_f32(*t, out)) { err = "unsupported dtype / short read for tensor " + name; return false; }
        if (ne) ne->assign(t->ne, t->ne + 4);
        return true;
    }
    // gemm_weight::store of a matrix: 0 F32, 1 F16, 2 BF16
    int store_of(const std::string & name) {
        const gguf_tensor * t = g.find(name);
        return !t ? 0 : t->dtype == GT_F16 ? 1 : t->dtype == GT_BF16 ? 2 : 0;
    }
    float * vec(const std::string & name, size_t expect) {
        std::vector<float> h;
        if (!get(name, h)) return nullptr;
        if (h.size() != expect) { err = "tensor " + name + " has wrong size"; return nullptr; }
        float * d = nullptr;
        std::string e = upload_f32(h, &d);
        if (!e.empty()) { err = e; return nullptr; }
        E.w_allocs.push_back(d);
        return d;
    }
    float * vec_host(const std::vector<float> & h) {
        float * d = nullptr;
        std::string e = upload_f32(h, &d);
        if (!e.empty()) { err = e; return nullptr; }
        E.w_allocs.push_back(d);
        return d;
    }
    // B[N][Ktot] fp32 on the host -> split fp16 on the device (+ tensor maps)
    // store 1 / 2 (F16 / BF16-stored matrix): one pass on operands of that type; the values are exactly representable
    bool pack(gemm_weight & W, const std::vector<float> & B, int N, int K, int taps, int npass, int store = 0) {
        if (store != 0) npass = 1;
        // refused here, at load time, with a message -- not as a bare cudaErrorInvalidValue from gemm() in the middle of a decode
        if (store == 2 && taps != 1) { err = "unsupported: a BF16-stored matrix cannot be a " + std::to_string(taps) + "-tap convolution kernel (F32 or F16 only)"; return false; }
        W.N = N; W.K = K; W.taps = taps; W.npass = npass; W.store = store;
        const size_t n = (size_t) N * K * taps;
        if (B.size() != n) { err = "internal: weight size mismatch"; return false; }
        float scale = 1.f;
        if (npass == 3) {   // power-of-two pre-scale keeps the lo parts in fp16's normal range (exact to undo)
            float mx = 0.f;
            for (float v : B) mx = std::max(mx, std::fabs(v));
            if (mx > 0.f && std::isfinite(mx)) {
                int e = (int) std::floor(std::log2(256.0 / (double) mx));
                e = std::max(-8, std::min(20, e));
                scale = std::ldexp(1.0f, e);
            }
        }
        W.out_scale = 1.0f / scale;
        float * tmp = nullptr;
        if (cudaMalloc(&tmp, n * sizeof(float)) != cudaSuccess) { err = "cudaMalloc failed (weight staging)"; return false; }
        cudaMemcpy(tmp, B.data(), n * sizeof(float), cudaMemcpyHostToDevice);
        if (cudaMalloc(&W.hi, n * sizeof(__half)) != cudaSuccess) { err = "cudaMalloc failed (weights)"; return false; }
        E.w_allocs.push_back(W.hi);
        if (npass == 3) {
            if (cudaMalloc(&W.lo, n * sizeof(__half)) != cudaSuccess) { err = "cudaMalloc failed (weights)"; return false; }
            E.w_allocs.push_back(W.lo);
        }
        if (store == 2) bf16_weights_kernel<<<296, 256>>>(tmp, n, W.hi);
        else split_weights_kernel<<<296, 256>>>(tmp, n, scale, W.hi, W.lo);
        cudaError_t ce = cudaDeviceSynchronize();
        cudaFree(tmp);
        if (ce != cudaSuccess) { err = std::string("split_weights_kernel: ") + cudaGetErrorString(ce); return false; }
        std::string e;
        // 256-wide tiles whenever they pad N no further than 128-wide ones would (e.g. the head, N = 394 -> 512 either
        // way): half as many passes over the A operand
        W.bn2 = ((N + 255) / 256 * 256 == (N + 127) / 128 * 128) ? 256 : 128;
        if (e.empty()) e = make_map(&W.tm2_hi, W.hi, N, K * taps, W.bn2 / 2);
        if (e.empty() && W.lo) e = make_map(&W.tm2_lo, W.lo, N, K * taps, W.bn2 / 2);
        if (e.empty() && !W.lo) W.tm2_lo = W.tm2_hi;
        if (e.empty()) e = make_map(&W.tm128_hi, W.hi, N, K * taps, 64);
        if (e.empty() && W.lo) e = make_map(&W.tm128_lo, W.lo, N, K * taps, 64);
        if (e.empty() && !W.lo) W.tm128_lo = W.tm128_hi;
        if (e.empty()) e = make_map(&W.tm64_hi, W.hi, N, K * taps, 32);
        if (e.empty() && W.lo) e = make_map(&W.tm64_lo, W.lo, N, K * taps, 32);
        if (e.empty() && !W.lo) W.tm64_lo = W.tm64_hi;
        if (!e.empty()) { err = e; return false; }
        return true;
    }
    // Linear weight(s): ggml ne [in, out] == row-major [out][in]; several are stacked along N
    bool linear(gemm_weight & W, const std::vector<std::string> & names, int K, bool interleave2 = false) {
        std::vector<std::vector<float>> parts(names.size());
        int N = 0;
        const int store = store_of(names[0]);
        for (size_t i = 0; i < names.size(); i++) {
            std::vector<int64_t> ne;
            if (!get(names[i], parts[i], &ne)) return false;
            if (ne[0] != K) { err = "tensor " + names[i] + ": unexpected input width"; return false; }
            N += (int) ne[1];
            // the matrices of one stacked projection run as ONE product, so they must agree on the operand type
            if (store_of(names[i]) != store) { err = "tensors " + names[0] + " and " + names[i] + " are stored in different types (F32/F16/BF16); matrices of one stacked projection must share a type"; return false; }
        }
        std::vector<float> B((size_t) N * K);
        if (interleave2) {   // two equally sized matrices, rows interleaved (gate/up)
            const int half = N / 2;
            for (int j = 0; j < half; j++) {
                memcpy(&B[(size_t) (2 * j) * K], &parts[0][(size_t) j * K], K * sizeof(float));
                memcpy(&B[(size_t) (2 * j + 1) * K], &parts[1][(size_t) j * K], K * sizeof(float));
            }
        } else {
            size_t off = 0;
            for (auto & p : parts) { memcpy(&B[off], p.data(), p.size() * sizeof(float)); off += p.size(); }
        }
        // F16/BF16-stored matrices: the reference rounds the activations to that type and multiplies once
        // (ggml-cpu.c:1291-1326); pack() then selects the single-pass product on operands of that type
        return pack(W, B, N, K, 1, 3, store);
    }
    // Conv1d k=3: ggml ne [3, IC, OC] == row-major [OC][IC][3] -> B[oc][tap*IC + ic]; operands rounded to fp16 (npass 1)
    bool conv3(gemm_weight & W, const std::string & name, int C) {
        std::vector<float> h; std::vector<int64_t> ne;
        if (!get(name, h, &ne)) return false;
        if (ne[0] != 3 || ne[1] != C || ne[2] != C) { err = "tensor " + name + ": expected [3," + std::to_string(C) + "," + std::to_string(C) + "]"; return false; }
        std::vector<float> B((size_t) C * 3 * C);
        for (int oc = 0; oc < C; oc++)
            for (int ic = 0; ic < C; ic++)
                for (int t = 0; t < 3; t++) B[(size_t) oc * 3 * C + (size_t) t * C + ic] = h[((size_t) oc * C + ic) * 3 + t];
        return pack(W, B, C, C, 3, 1);
    }
    // ConvTranspose1d: ggml ne [k, Cout, Cin] == row-major [Cin][Cout][k]; stride f, then trim = (k-f)/2 samples dropped
    // at each end (src/miocodec.cpp:683-696), i.e. PyTorch's ConvTranspose1d(padding=trim).  As a tap-GEMM over the
    // INPUT rows, output row f*q + r (r < f) of N = f*Cout columns:
    //   y[f*q + r][co] = b[co] + sum_{d=-D..D} sum_ci x[q+d][ci] * W[ci][co][r + trim - f*d]     (kernel index in [0, k))
    // with D = floor((f - 1 + trim) / f): 0 for k == f, 1 for k <= 3f, 2 for k <= 5f.  Kernel positions a (r, d) pair
    // does not have are zero columns of B; input rows outside the utterance are the zero pad rows of the row space.
    bool convT(gemm_weight & W, const std::string & name, int Cin, int Cout, int k, int f) {
        std::vector<float> h; std::vector<int64_t> ne;
        if (!get(name, h, &ne)) return false;
        if (ne[0] != k || ne[1] != Cout || ne[2] != Cin) { err = "tensor " + name + ": unexpected ConvTranspose shape"; return false; }
        if (f < 1 || k < f || (k - f) % 2 != 0) {
            // the reference's trim (k-f)/2 per side only restores the length f*L when k - f is even (src/miocodec.cpp:689-696)
            err = "ConvTranspose1d " + name + ": kernel " + std::to_string(k) + " / stride " + std::to_string(f) + " needs k >= f and (k - f) even";
            return false;
        }
        const int trim = (k - f) / 2;
        const int D = (f - 1 + trim) / f;
        if (D > 2) { err = "ConvTranspose1d " + name + ": kernel " + std::to_string(k) + " wider than 5 x stride " + std::to_string(f) + " is not supported"; return false; }
        const int taps = 2 * D + 1;
        const int N = f * Cout;
        const int store = store_of(name);
        // ggml has no BF16 ConvTranspose1d kernel either (ops.cpp:6069-6082 dispatches on F16 / F32 only and aborts)
        if (store == 2) { err = "ConvTranspose1d " + name + ": BF16 kernels are not supported (the reference aborts on them too)"; return false; }
        std::vector<float> B((size_t) N * taps * Cin, 0.f);
        for (int r = 0; r < f; r++)
            for (int co = 0; co < Cout; co++)
                for (int t = 0; t < taps; t++) {
                    const int d = t - D;
                    const int kk = r + trim - f * d;
                    if (kk < 0 || kk >= k) continue;
                    float * dst = &B[((size_t) (r * Cout + co) * taps + t) * Cin];
                    for (int ci = 0; ci < Cin; ci++) dst[ci] = h[((size_t) ci * Cout + co) * k + kk];
                }
        // F16-stored kernel: activations rounded to F16, one pass (ggml-cpu/ops.cpp:5909-5995)
        return pack(W, B, N, Cin, taps, 3, store);
    }
    bool resblock(resblock_w & R, const std::string & p, int C) {
        R.C = C;
        R.n1w = vec(p + "norm1.weight", C); R.n1b = vec(p + "norm1.bias", C);
        R.n2w = vec(p + "norm2.weight", C); R.n2b = vec(p + "norm2.bias", C);
        R.c1b = vec(p + "conv1.bias", C);   R.c2b = vec(p + "conv2.bias", C);
        if (!R.n1w || !R.n1b || !R.n2w || !R.n2b || !R.c1b || !R.c2b) return false;
        return conv3(R.c1, p + "conv1.weight", C) && conv3(R.c2, p + "conv2.weight", C);
    }
};
}  // namespace

std::string engine::load(const std::string & path, int dev) {
    device = dev;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) return "no CUDA device available (this library has no CPU fallback)";
    if (dev < 0 || dev >= n_dev) return "invalid CUDA device ordinal";
    CK(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, dev));
    if (prop.major != 10) return std::string("device '") + prop.name + "' is not sm_100 (B200); this build contains sm_100a code only";
    num_sms = prop.multiProcessorCount;

    gguf_file g;
    std::string e = g.open(path);
    if (!e.empty()) return e;
    // hyper-parameters (src/miocodec.cpp:447-474)
    hp.sample_rate = g.u32_or("miocodec.sample_rate", 44100);
    hp.n_fft = g.u32_or("miocodec.n_fft", 392);
    hp.hop = g.u32_or("miocodec.hop_length", 98);
    hp.n_freq = hp.n_fft / 2 + 1;
    hp.samples_per_token = g.u32_or("miocodec.samples_per_token", 1764);
    hp.head_out = g.u32_or("embedding_length_out", 394);
    hp.prenet_layers = g.u32_or("miocodec.prenet_layers", 6);
    hp.prenet_dim = g.u32_or("miocodec.prenet_dim", 768);
    hp.prenet_heads = g.u32_or("miocodec.prenet_heads", 12);
    hp.prenet_ff = g.u32_or("miocodec.prenet_ff", 2048);
    hp.prenet_window = g.u32_or("miocodec.prenet_window", 65);
    hp.dec_layers = g.u32_or("miocodec.decoder_layers", 8);
    hp.dec_dim = g.u32_or("miocodec.decoder_dim", 512);
    hp.dec_heads = g.u32_or("miocodec.decoder_heads", 8);
    hp.dec_ff = g.u32_or("miocodec.decoder_ff", 1536);
    hp.dec_window = g.u32_or("miocodec.decoder_window", 65);
    hp.adaln_dim = g.u32_or("miocodec.decoder_adanorm_dim", 128);
    hp.resnet_blocks = g.u32_or("miocodec.resnet_blocks", 2);
    hp.resnet_groups = g.u32_or("miocodec.resnet_groups", 32);
    hp.up_stages = g.u32_or("miocodec.wave_upsampler_layers", 2);
    hp.rope_theta = g.f32_or("miocodec.rope_theta", 10000.0f);
    hp.norm_eps = g.f32_or("miocodec.norm_eps", 1e-5f);
    hp.gn_eps = g.f32_or("miocodec.group_norm_eps", 1e-6f);

    // shapes this build has kernels for (the reference's fixed architecture)
    if (hp.n_fft != 392 || hp.hop != 98 || hp.head_out != 394) return "unsupported STFT geometry (kernels are built for n_fft=392, hop=98)";
    if (hp.prenet_dim != 768 || hp.dec_dim != 512 || hp.prenet_dim / hp.prenet_heads != 64 || hp.dec_dim / hp.dec_heads != 64)
        return "unsupported transformer widths (kernels are built for 768/512 channels, head_dim 64)";
    if (hp.prenet_window != 65 || hp.dec_window != 65) return "unsupported attention window (kernels are built for 65)";
    if (hp.resnet_groups != 32 || hp.up_stages != 2) return "unsupported resnet groups / upsampler depth";
    if (hp.prenet_ff % 64 || hp.dec_ff % 64 || hp.adaln_dim % 4) return "unsupported feed-forward / conditioning widths";

    loader L(*this, g);
    // upsampler geometry from the I32 tensors (src/miocodec.cpp:476-481)
    {
        const gguf_tensor * tf = g.find("miocodec.wave_upsampler.factors");
        const gguf_tensor * tk = g.find("miocodec.wave_upsampler.kernel_sizes");
        if (!tf || !tk || tf->dtype != GT_I32 || tk->dtype != GT_I32 || tf->nelements() < hp.up_stages || tk->nelements() < hp.up_stages)
            return "missing/invalid miocodec.wave_upsampler.factors / kernel_sizes";
        int32_t f[4] = {0}, k[4] = {0};
        if (!g.read_raw(*tf, f, sizeof(int32_t) * hp.up_stages) || !g.read_raw(*tk, k, sizeof(int32_t) * hp.up_stages)) return "short read (upsampler geometry)";
        int total = 2;     // wave_upsample doubles the rate first (stride 2 is fixed in the graph, src/miocodec.cpp:624)
        for (int s = 0; s < hp.up_stages; s++) {
            if (f[s] < 1 || f[s] > 64 || k[s] < 1) return "invalid upsampler factor / kernel size";
            hp.up_factor[s] = f[s]; hp.up_kernel[s] = k[s]; total *= f[s];
            hp.up_res[s] = total;
        }
        hp.frames_per_code = total;
        if (hp.frames_per_code > 4096) return "invalid upsampler factors";
        if (hp.frames_per_code * hp.hop != hp.samples_per_token) return "inconsistent samples_per_token (2 * prod(factors) * hop_length)";
    }
    {
        const gguf_tensor * te = g.find("token_embd");
        if (!te) return "missing tensor: token_embd";
        if (te->ne[0] != hp.prenet_dim) return "token_embd has the wrong width";
        hp.vocab = (int) te->ne[1];
        w.token_embd = L.vec("token_embd", (size_t) hp.vocab * hp.prenet_dim);
        if (!w.token_embd) return L.err;
    }
    const int D = hp.prenet_dim, C = hp.dec_dim;
    w.prenet.resize(hp.prenet_layers);
    for (int i = 0; i < hp.prenet_layers; i++) {
        const std::string p = "wave_prenet.blk." + std::to_string(i) + ".";
        prenet_layer_w & l = w.prenet[i];
        l.an_w = L.vec(p + "attn_norm.weight", D); l.an_b = L.vec(p + "attn_norm.bias", D);
        l.fn_w = L.vec(p + "ffn_norm.weight", D);  l.fn_b = L.vec(p + "ffn_norm.bias", D);
        if (!l.an_w || !l.an_b || !l.fn_w || !l.fn_b) return L.err;
        if (!L.linear(l.qkv, {p + "attn_q.weight", p + "attn_k.weight", p + "attn_v.weight"}, D)) return L.err;
        if (!L.linear(l.wo, {p + "attn_output.weight"}, D)) return L.err;
        if (!L.linear(l.gu, {p + "ffn_gate.weight", p + "ffn_up.weight"}, D, true)) return L.err;
        if (!L.linear(l.down, {p + "ffn_down.weight"}, hp.prenet_ff)) return L.err;
    }
    w.pre_norm_w = L.vec("wave_prenet.norm.weight", D); w.pre_norm_b = L.vec("wave_prenet.norm.bias", D);
    w.pre_out_b = L.vec("wave_prenet.output.bias", C);
    if (!w.pre_norm_w || !w.pre_norm_b || !w.pre_out_b) return L.err;
    if (!L.linear(w.pre_out, {"wave_prenet.output.weight"}, D)) return L.err;
    if (!L.convT(w.up0, "wave_upsample.weight", C, C, 2, 2)) return L.err;
    w.up0_b = L.vec("wave_upsample.bias", C);
    if (!w.up0_b) return L.err;
    w.prior.resize(hp.resnet_blocks); w.post.resize(hp.resnet_blocks);
    for (int b = 0; b < hp.resnet_blocks; b++) {
        if (!L.resblock(w.prior[b], "wave_prior." + std::to_string(b) + ".", C)) return L.err;
        if (!L.resblock(w.post[b], "wave_post." + std::to_string(b) + ".", C)) return L.err;
    }
    // decoder + stacked conditioning
    w.dec.resize(hp.dec_layers);
    w.cond_rows = hp.dec_layers * 2 * 3 * C + 2 * C;
    std::vector<float> cw((size_t) w.cond_rows * hp.adaln_dim), cb(w.cond_rows);
    std::vector<uint8_t> cstore(w.cond_rows, 0);     // F16/BF16-stored conditioning matrices: silu(emb) rounded to that type
    for (int i = 0; i < hp.dec_layers; i++) {
        const std::string p = "wave_decoder.blk." + std::to_string(i) + ".";
        dec_layer_w & l = w.dec[i];
        if (!L.linear(l.qkv, {p + "attn_q.weight", p + "attn_k.weight", p + "attn_v.weight"}, C)) return L.err;
        if (!L.linear(l.wo, {p + "attn_output.weight"}, C)) return L.err;
        if (!L.linear(l.gu, {p + "ffn_gate.weight", p + "ffn_up.weight"}, C, true)) return L.err;
        if (!L.linear(l.down, {p + "ffn_down.weight"}, hp.dec_ff)) return L.err;
        const char * nm[2] = {"attn_cond", "ffn_cond"};
        for (int j = 0; j < 2; j++) {
            std::vector<float> hw, hb;
            if (!L.get(p + nm[j] + ".weight", hw) || !L.get(p + nm[j] + ".bias", hb)) return L.err;
            if (hw.size() != (size_t) 3 * C * hp.adaln_dim || hb.size() != (size_t) 3 * C) return "bad cond tensor size in " + p;
            const size_t row0 = (size_t) (i * 2 + j) * 3 * C;
            memcpy(&cw[row0 * hp.adaln_dim], hw.data(), hw.size() * sizeof(float));
            memcpy(&cb[row0], hb.data(), hb.size() * sizeof(float));
            std::fill(cstore.begin() + row0, cstore.begin() + row0 + 3 * C, (uint8_t) L.store_of(p + nm[j] + ".weight"));
        }
    }
    {
        std::vector<float> hw, hb;
        if (!L.get("wave_decoder.norm_cond.weight", hw) || !L.get("wave_decoder.norm_cond.bias", hb)) return L.err;
        if (hw.size() != (size_t) 2 * C * hp.adaln_dim || hb.size() != (size_t) 2 * C) return "bad norm_cond tensor size";
        const size_t row0 = (size_t) hp.dec_layers * 2 * 3 * C;
        memcpy(&cw[row0 * hp.adaln_dim], hw.data(), hw.size() * sizeof(float));
        memcpy(&cb[row0], hb.data(), hb.size() * sizeof(float));
        std::fill(cstore.begin() + row0, cstore.end(), (uint8_t) L.store_of("wave_decoder.norm_cond.weight"));
    }
    w.cond_w = L.vec_host(cw); w.cond_b = L.vec_host(cb);
    if (!w.cond_w || !w.cond_b) return L.err;
    CK(cudaMalloc((void **) &w.cond_store, cstore.size()));
    w_allocs.push_back(w.cond_store);
    CK(cudaMemcpy(w.cond_store, cstore.data(), cstore.size(), cudaMemcpyHostToDevice));

    // upsampler stages
    w.upT.resize(hp.up_stages); w.upT_b.resize(hp.up_stages); w.snake_ea.resize(hp.up_stages); w.snake_eb.resize(hp.up_stages);
    w.up_res.resize(hp.up_stages);
    int cin = C;
    for (int s = 0; s < hp.up_stages; s++) {
        const std::string un = "wave_upsampler.up." + std::to_string(s);
        const gguf_tensor * t = g.find(un + ".weight");
        if (!t) return "missing tensor: " + un + ".weight";
        const int cout = (int) t->ne[1];
        hp.up_ch[s] = cout;
        if (cout < 64 || cout % 64 != 0 || cout > 4096) return "unsupported upsampler width " + std::to_string(cout) + " (channel counts must be multiples of 64)";
        if (!L.convT(w.upT[s], un + ".weight", cin, cout, hp.up_kernel[s], hp.up_factor[s])) return L.err;
        w.upT_b[s] = L.vec(un + ".bias", cout);
        std::vector<float> a, b;
        if (!w.upT_b[s] || !L.get("wave_upsampler.snake." + std::to_string(s) + ".alpha", a) ||
            !L.get("wave_upsampler.snake." + std::to_string(s) + ".beta", b)) return L.err;
        if ((int) a.size() != cout || (int) b.size() != cout) return "bad snake parameter size";
        for (auto & v : a) v = expf(v);
        for (auto & v : b) v = expf(v);
        w.snake_ea[s] = L.vec_host(a); w.snake_eb[s] = L.vec_host(b);
        if (!L.resblock(w.up_res[s], "wave_upsampler.resblk." + std::to_string(s) + ".", cout)) return L.err;
        cin = cout;
    }
    {
        const gguf_tensor * t = g.find("wave_upsampler.out_proj.weight");
        if (!t) return "missing tensor: wave_upsampler.out_proj.weight";
        hp.out_proj_dim = (int) t->ne[1];
        if (hp.out_proj_dim < 64 || hp.out_proj_dim % 64 != 0 || hp.out_proj_dim > 4096) return "unsupported out_proj width (must be a multiple of 64)";
        if (!L.linear(w.out_proj, {"wave_upsampler.out_proj.weight"}, cin)) return L.err;
        w.out_proj_b = L.vec("wave_upsampler.out_proj.bias", hp.out_proj_dim);
        std::vector<float> a, b;
        if (!w.out_proj_b || !L.get("wave_upsampler.out_snake.alpha", a) || !L.get("wave_upsampler.out_snake.beta", b)) return L.err;
        for (auto & v : a) v = expf(v);
        for (auto & v : b) v = expf(v);
        w.out_snake_ea = L.vec_host(a); w.out_snake_eb = L.vec_host(b);
    }
    {   // head: interleave (log-mag f, phase f) so the epilogue sees pairs; output column 2f/2f+1 = re/im of bin f
        std::vector<float> hw, hb;
        if (!L.get("istft_head.out.weight", hw) || !L.get("istft_head.out.bias", hb)) return L.err;
        const int P = hp.out_proj_dim, NF = hp.n_freq;
        if (hw.size() != (size_t) 2 * NF * P || hb.size() != (size_t) 2 * NF) return "bad istft_head tensor size";
        std::vector<float> B((size_t) 2 * NF * P), bb((2 * NF + 31) / 32 * 32, 0.f);   // bias padded to whole 32-column chunks
        for (int f = 0; f < NF; f++) {
            memcpy(&B[(size_t) (2 * f) * P], &hw[(size_t) f * P], P * sizeof(float));
            memcpy(&B[(size_t) (2 * f + 1) * P], &hw[(size_t) (NF + f) * P], P * sizeof(float));
            bb[2 * f] = hb[f]; bb[2 * f + 1] = hb[NF + f];
        }
        if (!L.pack(w.head, B, 2 * NF, P, 1, 3, L.store_of("istft_head.out.weight"))) return L.err;
        w.head_b = L.vec_host(bb);
    }
    {   // periodic Hann window (istft_cache, src/istft.cpp:29-31)
        std::vector<float> hn(hp.n_fft);
        for (int i = 0; i < hp.n_fft; i++) hn[i] = 0.5f * (1.0f - cosf(2.0f * (float) M_PI * i / hp.n_fft));
        w.hann = L.vec_host(hn);
        std::vector<float> tw(784);
        istft_twiddles(tw.data());
        w.istft_tw = L.vec_host(tw);
    }
    if (!L.err.empty()) return L.err;
    CK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    own_stream = true;
    if (!(e = ensure_rope(8192)).empty()) return e;
    // > 48 KB dynamic shared memory opt-ins are per device: set for THIS context's device
    CK(gemm_tc3_init_device());
    CK(attention_init());
    CK(istft_init_device());
    { const char * ev = getenv("MIO_ATTN"); attn_simt = ev && !strcmp(ev, "simt"); }
    { const char * ev = getenv("MIO_SERPENTINE"); if (ev && ev[0] == '0') serpentine = false; }
    { const char * ev = getenv("MIO_FUSE_TAIL"); if (ev) fuse_resblock_tail = ev[0] != '0'; }
    { const char * ev = getenv("MIO_EXP_SKIP_NORMS"); exp_skip_norms = ev && ev[0] == '1'; }
    { const char * ev = getenv("MIO_EXP_SLOT_ALIGN"); if (ev && atoi(ev) >= 1 && atoi(ev) <= 64) exp_slot_align = atoi(ev); }
    { const char * ev = getenv("MIO_CHUNK_ACC"); if (ev) chunk_mode = !strcmp(ev, "off") ? 0 : (!strcmp(ev, "all") ? 2 : 1); }
    { const char * ev = getenv("MIO_CHUNK_KB"); if (ev && atoi(ev) >= 1 && atoi(ev) <= 64) chunk_kb = atoi(ev); }
    CK(cudaDeviceSynchronize());
    return {};
}

// RoPE table for positions [0, max_pos): theta_i = pos * s^i built by repeated multiplication in f32, cosf/sinf of libm —
// exactly the reference's cache (ggml_rope_cache_init, ggml-cpu/ops.cpp:5612-5626; theta_scale = powf(base, -2/n_dims)).
// Grown on demand (the reference computes the phases for any position): upload() asks for 2 * (longest utterance).
std::string engine::ensure_rope(int max_pos) {
    if (max_pos <= w.rope_max_pos) return {};
    int n = std::max(8192, w.rope_max_pos);
    while (n < max_pos) n *= 2;
    std::vector<float> r((size_t) n * 64);
    const float theta_scale = powf(hp.rope_theta, -2.0f / 64.0f);
    for (int pos = 0; pos < n; pos++) {
        float theta = (float) pos;
        for (int i = 0; i < 32; i++) {
            r[((size_t) pos * 32 + i) * 2] = cosf(theta);
            r[((size_t) pos * 32 + i) * 2 + 1] = sinf(theta);
            theta *= theta_scale;
        }
    }
    // the same values position-minor for the tcgen05 epilogue, where lane = row = consecutive positions:
    // rope_t[half h][i][pos] = the four floats (cos, sin, cos, sin) of pairs h*16 + 2i, 2i+1 -> a warp's float4 load
    // covers 512 contiguous bytes instead of 32 separate lines
    std::vector<float> rt(r.size());
    for (int h = 0; h < 2; h++)
        for (int i = 0; i < 8; i++)
            for (int pos = 0; pos < n; pos++)
                for (int j = 0; j < 4; j++)
                    rt[((size_t) (h * 8 + i) * n + pos) * 4 + j] = r[((size_t) pos * 32 + h * 16 + 2 * i) * 2 + j];
    CK(cudaSetDevice(device));
    if (stream) CK(cudaStreamSynchronize(stream));     // kernels of the previous batch may still read the old tables
    if (w.rope_pre) cudaFree(w.rope_pre);
    if (w.rope_t) cudaFree(w.rope_t);
    w.rope_pre = w.rope_t = nullptr;
    w.rope_max_pos = 0;
    std::string e = upload_f32(r, &w.rope_pre);
    if (e.empty()) e = upload_f32(rt, &w.rope_t);
    if (!e.empty()) return e;
    w.rope_max_pos = n;
    return {};
}

// A context without decoder weights: istft() of the reference is a free function of (spec, istft_cache)
// (src/istft.h:38-42), so the drop-in needs an ISTFT that works before / without miocodec_load.
std::string engine::init_istft_only(int dev) {
    device = dev;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) return "no CUDA device available (this library has no CPU fallback)";
    if (dev < 0 || dev >= n_dev) return "invalid CUDA device ordinal";
    CK(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, dev));
    if (prop.major != 10) return std::string("device '") + prop.name + "' is not sm_100 (B200); this build contains sm_100a code only";
    istft_only = true;
    std::vector<float> hn(hp.n_fft);   // periodic Hann window (istft_cache, src/istft.cpp:29-31)
    for (int i = 0; i < hp.n_fft; i++) hn[i] = 0.5f * (1.0f - cosf(2.0f * (float) M_PI * i / hp.n_fft));
    std::string e = upload_f32(hn, &w.hann);
    if (!e.empty()) return e;
    w_allocs.push_back(w.hann);
    {
        std::vector<float> tw(784);
        istft_twiddles(tw.data());
        e = upload_f32(tw, &w.istft_tw);
        if (!e.empty()) return e;
        w_allocs.push_back(w.istft_tw);
    }
    CK(istft_init_device());
    CK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    own_stream = true;
    return {};
}

// ---------------------------------------------------------------------------------------------
// workspace
// ---------------------------------------------------------------------------------------------
static std::string make_view(a_view & v, __half * hi, __half * lo, int rows, int K) {
    v.hi = hi; v.lo = lo; v.rows = rows; v.K = K;
    std::string e = make_map(&v.tm_hi, hi, rows, K, 128);
    if (e.empty() && lo) e = make_map(&v.tm_lo, lo, rows, K, 128);
    if (e.empty() && !lo) v.tm_lo = v.tm_hi;
    return e;
}

std::string engine::ensure_workspace() {
    if (ws_ready) return {};
    CK(cudaSetDevice(device));
    // rows per code: 1 (prenet), 2 (decoder), 2*f0 and 2*f0*f1 = frames per code (upsampler stages)
    const size_t R1 = (size_t) max_rows1, R2 = 2 * R1, Ru0 = (size_t) hp.up_res[0] * R1, Ru1 = (size_t) hp.up_res[1] * R1;
    const size_t c0 = (size_t) hp.up_ch[0], c1 = (size_t) hp.up_ch[1], P = (size_t) hp.out_proj_dim;
    const size_t n_ff = std::max({R1 * 3072, Ru0 * c0, Ru1 * c1}), n_c16 = std::max({R2 * 512, Ru0 * c0, Ru1 * c1});
    std::string e;
#define A(ptr, n) if (!(e = dev_alloc(&ptr, (size_t) (n))).empty()) return e;
    if (istft_only) {   // mio_istft_ctx_create: no decoder workspace (run_istft_only grows its own staging buffer)
        A(err_flag, 1);
        CK(cudaMemset(err_flag, 0, sizeof(int)));
        ws_ready = true;
        return {};
    }
    A(x768, R1 * 768); A(qkv_hi, std::max(R1 * 2304, R2 * 1536)); A(qkv_lo, std::max(R1 * 2304, R2 * 1536));
    A(x512, R2 * 512); A(y512, R2 * 512); A(xu0, Ru0 * c0); A(yu0, Ru0 * c0); A(xu1, Ru1 * c1); A(yu1, Ru1 * c1);
    A(spec, Ru1 * SPEC_LD);
    A(aA_hi, R1 * 1024); A(aA_lo, R1 * 1024); A(aFF_hi, n_ff); A(aFF_lo, n_ff);
    A(c16, n_c16); A(hH_hi, Ru1 * P); A(hH_lo, Ru1 * P);
    A(gn_part, (size_t) (max_rows1 / GN_SEG_CODES + max_utts + 1) * 32);
    A(gn_mr, (size_t) max_utts * 32);
    A(attn_b1, R1); A(attn_b2, R2);
    A(mod, (size_t) max_voices * w.cond_rows);
    A(voice_silu, (size_t) max_voices * hp.adaln_dim);
    A(peak, (size_t) max_utts);
    A(err_flag, 1);
#undef A
    CK(cudaMemset(err_flag, 0, sizeof(int)));
    // stale rows beyond the current batch are only ever multiplied into masked outputs, but start finite
    CK(cudaMemset(aA_hi, 0, R1 * 1024 * 2)); CK(cudaMemset(aA_lo, 0, R1 * 1024 * 2));
    CK(cudaMemset(aFF_hi, 0, n_ff * 2)); CK(cudaMemset(aFF_lo, 0, n_ff * 2));
    CK(cudaMemset(c16, 0, n_c16 * 2));
    CK(cudaMemset(hH_hi, 0, Ru1 * P * 2)); CK(cudaMemset(hH_lo, 0, Ru1 * P * 2));
    if (!(e = make_view(vA768, aA_hi, aA_lo, (int) R1, 768)).empty()) return e;
    if (!(e = make_view(vA512, aA_hi, aA_lo, (int) R2, 512)).empty()) return e;
    if (!(e = make_view(vFF2048, aFF_hi, aFF_lo, (int) R1, 2048)).empty()) return e;
    if (!(e = make_view(vFF1536, aFF_hi, aFF_lo, (int) R2, 1536)).empty()) return e;
    if (!(e = make_view(vFFu0, aFF_hi, aFF_lo, (int) Ru0, (int) c0)).empty()) return e;
    if (!(e = make_view(vFFu1, aFF_hi, aFF_lo, (int) Ru1, (int) c1)).empty()) return e;
    if (!(e = make_view(vC512, c16, nullptr, (int) R2, 512)).empty()) return e;
    if (!(e = make_view(vCu0, c16, nullptr, (int) Ru0, (int) c0)).empty()) return e;
    if (!(e = make_view(vCu1, c16, nullptr, (int) Ru1, (int) c1)).empty()) return e;
    if (!(e = make_view(vH, hH_hi, hH_lo, (int) Ru1, (int) P)).empty()) return e;
    ws_ready = true;
    return {};
}

row_map engine::rmap() const {
    row_map r;
    r.row_utt = d_row_utt + cur_row0; r.start = d_start + cur_utt0; r.len = d_len + cur_utt0; r.voice = d_voice + cur_utt0;
    r.rows1 = cur_rows1; r.n_utts = cur_n_utts;
    return r;
}

// ---------------------------------------------------------------------------------------------
// batch upload: builds the padded row space, split into chunks of whole utterances that fit the workspace
// ---------------------------------------------------------------------------------------------
std::string engine::upload(const int32_t * codes, const int32_t * cu, int n_utts, const float * voice_table, int nv,
                           const int32_t * voice_ids, int chunk_rows_limit) {
    batch_uploaded = false; batch_ran = false;
    // chunk_rows_limit > 0: cut chunks at that many rows (pipelined host delivery), never above the workspace
    const int row_cap = (chunk_rows_limit > 0 && chunk_rows_limit < max_rows1) ? chunk_rows_limit : max_rows1;
    if (istft_only) return "bad argument: this context was created by mio_istft_ctx_create and has no decoder weights";
    if (n_utts < 0 || (n_utts > 0 && (!codes || !cu || !voice_table || !voice_ids)) || nv <= 0) return "bad argument";
    if (nv > max_voices) return "too many voices (max " + std::to_string(max_voices) + ")";
    std::string e = ensure_workspace();
    if (!e.empty()) return e;
    CK(cudaSetDevice(device));
    lay = batch_layout();
    lay.n_utts = n_utts;
    chunks.clear();
    const int fpc = hp.frames_per_code;
    int64_t rows_total = 0;
    std::vector<int32_t> chunk_cu_all;   // per chunk: local prefix of istft chunk counts (n+1 entries each)
    std::vector<int32_t> seg_cu_all;     // per chunk: local prefix of GroupNorm segment counts (n+1 entries each)
    // Every utterance starts on a multiple of ROW_ALIGN rows (code resolution) of its chunk and is followed by at least
    // one zero row: the 16-row groups of the attention kernel and the GroupNorm segments then sit at utterance-relative
    // positions, which makes an utterance's result bit-identical wherever it lies in a batch (the reference is
    // batch-invariant by construction: it decodes one utterance per call).
    // (exp_slot_align: measurement-only override, MIO_EXP_SLOT_ALIGN -- anything but ROW_ALIGN gives up batch invariance)
    const int slot_align = exp_slot_align > 0 ? exp_slot_align : ROW_ALIGN;
    auto slot_rows = [slot_align](int T) { return (T + 1 + slot_align - 1) / slot_align * slot_align; };
    // Pipelined delivery: chunk i's D2H copy runs while chunk i+1 computes, so only the LAST copy is exposed.  Every
    // extra chunk costs ~1 ms of launch tails (160 kernels), so the batch is cut as rarely as the workspace allows and
    // the final chunk is split once more into (100 - p) % and p %: with p = fetch_tail_percent ~ D2H time / (compute +
    // D2H time) the small tail computes exactly while the big piece is copied, and only the tail's own copy is exposed.
    int64_t rows_left = 0;
    for (int i = 0; i < n_utts; i++) rows_left += (int64_t) slot_rows(std::max(0, cu[i + 1] - cu[i]));
    bool tail_split_done = false;
    big_piece_chunk = -1;
    int t_max = 0;
    int u = 0;
    while (u < n_utts) {
        chunk_info c;
        c.utt0 = u; c.row0 = (int) rows_total; c.cu0 = (int) chunk_cu_all.size(); c.scu0 = (int) seg_cu_all.size();
        int rows = 0, blocks = 0, segs = 0;
        chunk_cu_all.push_back(0);
        seg_cu_all.push_back(0);
        int row_cap_c = row_cap;
        const int tail_pct = fetch_tail_percent >= 0 ? fetch_tail_percent : (int) (tail_auto_pct + 0.5f);
        if (chunk_rows_limit > 0 && tail_pct > 0 && !tail_split_done && rows_left <= row_cap && rows_left >= 4096) {
            row_cap_c = (int) (rows_left - rows_left * tail_pct / 100);
            tail_split_done = true;
            big_piece_chunk = (int) chunks.size();
        }
        while (u < n_utts) {
            const int T = cu[u + 1] - cu[u];
            if (T < 0) return "cu_seqlens must be non-decreasing";
            const int slot = slot_rows(T);
            if (slot > max_rows1) return "utterance of " + std::to_string(T) + " codes exceeds max_chunk_codes";
            t_max = std::max(t_max, T);
            if (voice_ids[u] < 0 || voice_ids[u] >= nv) return "voice id out of range";
            if ((rows + slot > row_cap_c && rows > 0) || rows + slot > max_rows1 || (u - c.utt0) >= max_utts) break;
            lay.start.push_back(rows); lay.len.push_back(T); lay.voice.push_back(voice_ids[u]);
            lay.out_off.push_back((int64_t) cu[u] * hp.samples_per_token);
            for (int i = 0; i < T; i++) {
                const int32_t code = codes[cu[u] + i];
                if (code < 0 || code >= hp.vocab)   // the reference aborts in GGML_ASSERT (ggml-cpu/ops.cpp:4811)
                    return "RANGE: code " + std::to_string(code) + " at position " + std::to_string(i) + " of utterance " +
                           std::to_string(u) + " is outside [0, " + std::to_string(hp.vocab) + ")";
                lay.codes_padded.push_back(code); lay.row_utt.push_back(u - c.utt0);
            }
            for (int i = T; i < slot; i++) { lay.codes_padded.push_back(-1); lay.row_utt.push_back(-1); }
            rows += slot;
            blocks += (T * fpc + ISTFT_SEG_PER_BLOCK - 1) / ISTFT_SEG_PER_BLOCK;
            chunk_cu_all.push_back(blocks);
            segs += (T + GN_SEG_CODES - 1) / GN_SEG_CODES;
            seg_cu_all.push_back(segs);
            u++;
        }
        c.n_utts = u - c.utt0; c.rows1 = rows; c.istft_blocks = blocks; c.n_segs = segs;
        rows_total += rows; rows_left -= rows;
        chunks.push_back(c);
    }
    if (!(e = ensure_rope(2 * t_max)).empty()) return e;    // decoder positions run to 2 T - 1
    lay.rows1 = (int) rows_total;
    lay.total_codes = n_utts ? cu[n_utts] : 0;
    wave_samples = lay.total_codes * hp.samples_per_token;

    // device arrays (grown on demand)
    auto grow = [&](auto ** p, size_t & cap, size_t need) -> std::string {
        if (need <= cap && *p) return {};
        if (*p) cudaFree(*p);
        *p = nullptr;
        cap = need + need / 4 + 64;
        return dev_alloc(p, cap);
    };
    if (!(e = grow(&d_codes, cap_rows, (size_t) rows_total)).empty()) return e;
    if (!(e = grow(&d_row_utt, cap_rows_ru, (size_t) rows_total)).empty()) return e;
    if ((size_t) n_utts + 1 > cap_utts) {
        for (void * p : {(void *) d_start, (void *) d_len, (void *) d_voice, (void *) d_out_off}) if (p) cudaFree(p);
        cap_utts = (size_t) n_utts + n_utts / 4 + 64;
        if (!(e = dev_alloc(&d_start, cap_utts)).empty() || !(e = dev_alloc(&d_len, cap_utts)).empty() ||
            !(e = dev_alloc(&d_voice, cap_utts)).empty() || !(e = dev_alloc(&d_out_off, cap_utts)).empty()) return e;
    }
    if (!(e = grow(&d_chunk_cu, cap_ccu, chunk_cu_all.size())).empty()) return e;
    if (!(e = grow(&d_seg_cu, cap_scu, seg_cu_all.size())).empty()) return e;
    if (!(e = grow(&d_voice_table, cap_vt, (size_t) nv * hp.adaln_dim)).empty()) return e;
    if ((int64_t) wave_cap < wave_samples) {
        // first allocation: exact.  A context whose batches keep growing (a streaming session re-decodes ever longer
        // prefixes: 18 sizes for a 700-code stream, and freeing + allocating gigabytes stalls the device for tens to
        // hundreds of ms each time, MIO_STREAM_TIMING=1) grows geometrically instead.
        const int64_t grown = d_wave ? wave_cap + wave_cap / 2 : 0;
        if (d_wave) cudaFree(d_wave);
        d_wave = nullptr;
        wave_cap = std::max<int64_t>(wave_samples + 64, grown);
        if (!(e = dev_alloc(&d_wave, (size_t) wave_cap)).empty()) {
            (void) cudaGetLastError();                 // the failed attempt must not surface at the next launch check
            wave_cap = wave_samples + 64;              // the headroom did not fit: exact size
            if (!(e = dev_alloc(&d_wave, (size_t) wave_cap)).empty()) { wave_cap = 0; return e; }
        }
    }
    // one pinned staging buffer, async copies on the context stream
    const size_t b_codes = (size_t) rows_total * 4, b_utts = (size_t) n_utts * 4, b_off = (size_t) n_utts * 8,
                 b_ccu = chunk_cu_all.size() * 4, b_vt = (size_t) nv * hp.adaln_dim * 4;
    const size_t need = 2 * b_codes + 3 * b_utts + b_off + 2 * b_ccu + b_vt + 256;
    if (need > h_stage_bytes) {
        if (h_stage) cudaFreeHost(h_stage);
        h_stage = nullptr;
        h_stage_bytes = need + need / 4;
        CK(cudaMallocHost(&h_stage, h_stage_bytes));
    }
    CK(cudaStreamSynchronize(stream));   // previous async copies out of the staging buffer are done
    uint8_t * hs = (uint8_t *) h_stage;
    size_t off = 0;
    auto put = [&](void * dst, const void * src, size_t bytes) -> cudaError_t {
        if (bytes == 0) return cudaSuccess;
        memcpy(hs + off, src, bytes);
        cudaError_t ce = cudaMemcpyAsync(dst, hs + off, bytes, cudaMemcpyHostToDevice, stream);
        off += (bytes + 15) / 16 * 16;
        return ce;
    };
    CK(put(d_codes, lay.codes_padded.data(), b_codes));
    CK(put(d_row_utt, lay.row_utt.data(), b_codes));
    CK(put(d_start, lay.start.data(), b_utts));
    CK(put(d_len, lay.len.data(), b_utts));
    CK(put(d_voice, lay.voice.data(), b_utts));
    CK(put(d_out_off, lay.out_off.data(), b_off));
    CK(put(d_chunk_cu, chunk_cu_all.data(), b_ccu));
    CK(put(d_seg_cu, seg_cu_all.data(), b_ccu));
    CK(put(d_voice_table, voice_table, b_vt));
    n_voices = nv;
    h2d_bytes = 2 * b_codes + 3 * b_utts + b_off + 2 * b_ccu + b_vt;
    batch_uploaded = true;
    return {};
}

// ---------------------------------------------------------------------------------------------
// debug snapshots
// ---------------------------------------------------------------------------------------------
// A tap is a named point of the kernel sequence.  With option "debug_snapshots" the buffer is copied to the host
// (mio_debug_read); a pending mio_debug_write of the same name REPLACES the buffer's leading rows first, so the
// kernels behind the tap run on the injected values (teacher-forced per-stage parity tests: every stage is fed the
// oracle's input for that stage and compared with the oracle's output at fp32 accuracy).
void engine::tap_f32(const std::string & name, float * d, size_t rows, size_t cols, size_t ld) {
    if (ld == 0) ld = cols;
    auto ov = overrides.find(name);
    if (ov != overrides.end() && !ov->second.empty()) {
        const size_t r = std::min(rows, ov->second.size() / cols);
        cudaMemcpy2DAsync(d, ld * sizeof(float), ov->second.data(), cols * sizeof(float), cols * sizeof(float), r, cudaMemcpyHostToDevice, stream);
        cudaStreamSynchronize(stream);
    }
    if (!snap_wanted(name)) return;
    cudaStreamSynchronize(stream);
    std::vector<float> & v = snaps[name];
    v.resize(rows * cols);
    cudaMemcpy2D(v.data(), cols * sizeof(float), d, ld * sizeof(float), cols * sizeof(float), rows, cudaMemcpyDeviceToHost);
}
void engine::tap_f16(const std::string & name, __half * hi, __half * lo, size_t n) {
    auto ov = overrides.find(name);
    if (ov != overrides.end() && !ov->second.empty()) {
        const size_t m = std::min(n, ov->second.size());
        std::vector<__half> h(m), l(m);
        for (size_t i = 0; i < m; i++) {
            const float x = ov->second[i];
            h[i] = __float2half_rn(x);
            l[i] = __float2half_rn(x - __half2float(h[i]));
        }
        cudaMemcpyAsync(hi, h.data(), m * 2, cudaMemcpyHostToDevice, stream);
        if (lo) cudaMemcpyAsync(lo, l.data(), m * 2, cudaMemcpyHostToDevice, stream);
        cudaStreamSynchronize(stream);
    }
    if (!snap_wanted(name)) return;
    cudaStreamSynchronize(stream);
    std::vector<uint16_t> a(n), b(n);
    cudaMemcpy(a.data(), hi, n * 2, cudaMemcpyDeviceToHost);
    if (lo) cudaMemcpy(b.data(), lo, n * 2, cudaMemcpyDeviceToHost);
    std::vector<float> & v = snaps[name];
    v.resize(n);
    for (size_t i = 0; i < n; i++) v[i] = fp16_bits_to_float(a[i]) + (lo ? fp16_bits_to_float(b[i]) : 0.f);
}

// ---------------------------------------------------------------------------------------------
// GEMM dispatch
// ---------------------------------------------------------------------------------------------
cudaError_t engine::gemm(const gemm_weight & W, const a_view & A, int M, gemm_params p) {
    p.M = M; p.N = W.N; p.K = W.K; p.taps = W.taps; p.npass = W.npass; p.out_scale = W.out_scale;
    p.reverse = next_dir() ? 1 : 0;
    // chunked accumulation (gemm3.cu): callers request it (p.chunk = 1) for the layers in front of the first conv1d
    p.chunk = (W.npass == 3 && (chunk_mode == 2 || (chunk_mode == 1 && p.chunk > 0))) ? chunk_kb : 0;
    { static const int order = getenv("MIO_MMA_ORDER") ? atoi(getenv("MIO_MMA_ORDER")) : 2; p.mma_order = order; }
    if (p.bias_mod <= 1) p.bias_mod = W.N;
    p.a_hi = A.hi; p.a_lo = A.lo; p.lda = A.K; p.a_rows = A.rows;
    p.b_hi = W.hi; p.b_lo = W.lo;
    if (A.K != W.K) return cudaErrorInvalidValue;
    if (W.store == 2) {   // BF16-stored matrix: the activation pair becomes one BF16 operand, in place (every operand buffer feeds one product)
        if (W.taps != 1) return cudaErrorInvalidValue;
        cudaError_t e = launch_pair_to_bf16(A.hi, A.lo, (size_t) M * A.K, stream);
        if (e != cudaSuccess) return e;
        launches++;
        p.bf16 = 1;
    }
    launches++;
    // algorithmic FLOPs: M counts the slot rows (16-row alignment + the zero row behind every utterance); pad rows are not work
    prof_begin(W.npass == 3 ? CAT_GEMM_SPLIT : CAT_GEMM_F16, 2.0 * M * cur_live_frac * (double) W.N * W.K * W.taps, 0.0);
    cudaError_t ce;
    if (use_simt) ce = launch_gemm_simt(p, stream);
    else {
        const bool f16 = (p.epi == EPI_SWIGLU || p.epi == EPI_SPLIT || p.epi == EPI_QKV);
        const int ncols = (p.epi == EPI_SWIGLU) ? W.N / 2 : W.N;
        CUtensorMap o0, o1;
        bool ok;
        if (f16) {
            ok = out_map(&o0, p.out_hi, 2, M, ncols, (size_t) p.ld_h * 2) && out_map(&o1, p.out_lo, 2, M, ncols, (size_t) p.ld_h * 2);
        } else {
            ok = out_map(&o0, p.out_f32, 4, M, ncols, (size_t) p.ld_out * 4);
            o1 = o0;
        }
        if (!ok) { prof_end(); return cudaErrorInvalidValue; }
        // tile width: of {bn2, 128, 64} the one with the least (waves of tiles over the CTA pairs) x (tile width), with
        // an 8 % handicap per halving (narrower tiles re-read A and pay the epilogue more often).  Large M keeps bn2;
        // small M (a short utterance, the tail piece of the delivery pipeline) gets tiles that fill the machine.
        // SwiGLU pairs need >= 64 accumulator columns per epilogue warp, i.e. BN >= 128.
        const int tiles_m = (M + 255) / 256, pairs = std::max(1, num_sms / 2);
        const int bn_min = (p.epi == EPI_SWIGLU) ? 128 : 64;
        int bn = W.bn2;
        double best = 1e30;
        for (int cand = W.bn2; cand >= bn_min; cand >>= 1) {
            const int tiles = tiles_m * ((W.N + cand - 1) / cand);
            const double waves = (double) ((tiles + pairs - 1) / pairs);
            double cost = waves * cand;
            for (int c = cand; c < W.bn2; c <<= 1) cost *= 1.08;
            if (cost < best) { best = cost; bn = cand; }
        }
        const CUtensorMap & bh = bn == W.bn2 ? W.tm2_hi : (bn == 128 ? W.tm128_hi : W.tm64_hi);
        const CUtensorMap & bl = bn == W.bn2 ? W.tm2_lo : (bn == 128 ? W.tm128_lo : W.tm64_lo);
        ce = launch_gemm_tc3(p, A.tm_hi, A.tm_lo, bh, bl, o0, o1, bn, num_sms, stream);
    }
    prof_end();
    return ce;
}

// One bare GEMM through the product kernel (mio_debug_gemm): out[M][N] = A[M][K] * B[N][K]^T with the operands split /
// rounded exactly as the decoder does (npass 3: fp16 hi/lo pairs, npass 1: fp16).  Per-op parity tests and the
// accumulation study (tests/accuracy_study.py) compare it with an fp64 product of the same operands.
std::string engine::debug_gemm(int M, int N, int K, const float * A, const float * B, int npass, float * out) {
    if (M <= 0 || N <= 0 || K <= 0 || K % 64 != 0 || !A || !B || !out || (npass != 1 && npass != 3)) return "bad argument";
    if (istft_only) return "bad argument: this context has no decoder";
    CK(cudaSetDevice(device));
    cur_live_frac = 1.0;             // a bare product has no pad rows
    gguf_file none;
    loader L(*this, none);
    gemm_weight W;
    const size_t n_before = w_allocs.size();
    std::vector<float> Bv(B, B + (size_t) N * K);
    if (!L.pack(W, Bv, N, K, 1, npass)) return L.err;
    float *dA = nullptr, *dO = nullptr;
    __half *ah = nullptr, *al = nullptr;
    std::string e;
    auto cleanup = [&]() {
        cudaStreamSynchronize(stream);
        for (void * p : {(void *) dA, (void *) dO, (void *) ah, (void *) al}) if (p) cudaFree(p);
        while (w_allocs.size() > n_before) { cudaFree(w_allocs.back()); w_allocs.pop_back(); }
        out_maps.clear();                                // the freed buffers' addresses may be reused
    };
    const size_t na = (size_t) M * K;
    const int ldo = (N + 3) / 4 * 4;                     // TMA store pitch: a multiple of 16 bytes
    if (!(e = dev_alloc(&dA, na)).empty() || !(e = dev_alloc(&dO, (size_t) M * ldo)).empty() || !(e = dev_alloc(&ah, na)).empty() ||
        !(e = dev_alloc(&al, na)).empty()) { cleanup(); return e; }
    cudaMemcpy(dA, A, na * sizeof(float), cudaMemcpyHostToDevice);
    split_weights_kernel<<<296, 256>>>(dA, na, 1.0f, ah, npass == 3 ? al : nullptr);
    if (cudaDeviceSynchronize() != cudaSuccess) { cleanup(); return "split kernel failed"; }
    a_view v;
    if (!(e = make_view(v, ah, npass == 3 ? al : nullptr, M, K)).empty()) { cleanup(); return e; }
    gemm_params p;
    p.epi = EPI_STORE; p.out_f32 = dO; p.ld_out = ldo; p.chunk = 1;     // chunked unless option chunk_acc = off
    cudaError_t ce = gemm(W, v, M, p);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(stream);
    if (ce == cudaSuccess) ce = cudaMemcpy2D(out, (size_t) N * sizeof(float), dO, (size_t) ldo * sizeof(float), (size_t) N * sizeof(float), M, cudaMemcpyDeviceToHost);
    cleanup();
    if (ce != cudaSuccess) return std::string("debug_gemm: ") + cudaGetErrorString(ce);
    return {};
}

// ---------------------------------------------------------------------------------------------
// per-category device timing (CUDA events on the launching stream), used by bench.py for the roofline lines
// ---------------------------------------------------------------------------------------------
void engine::prof_begin(int cat, double flops, double bytes) {
    if (!profiling) return;
    if (prof_used + 2 > prof_events.size()) {
        const size_t old = prof_events.size();
        prof_events.resize(old + 256);
        for (size_t i = old; i < prof_events.size(); i++) cudaEventCreate(&prof_events[i]);
    }
    prof_rec r; r.cat = cat; r.flops = flops; r.bytes = bytes; r.ev = prof_used;
    prof_recs.push_back(r);
    cudaEventRecord(prof_events[prof_used], stream);
    prof_used += 2;
}
void engine::prof_end() {
    if (!profiling) return;
    cudaEventRecord(prof_events[prof_recs.back().ev + 1], stream);
}
void engine::prof_collect() {
    // call after the stream has been synchronised
    for (const prof_rec & r : prof_recs) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, prof_events[r.ev], prof_events[r.ev + 1]) != cudaSuccess) continue;
        prof_tot[r.cat].ms += ms; prof_tot[r.cat].flops += r.flops; prof_tot[r.cat].bytes += r.bytes; prof_tot[r.cat].launches++;
    }
    prof_recs.clear();
    prof_used = 0;
}

// ---------------------------------------------------------------------------------------------
// the decoder forward of one chunk + ISTFT (stage numbers follow src/miocodec.cpp:598-740)
// ---------------------------------------------------------------------------------------------
#define CKR(cat, bytes, expr)                                                                          \
    do {                                                                                               \
        prof_begin(cat, 0.0, bytes);                                                                   \
        cudaError_t _e = (expr);                                                                       \
        prof_end();                                                                                    \
        launches++;                                                                                    \
        if (_e != cudaSuccess) return std::string(#expr) + ": " + cudaGetErrorString(_e);              \
    } while (0)
// measurement-only switch (MIO_EXP_SKIP_NORMS=1, profiles/r3_norm_upper_bound.md): after the first two runs of a context
// have filled the operand buffers with realistic values, every LayerNorm / GroupNorm launch that feeds a GEMM operand is skipped (65 of 66; the final
// AdaLN, which rewrites the fp32 stream in place, stays so that the head sees bounded values).  Results are WRONG;
// the step time is the floor any fusion of the norms into neighbouring kernels could reach.
#define CKN(cat, bytes, expr)                                                                          \
    do {                                                                                               \
        if (!(exp_skip_norms && exp_runs > 2)) CKR(cat, bytes, expr);                                  \
    } while (0)
#define CKG(expr)                                                                                      \
    do {                                                                                               \
        cudaError_t _e = (expr);                                                                       \
        if (_e != cudaSuccess) return std::string(#expr) + ": " + cudaGetErrorString(_e);              \
    } while (0)

std::string engine::run_resblock(const resblock_w & R, float * x, float * y, const a_view & vc, int rows, int row_div,
                                 const std::string & tag, const a_view * split_out) {
    const row_map rm = rmap();
    const int C = R.C;
    const int32_t * scu = d_seg_cu + cur_scu0;
    gemm_params p;
    // GN -> SiLU -> conv1 + bias
    CKN(CAT_GN_STATS, (double) (rows) * 4.0 * (C), launch_gn_stats(x, C, rm, row_div, scu, cur_n_segs, gn_part, stream, next_dir()));
    CKN(CAT_GN_APPLY, (double) (rows) * 6.0 * (C), launch_gn_apply_silu(x, rows, C, rm, row_div, gn_part, scu, hp.gn_eps, R.n1w, R.n1b, gn_mr, vc.hi, stream, next_dir())); launches++;
    if (taps_on()) tap_f16(tag + "_h1", vc.hi, nullptr, (size_t) rows * C);
    p = gemm_params(); p.epi = EPI_STORE; p.out_f32 = y; p.ld_out = C; p.bias = R.c1b;
    CKG(gemm(R.c1, vc, rows, p));
    if (taps_on()) tap_f32(tag + "_c1", y, rows, C);
    // GN -> SiLU -> conv2 + bias + residual
    CKN(CAT_GN_STATS, (double) (rows) * 4.0 * (C), launch_gn_stats(y, C, rm, row_div, scu, cur_n_segs, gn_part, stream, next_dir()));
    CKN(CAT_GN_APPLY, (double) (rows) * 6.0 * (C), launch_gn_apply_silu(y, rows, C, rm, row_div, gn_part, scu, hp.gn_eps, R.n2w, R.n2b, gn_mr, vc.hi, stream, next_dir())); launches++;
    if (taps_on()) tap_f16(tag + "_h2", vc.hi, nullptr, (size_t) rows * C);
    if (split_out) {   // x + conv2(...) goes straight to the next product's operand buffers, zeros on pad rows
        p = gemm_params(); p.epi = EPI_SPLIT; p.out_hi = split_out->hi; p.out_lo = split_out->lo; p.ld_h = C; p.bias = R.c2b;
        p.resid = x; p.ld_resid = C; p.row_utt = rm.row_utt; p.row_div = row_div;
        CKG(gemm(R.c2, vc, rows, p));
        return {};
    }
    p = gemm_params(); p.epi = EPI_RESID; p.out_f32 = x; p.ld_out = C; p.bias = R.c2b;
    CKG(gemm(R.c2, vc, rows, p));
    if (taps_on()) tap_f32(tag + "_out", x, rows, C);
    return {};
}

// QKV projection epilogue: RoPE on q|k with utterance-relative positions, q / sqrt(head_dim), fp16 hi/lo split
gemm_params engine::qkv_params(const row_map & rm, int D, int row_div) {
    gemm_params p;
    p.epi = EPI_QKV; p.out_hi = qkv_hi; p.out_lo = qkv_lo; p.ld_h = 3 * D;
    p.row_utt = rm.row_utt; p.row_div = row_div; p.utt_start = rm.start;
    p.rope = w.rope_pre; p.rope_t = w.rope_t; p.rope_max_pos = w.rope_max_pos; p.rope_cols = 2 * D; p.q_cols = D; p.q_scale = 0.125f;   // 1/sqrt(64)
    return p;
}

// banded attention over the q|k|v written by the QKV epilogue; bytes: 12*D read (hi+lo of q,k,v) + 4*D written per row
std::string engine::run_attention(int rows, int D, int n_head, int window, int row_div, const a_view & out, const int2 * bounds) {
    const row_map rm = rmap();
    if (attn_simt) {
        CKR(CAT_ATTN, (double) rows * 16.0 * D, launch_attention_simt(qkv_hi, qkv_lo, rows, D, n_head, window, rm, row_div, out.hi, out.lo, stream));
        return {};
    }
    CUtensorMap ih, il, oh, ol;
    if (!out_map(&ih, qkv_hi, 2, rows, 3 * D, (size_t) 3 * D * 2, 64, 64) || !out_map(&il, qkv_lo, 2, rows, 3 * D, (size_t) 3 * D * 2, 64, 64) ||
        !out_map(&oh, out.hi, 2, rows, D, (size_t) D * 2, 64, 16) || !out_map(&ol, out.lo, 2, rows, D, (size_t) D * 2, 64, 16)) return err;
    CKR(CAT_ATTN, (double) rows * 16.0 * D, launch_attention_tc(ih, il, oh, ol, rows, D, n_head, window, bounds, stream, next_dir()));
    return {};
}

std::string engine::run_chunk(const chunk_info & c) {
    cur_row0 = c.row0; cur_utt0 = c.utt0; cur_rows1 = c.rows1; cur_n_utts = c.n_utts; cur_scu0 = c.scu0; cur_n_segs = c.n_segs;
    {
        int64_t live = 0;
        for (int u = 0; u < c.n_utts; u++) live += lay.len[c.utt0 + u];
        cur_live_frac = c.rows1 > 0 ? (double) live / (double) c.rows1 : 1.0;
    }
    const row_map rm = rmap();
    const int R1 = c.rows1, R2 = 2 * R1, rd0 = hp.up_res[0], rd1 = hp.up_res[1], Ru0 = rd0 * R1, Ru1 = rd1 * R1;
    const int c0 = hp.up_ch[0], c1 = hp.up_ch[1], P = hp.out_proj_dim, f0 = hp.up_factor[0], f1 = hp.up_factor[1];
    const int D = hp.prenet_dim, C = hp.dec_dim;
    gemm_params p;
    std::string e;

    // valid key ranges of the banded attention, once per chunk and resolution (all layers reuse them)
    if (!attn_simt) {
        CKR(CAT_ATTN, 0.0, launch_attn_bounds(rm, R1, 1, hp.prenet_window, attn_b1, stream));
        CKR(CAT_ATTN, 0.0, launch_attn_bounds(rm, R2, 2, hp.dec_window, attn_b2, stream));
    }
    // 1. token gather
    CKR(CAT_GATHER, (double) (R1) * (4.0 + 8.0 * (D)), launch_gather(w.token_embd, d_codes + c.row0, x768, R1, D, hp.vocab, err_flag, stream));
    if (taps_on()) tap_f32("embed", x768, R1, D);
    // 2. prenet
    for (int i = 0; i < hp.prenet_layers; i++) {
        const prenet_layer_w & l = w.prenet[i];
        CKN(CAT_LN, (double) (R1) * 8.0 * (D), launch_layernorm(x768, R1, D, hp.norm_eps, l.an_w, l.an_b, nullptr, 0, 0, 0, rm, 1, vA768.hi, vA768.lo, nullptr, stream, next_dir()));
        p = qkv_params(rm, D, 1);
        p.chunk = 1;
        CKG(gemm(l.qkv, vA768, R1, p));
        if (i == 0 && taps_on()) tap_f16("pre0_qkv_rope", qkv_hi, qkv_lo, (size_t) R1 * 3 * D);
        if (!(e = run_attention(R1, D, hp.prenet_heads, hp.prenet_window, 1, vA768, attn_b1)).empty()) return e;
        if (i == 0 && taps_on()) tap_f16("pre0_attn", vA768.hi, vA768.lo, (size_t) R1 * D);
        p = gemm_params(); p.epi = EPI_RESID; p.out_f32 = x768; p.ld_out = D;
        p.chunk = 1;
        CKG(gemm(l.wo, vA768, R1, p));
        if (i == 0 && taps_on()) tap_f32("pre0_x_attn", x768, R1, D);
        CKN(CAT_LN, (double) (R1) * 8.0 * (D), launch_layernorm(x768, R1, D, hp.norm_eps, l.fn_w, l.fn_b, nullptr, 0, 0, 0, rm, 1, vA768.hi, vA768.lo, nullptr, stream, next_dir()));
        p = gemm_params(); p.epi = EPI_SWIGLU; p.out_hi = vFF2048.hi; p.out_lo = vFF2048.lo; p.ld_h = hp.prenet_ff;
        p.chunk = 1;
        CKG(gemm(l.gu, vA768, R1, p));
        p = gemm_params(); p.epi = EPI_RESID; p.out_f32 = x768; p.ld_out = D;
        p.chunk = 1;
        CKG(gemm(l.down, vFF2048, R1, p));
        if (taps_on()) tap_f32("pre" + std::to_string(i) + "_out", x768, R1, D);
    }
    if (taps_on()) tap_f32("prenet_x", x768, R1, D);
    CKN(CAT_LN, (double) (R1) * 8.0 * (D), launch_layernorm(x768, R1, D, hp.norm_eps, w.pre_norm_w, w.pre_norm_b, nullptr, 0, 0, 0, rm, 1, vA768.hi, vA768.lo, nullptr, stream, next_dir()));
    // 2b. 768 -> 512 (+bias) written straight as the split operand of the ConvTranspose GEMM
    p = gemm_params(); p.epi = EPI_SPLIT; p.out_hi = vA512.hi; p.out_lo = vA512.lo; p.ld_h = C; p.bias = w.pre_out_b;
    // the prenet LN output lives in aA viewed as [R1][768]; the 512-wide result must not overwrite it while other
    // CTAs still read it -> stage through aFF
    p.out_hi = aFF_hi; p.out_lo = aFF_lo;
    p.chunk = 1;
    CKG(gemm(w.pre_out, vA768, R1, p));
    if (taps_on()) tap_f16("prenet_out", aFF_hi, aFF_lo, (size_t) R1 * C);
    // 3. wave_upsample ConvTranspose1d(k=2, s=2): [R1][2*512] == [R2][512]
    {
        a_view v = vFF1536;   // aFF viewed as [.][512]
        if (!(e = view512_ff(v)).empty()) return e;
        p = gemm_params(); p.epi = EPI_STORE; p.out_f32 = x512; p.ld_out = 2 * C; p.bias = w.up0_b; p.bias_mod = C;
        p.chunk = 1;
        CKG(gemm(w.up0, v, R1, p));
    }
    if (taps_on()) tap_f32("upsample", x512, R2, C);
    // 4. prior resnet blocks
    for (int b = 0; b < hp.resnet_blocks; b++)
        if (!(e = run_resblock(w.prior[b], x512, y512, vC512, R2, 2, "prior" + std::to_string(b))).empty()) return e;
    if (taps_on()) tap_f32("prior", x512, R2, C);
    // 5. decoder: 8 AdaLN-Zero layers (conditioning vectors per voice were computed once in run())
    for (int i = 0; i < hp.dec_layers; i++) {
        const dec_layer_w & l = w.dec[i];
        const int oa = (i * 2 + 0) * 3 * C, of = (i * 2 + 1) * 3 * C;     // shift | scale | gate
        CKN(CAT_LN, (double) (R2) * 8.0 * (C), launch_layernorm(x512, R2, C, hp.norm_eps, nullptr, nullptr, mod, w.cond_rows, oa, oa + C, rm, 2, vA512.hi, vA512.lo, nullptr, stream, next_dir()));
        p = qkv_params(rm, C, 2);
        CKG(gemm(l.qkv, vA512, R2, p));
        if (i == 0 && taps_on()) tap_f16("dec0_qkv_rope", qkv_hi, qkv_lo, (size_t) R2 * 3 * C);
        if (!(e = run_attention(R2, C, hp.dec_heads, hp.dec_window, 2, vA512, attn_b2)).empty()) return e;
        p = gemm_params(); p.epi = EPI_RESID; p.out_f32 = x512; p.ld_out = C;
        p.gate = mod + oa + 2 * C; p.gate_ld = w.cond_rows; p.row_utt = rm.row_utt; p.row_div = 2; p.gate_voice = rm.voice;
        CKG(gemm(l.wo, vA512, R2, p));
        CKN(CAT_LN, (double) (R2) * 8.0 * (C), launch_layernorm(x512, R2, C, hp.norm_eps, nullptr, nullptr, mod, w.cond_rows, of, of + C, rm, 2, vA512.hi, vA512.lo, nullptr, stream, next_dir()));
        p = gemm_params(); p.epi = EPI_SWIGLU; p.out_hi = vFF1536.hi; p.out_lo = vFF1536.lo; p.ld_h = hp.dec_ff;
        CKG(gemm(l.gu, vA512, R2, p));
        p = gemm_params(); p.epi = EPI_RESID; p.out_f32 = x512; p.ld_out = C;
        p.gate = mod + of + 2 * C; p.gate_ld = w.cond_rows; p.row_utt = rm.row_utt; p.row_div = 2; p.gate_voice = rm.voice;
        CKG(gemm(l.down, vFF1536, R2, p));
        if (taps_on()) tap_f32("dec" + std::to_string(i) + "_out", x512, R2, C);
    }
    if (taps_on()) tap_f32("decoder_x", x512, R2, C);
    // 6. final AdaLN (shift, scale only) -> fp32 in place
    {
        const int on = hp.dec_layers * 2 * 3 * C;
        CKR(CAT_LN, (double) (R2) * 8.0 * (C), launch_layernorm(x512, R2, C, hp.norm_eps, nullptr, nullptr, mod, w.cond_rows, on, on + C, rm, 2, nullptr, nullptr, x512, stream, next_dir()));
    }
    if (taps_on()) tap_f32("final_norm", x512, R2, C);
    // 7. post resnet blocks
    // The last block of a stage feeds only the next product: unless a tap wants the fp32 tensor, its conv2 epilogue adds the
    // residual and writes the (hi, lo) operand directly (no reduce-add into x, no split pass): fuse_tail
    const bool fuse_tail = !taps_on() && fuse_resblock_tail;
    for (int b = 0; b < hp.resnet_blocks; b++)
        if (!(e = run_resblock(w.post[b], x512, y512, vC512, R2, 2, "post" + std::to_string(b),
                               (fuse_tail && b == hp.resnet_blocks - 1) ? &vA512 : nullptr)).empty()) return e;
    const bool post_fused = fuse_tail && hp.resnet_blocks > 0;
    if (taps_on()) tap_f32("post", x512, R2, C);
    // 8. upsampler stage 0: ConvT(512 -> c0, kernel k0, stride f0) + bias + trim + Snake, then resblock
    if (!post_fused) CKR(CAT_SPLIT, (double) (R2) * 8.0 * (C), launch_split(x512, R2, C, rm, 2, vA512.hi, vA512.lo, stream));
    p = gemm_params(); p.epi = EPI_SNAKE; p.out_f32 = xu0; p.ld_out = f0 * c0; p.bias = w.upT_b[0]; p.bias_mod = c0;
    p.snake_ea = w.snake_ea[0]; p.snake_eb = w.snake_eb[0];
    CKG(gemm(w.upT[0], vA512, R2, p));
    if (taps_on()) tap_f32("up0_snake", xu0, Ru0, c0);
    if (!(e = run_resblock(w.up_res[0], xu0, yu0, vCu0, Ru0, rd0, "upres0", fuse_tail ? &vFFu0 : nullptr)).empty()) return e;
    if (taps_on()) tap_f32("up0", xu0, Ru0, c0);
    //    stage 1: ConvT(c0 -> c1, kernel k1, stride f1)
    if (!fuse_tail) CKR(CAT_SPLIT, (double) (Ru0) * 8.0 * (c0), launch_split(xu0, Ru0, c0, rm, rd0, vFFu0.hi, vFFu0.lo, stream));
    p = gemm_params(); p.epi = EPI_SNAKE; p.out_f32 = xu1; p.ld_out = f1 * c1; p.bias = w.upT_b[1]; p.bias_mod = c1;
    p.snake_ea = w.snake_ea[1]; p.snake_eb = w.snake_eb[1];
    CKG(gemm(w.upT[1], vFFu0, Ru0, p));
    if (taps_on()) tap_f32("up1_snake", xu1, Ru1, c1);
    if (!(e = run_resblock(w.up_res[1], xu1, yu1, vCu1, Ru1, rd1, "upres1", fuse_tail ? &vFFu1 : nullptr)).empty()) return e;
    if (taps_on()) tap_f32("up1", xu1, Ru1, c1);
    // 8b. out_proj c1 -> P + bias + Snake -> split operand of the head
    if (!fuse_tail) CKR(CAT_SPLIT, (double) (Ru1) * 8.0 * (c1), launch_split(xu1, Ru1, c1, rm, rd1, vFFu1.hi, vFFu1.lo, stream));
    p = gemm_params(); p.epi = EPI_SPLIT; p.out_hi = vH.hi; p.out_lo = vH.lo; p.ld_h = P; p.bias = w.out_proj_b;
    p.snake_ea = w.out_snake_ea; p.snake_eb = w.out_snake_eb;
    CKG(gemm(w.out_proj, vFFu1, Ru1, p));
    if (taps_on()) tap_f16("out_snake", vH.hi, vH.lo, (size_t) Ru1 * P);
    // 9. head: P -> 394, exp/clamp/cos/sin -> interleaved (re, im)
    p = gemm_params(); p.epi = EPI_HEAD; p.out_f32 = spec; p.ld_out = SPEC_LD; p.bias = w.head_b;
    CKG(gemm(w.head, vH, Ru1, p));
    if (taps_on()) tap_f32("spec", spec, Ru1, 2 * hp.n_freq, SPEC_LD);
    // ISTFT
    if (!(e = run_istft(c, spec)).empty()) return e;
    return {};
}

std::string engine::run_istft(const chunk_info & c, const float * spec_dev) {
    istft_params ip;
    ip.spec = spec_dev; ip.spec_ld = SPEC_LD; ip.wave = d_wave;
    ip.start = d_start + c.utt0; ip.len = d_len + c.utt0; ip.out_off = d_out_off + c.utt0;
    ip.chunk_cu = d_chunk_cu + c.cu0; ip.hann = w.hann; ip.tw = w.istft_tw;
    ip.n_utts = c.n_utts; ip.frames_per_code = hp.frames_per_code; ip.total_chunks = c.istft_blocks;
    double istft_bytes = 0.0;   // 394*4 read + 98*4 written per frame
    for (int u = 0; u < c.n_utts; u++) istft_bytes += (double) lay.len[c.utt0 + u] * hp.frames_per_code * 1968.0;
    CKR(CAT_ISTFT, istft_bytes, launch_istft(ip, stream));
    return {};
}

std::string engine::view512_ff(a_view & v) {
    if (!vFF512_ready) {
        std::string e = make_view(vFF512, aFF_hi, aFF_lo, max_rows1 * 6, 512);
        if (!e.empty()) return e;
        vFF512_ready = true;
    }
    v = vFF512;
    return {};
}

std::string engine::ensure_pcm() {
    if (pcm_cap >= wave_samples && d_pcm) return {};
    if (d_pcm) cudaFree(d_pcm);
    d_pcm = nullptr;
    pcm_cap = wave_samples + 64;
    return dev_alloc(&d_pcm, (size_t) pcm_cap);
}

std::string engine::run(bool peak_normalize, float * host_dst, int16_t * host_pcm) {
    std::string e = run_impl(peak_normalize, host_dst, host_pcm);
    if (!e.empty() && (host_dst || host_pcm)) {
        // an error in the middle of the delivery pipeline: no D2H copy into the caller's buffer may still be in flight
        // when the call returns
        if (copy_stream) cudaStreamSynchronize(copy_stream);
        if (stream) cudaStreamSynchronize(stream);
    }
    return e;
}

std::string engine::run_impl(bool peak_normalize, float * host_dst, int16_t * host_pcm) {
    if (!batch_uploaded) return "mio_batch_run before mio_batch_upload";
    CK(cudaSetDevice(device));
    launches = 0;
    exp_runs++;
    snaps.clear();
    // AdaLN conditioning vectors of every voice of the batch (compute_adaln3, src/miocodec.cpp:310-319)
    if (lay.n_utts > 0) CKR(CAT_COND, 0.0, launch_cond(d_voice_table, n_voices, hp.adaln_dim, w.cond_w, w.cond_b, w.cond_rows, w.cond_store, voice_silu, mod, stream));
    if (lay.n_utts > 0) launches++;   // two kernels (silu of the embeddings, then the GEMV rows)
    if (host_pcm) { std::string e = ensure_pcm(); if (!e.empty()) return e; }
    if ((host_dst || host_pcm) && !copy_stream) {
        CK(cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&chunk_done, cudaEventDisableTiming));
        for (cudaEvent_t & ev : tune_ev) CK(cudaEventCreate(&ev));
    }
    // auto-tuned tail split: time the compute and the copy of the piece in front of the tail
    const bool tune = (host_dst || host_pcm) && fetch_tail_percent < 0 && big_piece_chunk >= 0 && big_piece_chunk + 1 < (int) chunks.size();
    for (size_t ci = 0; ci < chunks.size(); ci++) {
        const chunk_info & c = chunks[ci];
        const bool timed = tune && (int) ci == big_piece_chunk;
        if (timed) CK(cudaEventRecord(tune_ev[0], stream));
        std::string e = run_chunk(c);
        if (!e.empty()) return e;
        if (peak_normalize) {
            CKR(CAT_PEAK, (double) wave_samples * 12.0, launch_peak_normalize(d_wave, d_len + c.utt0, d_out_off + c.utt0, c.n_utts, hp.samples_per_token, peak, stream));
            launches++;   // two kernels
        }
        if (host_dst || host_pcm) {
            // deliver this chunk's waveform while the next chunk computes: the chunk's utterances are contiguous in
            // the output, and chunks write disjoint ranges of d_wave
            int64_t codes = 0;
            for (int u = 0; u < c.n_utts; u++) codes += lay.len[c.utt0 + u];
            const int64_t off = lay.out_off[c.utt0], n = codes * hp.samples_per_token;
            if (host_pcm) CKR(CAT_PEAK, (double) n * 6.0, launch_pcm16(d_wave, d_pcm, off, off + n, stream));
            if (timed) CK(cudaEventRecord(tune_ev[1], stream));
            CK(cudaEventRecord(chunk_done, stream));
            CK(cudaStreamWaitEvent(copy_stream, chunk_done, 0));
            if (timed) CK(cudaEventRecord(tune_ev[2], copy_stream));
            if (n > 0 && host_dst) CK(cudaMemcpyAsync(host_dst + off, d_wave + off, (size_t) n * sizeof(float), cudaMemcpyDeviceToHost, copy_stream));
            if (n > 0 && host_pcm) CK(cudaMemcpyAsync(host_pcm + off, d_pcm + off, (size_t) n * sizeof(int16_t), cudaMemcpyDeviceToHost, copy_stream));
            if (timed) CK(cudaEventRecord(tune_ev[3], copy_stream));
        }
    }
    batch_ran = true;
    if (host_dst || host_pcm) {
        CK(cudaStreamSynchronize(copy_stream));
        std::string e = sync();
        if (e.empty() && tune) {
            // the tail (fraction f of the rows) should compute exactly while the big piece is copied:
            // f * t_compute = (1 - f) * t_copy per row  ->  f = t_copy / (t_compute + t_copy)
            float t_c = 0.f, t_d = 0.f;
            if (cudaEventElapsedTime(&t_c, tune_ev[0], tune_ev[1]) == cudaSuccess &&
                cudaEventElapsedTime(&t_d, tune_ev[2], tune_ev[3]) == cudaSuccess && t_c > 0.f && t_d > 0.f) {
                const float f = 100.f * t_d / (t_c + t_d);
                tail_auto_pct = 0.5f * tail_auto_pct + 0.5f * std::min(40.f, std::max(5.f, f));
            }
        }
        return e;
    }
    return {};
}

std::string engine::sync() {
    CK(cudaSetDevice(device));
    CK(cudaStreamSynchronize(stream));
    if (profiling) prof_collect();
    int flag = 0;
    CK(cudaMemcpy(&flag, err_flag, sizeof(int), cudaMemcpyDeviceToHost));
    if (flag != 0) {
        cudaMemset(err_flag, 0, sizeof(int));
        return "RANGE: code index out of [0, " + std::to_string(hp.vocab) + ") at padded row " + std::to_string(flag - 1);
    }
    return {};
}

std::string engine::fetch_wave(float * out, int64_t cap) {
    if (!batch_ran) return "mio_batch_fetch before mio_batch_run";
    if (cap < wave_samples) return "output capacity too small";
    std::string e = sync();
    if (!e.empty()) return e;
    if (wave_samples > 0) CK(cudaMemcpy(out, d_wave, (size_t) wave_samples * sizeof(float), cudaMemcpyDeviceToHost));
    return {};
}

// PCM16 image of the resident waveform of the last batch (mio_batch_fetch_pcm16)
std::string engine::fetch_pcm16(int16_t * out, int64_t cap) {
    if (!batch_ran) return "mio_batch_fetch before mio_batch_run";
    if (cap < wave_samples) return "output capacity too small";
    CK(cudaSetDevice(device));
    std::string e = ensure_pcm();
    if (!e.empty()) return e;
    CKR(CAT_PEAK, (double) wave_samples * 6.0, launch_pcm16(d_wave, d_pcm, 0, wave_samples, stream));
    if (!(e = sync()).empty()) return e;
    if (wave_samples > 0) CK(cudaMemcpy(out, d_pcm, (size_t) wave_samples * sizeof(int16_t), cudaMemcpyDeviceToHost));
    return {};
}

// Pre-size everything a batch of `total_codes` codes in `n_utts` utterances (and a streaming round emitting `emit_samples`
// samples) will need, so that the first real call does not pay cudaMalloc / cudaMallocHost of hundreds of MB to GBs
// (measured: up to 1 s for one pinned block, MIO_STREAM_TIMING=1).  Implemented as the upload of a dummy batch of that
// shape (no kernels run) plus the streaming emit buffers.
std::string engine::reserve(int64_t total_codes, int n_utts, int64_t emit_samples) {
    if (total_codes < 0 || n_utts < 0 || emit_samples < 0) return "bad argument";
    if (istft_only) return "bad argument: this context has no decoder";
    CK(cudaSetDevice(device));
    std::string e;
    if (total_codes > 0 && n_utts > 0) {
        n_utts = (int) std::min<int64_t>(n_utts, total_codes);
        std::vector<int32_t> codes((size_t) total_codes, 0), cu((size_t) n_utts + 1), vid((size_t) n_utts, 0);
        const int64_t per = total_codes / n_utts;
        if (per + 1 >= max_rows1) return "bad argument: utterances of that length exceed max_chunk_codes";
        for (int u = 0; u <= n_utts; u++) cu[u] = (int32_t) (u == n_utts ? total_codes : per * u);
        std::vector<float> vt((size_t) hp.adaln_dim, 0.f);
        if (!(e = upload(codes.data(), cu.data(), n_utts, vt.data(), 1, vid.data())).empty()) return e;
        if (!(e = ensure_pcm()).empty()) return e;
        if (!(e = sync()).empty()) return e;
        batch_uploaded = false; batch_ran = false;
    }
    if (emit_samples > 0) {
        if (emit_cap < emit_samples) {
            if (d_emit) cudaFree(d_emit);
            d_emit = nullptr; emit_cap = 0;
            float * p = nullptr;
            if (!(e = dev_alloc(&p, (size_t) emit_samples)).empty()) return e;
            d_emit = p; emit_cap = emit_samples;
        }
        bool pooled = true;
        if (!acquire_emit_block(emit_samples, &pooled)) return "cudaMallocHost failed (emit block)";
        // the job table of a round: one entry per firing stream, at most max_voices streams per pass
        const int64_t want_jobs = (int64_t) std::max(max_voices, n_utts) + 64;
        if (emit_jobs_cap < want_jobs) {
            if (d_emit_jobs) cudaFree(d_emit_jobs);
            d_emit_jobs = nullptr; emit_jobs_cap = 0;
            if (cudaMalloc(&d_emit_jobs, (size_t) want_jobs * EMIT_JOB_BYTES) != cudaSuccess) { d_emit_jobs = nullptr; return "cudaMalloc failed (emit job table)"; }
            emit_jobs_cap = want_jobs;
        }
    }
    return {};
}

// PCM16 image of the resident waveform, left ON THE DEVICE (mio_batch_pcm16_device): converted on the context's stream,
// valid until the next upload; the multi-GPU gather sends it to the collecting rank over NVLink.
std::string engine::pcm16_device(int16_t ** out) {
    *out = nullptr;
    if (!batch_ran) return "mio_batch_pcm16_device before mio_batch_run";
    CK(cudaSetDevice(device));
    std::string e = ensure_pcm();
    if (!e.empty()) return e;
    CKR(CAT_PEAK, (double) wave_samples * 6.0, launch_pcm16(d_wave, d_pcm, 0, wave_samples, stream));
    *out = d_pcm;
    return {};
}

// The same conversion with the result stored straight into a caller-supplied DEVICE buffer -- which may be another GPU's
// memory mapped into this process (CUDA IPC / peer access): the stores then travel over NVLink and the sharded sweep's
// gather needs no separate transfer step (mio_batch_pcm16_to_device; sharding.P2PGather).
std::string engine::pcm16_to_device(int16_t * dst, int64_t cap) {
    if (!batch_ran) return "mio_batch_pcm16_to_device before mio_batch_run";
    if (!dst || cap < wave_samples) return "bad argument: destination too small";
    if (((uintptr_t) dst & 15) != 0) return "bad argument: destination must be 16-byte aligned (8 samples)";
    CK(cudaSetDevice(device));
    CKR(CAT_PEAK, (double) wave_samples * 6.0, launch_pcm16(d_wave, dst, 0, wave_samples, stream));
    return {};
}

// spectrogram of utterance `utt` of a single-chunk batch (mio_decode_spec)
std::string engine::fetch_spec(int utt, float * out, int64_t cap) {
    if (!batch_ran) return "fetch before run";
    if (chunks.size() != 1 || utt < 0 || utt >= lay.n_utts) return "fetch_spec needs a single-chunk batch";
    std::string e = sync();
    if (!e.empty()) return e;
    const int fpc = hp.frames_per_code;
    const size_t n = (size_t) lay.len[utt] * fpc * 2 * hp.n_freq;
    if ((int64_t) n > cap) return "output capacity too small";
    if (n) CK(cudaMemcpy2D(out, (size_t) 2 * hp.n_freq * sizeof(float), spec + (size_t) lay.start[utt] * fpc * SPEC_LD,
                           (size_t) SPEC_LD * sizeof(float), (size_t) 2 * hp.n_freq * sizeof(float), (size_t) lay.len[utt] * fpc,
                           cudaMemcpyDeviceToHost));
    return {};
}

// standalone ISTFT of a host spectrogram (mio_istft): one utterance of n_frames frames
std::string engine::run_istft_only(const float * spec_host, int n_frames) {
    std::string e;
    CK(cudaSetDevice(device));
    if (n_frames <= 0) return "bad argument";
    if (!err_flag) { if (!(e = dev_alloc(&err_flag, 1)).empty()) return e; CK(cudaMemset(err_flag, 0, sizeof(int))); }
    // own staging buffer, grown with the call: never the decoder's workspace (mio_istft on a codec context must not
    // need 8 GB of activations, and an ISTFT-only context stays a few MB)
    if ((int64_t) istft_spec_rows < (int64_t) n_frames) {
        if (stream) CK(cudaStreamSynchronize(stream));
        if (istft_spec) cudaFree(istft_spec);
        istft_spec = nullptr;
        istft_spec_rows = (size_t) n_frames + (size_t) n_frames / 4 + 64;
        if (!(e = dev_alloc(&istft_spec, istft_spec_rows * SPEC_LD)).empty()) { istft_spec_rows = 0; return e; }
    }
    float * const spec = istft_spec;                       // shadows the decoder's spectrogram buffer in this function
    batch_uploaded = false; batch_ran = false;
    // layout: a single pseudo-utterance whose length is counted in frames (frames_per_code = 1)
    const int32_t start0 = 0, len0 = n_frames;
    const int64_t off0 = 0;
    const int blocks = (n_frames + ISTFT_SEG_PER_BLOCK - 1) / ISTFT_SEG_PER_BLOCK;
    const int32_t ccu[2] = {0, blocks};
    if (!d_start) {
        cap_utts = 64;
        if (!(e = dev_alloc(&d_start, cap_utts)).empty() || !(e = dev_alloc(&d_len, cap_utts)).empty() ||
            !(e = dev_alloc(&d_voice, cap_utts)).empty() || !(e = dev_alloc(&d_out_off, cap_utts)).empty()) return e;
    }
    if (!d_chunk_cu) { cap_ccu = 64; if (!(e = dev_alloc(&d_chunk_cu, cap_ccu)).empty()) return e; }
    wave_samples = (int64_t) n_frames * hp.hop;
    if (wave_cap < wave_samples) {
        if (d_wave) cudaFree(d_wave);
        d_wave = nullptr; wave_cap = wave_samples + 64;
        if (!(e = dev_alloc(&d_wave, (size_t) wave_cap)).empty()) return e;
    }
    CK(cudaStreamSynchronize(stream));
    // all copies are ordered on the context stream (a blocking cudaMemcpy from pageable memory may return before
    // its DMA lands, and the context stream does not synchronise with the legacy default stream)
    CK(cudaMemcpyAsync(d_start, &start0, 4, cudaMemcpyHostToDevice, stream));
    CK(cudaMemcpyAsync(d_len, &len0, 4, cudaMemcpyHostToDevice, stream));
    CK(cudaMemcpyAsync(d_out_off, &off0, 8, cudaMemcpyHostToDevice, stream));
    CK(cudaMemcpyAsync(d_chunk_cu, ccu, 8, cudaMemcpyHostToDevice, stream));
    CK(cudaMemcpy2DAsync(spec, (size_t) SPEC_LD * sizeof(float), spec_host, (size_t) 2 * hp.n_freq * sizeof(float),
                         (size_t) 2 * hp.n_freq * sizeof(float), (size_t) n_frames, cudaMemcpyHostToDevice, stream));
    CK(cudaStreamSynchronize(stream));   // the small host sources above are locals
    istft_params ip;
    ip.spec = spec; ip.spec_ld = SPEC_LD; ip.wave = d_wave; ip.start = d_start; ip.len = d_len; ip.out_off = d_out_off; ip.chunk_cu = d_chunk_cu;
    ip.hann = w.hann; ip.tw = w.istft_tw; ip.n_utts = 1; ip.frames_per_code = 1; ip.total_chunks = blocks;
    const double istft_bytes = (double) n_frames * 1968.0;
    launches = 0;
    CKR(CAT_ISTFT, istft_bytes, launch_istft(ip, stream));
    lay = batch_layout();
    batch_ran = true;
    return {};
}

}  // namespace mio
