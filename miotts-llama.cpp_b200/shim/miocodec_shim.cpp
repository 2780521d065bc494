// C++ drop-in layer: the reference's miocodec.h / istft.h entry points on top of the C ABI.
// Error behaviour mirrors the reference: diagnostics on stderr, nullptr / empty vector / false to the caller,
// no exceptions (src/miocodec.cpp:433,441,486,562,751,757,787; src/istft.cpp:103-105).
#include "miocodec.h"
#include "istft.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "../../include/miocodec_b200.h"

struct miocodec_context {
    mio_ctx * c = nullptr;
};

namespace {
int device_ordinal() {
    const char * e = getenv("MIO_DEVICE");
    return e ? atoi(e) : 0;
}

// istft() of the reference is a pure, thread-safe free function of (spec, cache) (src/istft.h:38-42).  It therefore gets
// its OWN ISTFT-only device context per calling thread (a few MB: tables + a spectrogram staging buffer that grows with
// the call), never a codec context: a concurrent miocodec_decode / streaming session / miocodec_free on another thread
// cannot be disturbed by it, and it works before any codec is loaded.
struct istft_thread_ctx {
    mio_ctx * c = nullptr;
    int device = -1;
    ~istft_thread_ctx() { if (c) mio_ctx_destroy(c); }
};
thread_local istft_thread_ctx t_istft;
}  // namespace

miocodec_context * miocodec_load(const std::string & model_path) {
    mio_ctx * c = nullptr;
    if (mio_ctx_create(model_path.c_str(), device_ordinal(), &c) != MIO_OK) {
        fprintf(stderr, "miocodec: failed to load %s: %s\n", model_path.c_str(), mio_last_error());
        return nullptr;
    }
    miocodec_context * ctx = new miocodec_context();
    ctx->c = c;
    return ctx;
}

void miocodec_free(miocodec_context * ctx) {
    if (!ctx) return;
    mio_ctx_destroy(ctx->c);
    delete ctx;
}

std::vector<float> miocodec_decode(miocodec_context * ctx, const int * codes, int n_codes, const float * global_emb,
                                   int /*audio_length: ignored like the reference*/, int * out_n_frames) {
    if (out_n_frames) *out_n_frames = 0;
    if (!ctx || !codes || !global_emb || n_codes <= 0) {
        fprintf(stderr, "miocodec: decode called with invalid arguments\n");
        return {};
    }
    int n_frames = 0;
    if (mio_decode_spec(ctx->c, codes, n_codes, global_emb, nullptr, 0, &n_frames) != MIO_OK) return {};
    std::vector<float> spec((size_t) n_frames * (size_t) (mio_ctx_n_fft(ctx->c) + 2));
    if (mio_decode_spec(ctx->c, codes, n_codes, global_emb, spec.data(), (int64_t) spec.size(), &n_frames) != MIO_OK) {
        fprintf(stderr, "miocodec: decode failed: %s\n", mio_last_error());
        return {};
    }
    if (out_n_frames) *out_n_frames = n_frames;
    return spec;
}

int miocodec_sample_rate(const miocodec_context * ctx) { return ctx ? mio_ctx_sample_rate(ctx->c) : 0; }
int miocodec_n_fft(const miocodec_context * ctx) { return ctx ? mio_ctx_n_fft(ctx->c) : 0; }
int miocodec_hop_length(const miocodec_context * ctx) { return ctx ? mio_ctx_hop_length(ctx->c) : 0; }
int miocodec_samples_per_token(const miocodec_context * ctx) { return ctx ? mio_ctx_samples_per_token(ctx->c) : 0; }

std::vector<float> load_voice_embedding(const std::string & path) {
    std::vector<float> v(4096);
    const int n = mio_voice_load(path.c_str(), v.data(), (int) v.size());
    if (n < 0) {
        fprintf(stderr, "load_voice_embedding: %s: %s\n", path.c_str(), mio_last_error());
        return {};
    }
    v.resize((size_t) n);
    return v;
}

void miocodec_print_tensors(const std::string & path) {
    if (mio_print_tensors(path.c_str()) != MIO_OK) fprintf(stderr, "miocodec: %s\n", mio_last_error());
}

bool miocodec_decode_audio_batch(miocodec_context * ctx, const std::vector<std::vector<int>> & codes,
                                 const std::vector<std::vector<float>> & voice_embs, bool peak_normalize,
                                 std::vector<std::vector<float>> & out_audio) {
    out_audio.clear();
    if (!ctx || codes.size() != voice_embs.size()) return false;
    const int B = (int) codes.size();
    if (B == 0) return true;
    std::vector<int32_t> cat, cu(1, 0), vid(B);
    std::vector<float> vt;
    for (int b = 0; b < B; b++) {
        if (voice_embs[b].size() != 128) { fprintf(stderr, "miocodec: voice embedding %d is not 128 floats\n", b); return false; }
        cat.insert(cat.end(), codes[b].begin(), codes[b].end());
        cu.push_back((int32_t) cat.size());
        vid[b] = b;
        vt.insert(vt.end(), voice_embs[b].begin(), voice_embs[b].end());
    }
    const int spt = mio_ctx_samples_per_token(ctx->c);
    std::vector<float> wave((size_t) cat.size() * (size_t) spt);
    if (mio_decode_wave_batch(ctx->c, cat.data(), cu.data(), B, vt.data(), B, vid.data(), peak_normalize ? 1 : 0, wave.data(),
                              (int64_t) wave.size()) != MIO_OK) {
        fprintf(stderr, "miocodec: batch decode failed: %s\n", mio_last_error());
        return false;
    }
    out_audio.resize(B);
    for (int b = 0; b < B; b++) out_audio[b].assign(wave.begin() + (size_t) cu[b] * spt, wave.begin() + (size_t) cu[b + 1] * spt);
    return true;
}

// ------------------------------------------------------------------------------------------------ istft.h
// Table formulae follow src/istft.cpp:7-32 (angle formed in f32, cosf/sinf of libm).
istft_cache::istft_cache(int n_fft, int win_length) {
    n_fft_ = n_fft;
    win_length_ = win_length;
    n_freq_ = n_fft / 2 + 1;
    n_mid_ = std::max(0, n_freq_ - 2);
    const size_t cells = (size_t) std::max(0, n_fft) * (size_t) n_mid_;
    cos_table_.resize(cells);
    sin_table_.resize(cells);
    nyquist_sign_.resize((size_t) std::max(0, n_fft));
    hann_window_.resize((size_t) std::max(0, win_length));
    const float step = 2.0f * (float) M_PI / (float) n_fft;
    for (int n = 0; n < n_fft; n++) {
        nyquist_sign_[n] = (n % 2 == 0) ? 1.0f : -1.0f;
        float * c = cos_table_.data() + (size_t) n * n_mid_;
        float * s = sin_table_.data() + (size_t) n * n_mid_;
        for (int k = 1; k <= n_mid_; k++) {
            const float ang = step * (float) k * (float) n;
            c[k - 1] = cosf(ang);
            s[k - 1] = sinf(ang);
        }
    }
    for (int i = 0; i < win_length; i++) hann_window_[i] = 0.5f * (1.0f - cosf(2.0f * (float) M_PI * i / win_length));
}
int istft_cache::n_fft() const { return n_fft_; }
int istft_cache::win_length() const { return win_length_; }
int istft_cache::n_freq() const { return n_freq_; }
int istft_cache::n_mid() const { return n_mid_; }
const std::vector<float> & istft_cache::cos_table() const { return cos_table_; }
const std::vector<float> & istft_cache::sin_table() const { return sin_table_; }
const std::vector<float> & istft_cache::nyquist_sign() const { return nyquist_sign_; }
const std::vector<float> & istft_cache::hann_window() const { return hann_window_; }

std::vector<float> istft(const float * spec, int n_frames, int hop_length, const istft_cache & cache) {
    if (!spec || n_frames <= 0) return {};      // the reference returns {} when trim_end <= trim_start
    if (cache.n_fft() != 392 || cache.win_length() != 392 || hop_length != 98) {
        fprintf(stderr, "istft: unsupported geometry n_fft=%d win=%d hop=%d (device kernels are built for 392/392/98; no CPU fallback)\n",
                cache.n_fft(), cache.win_length(), hop_length);
        return {};
    }
    const int dev = device_ordinal();
    if (t_istft.c && t_istft.device != dev) { mio_ctx_destroy(t_istft.c); t_istft.c = nullptr; }
    if (!t_istft.c) {
        if (mio_istft_ctx_create(dev, &t_istft.c) != MIO_OK) {
            fprintf(stderr, "istft: %s\n", mio_last_error());
            t_istft.c = nullptr;
            return {};
        }
        t_istft.device = dev;
    }
    mio_ctx * c = t_istft.c;
    std::vector<float> wave((size_t) n_frames * (size_t) hop_length);
    int64_t n = 0;
    if (mio_istft(c, spec, n_frames, wave.data(), (int64_t) wave.size(), &n) != MIO_OK) {
        fprintf(stderr, "istft: %s\n", mio_last_error());
        return {};
    }
    wave.resize((size_t) n);
    return wave;
}
