// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// Thin extern "C" wrapper around the UNMODIFIED reference implementation of the
// hot path: miocodec_decode() (/root/reference/src/miocodec.cpp:519-810) and
// istft() (/root/reference/src/istft.cpp:68-108).  The reference sources are
// compiled where they lie under /root/reference by oracle/build_ref.sh; nothing
// from them is copied into this repository.  The resulting shared object lives
// in oracle/_ref/ (git-ignored) and is used by tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference arm as the checker and the
// timed CPU baseline.
#include "miocodec.h"   // reference header, found via -I/root/reference/src
#include "istft.h"      // reference header
#include "token-parser.h"   // reference header (pre-path ingest)
#include "wav-writer.h"     // reference header (post-path epilogue)
#include "test-to-speech.h" // reference header: TestToSpeech (streaming policy)

#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <unistd.h>
#include <fcntl.h>

namespace {
// The reference logs several lines to stderr on every decode
// (src/miocodec.cpp:551,747,784-785); silence them unless MIO_REF_VERBOSE is set.
struct stderr_mute {
    int saved = -1;
    stderr_mute() {
        if (getenv("MIO_REF_VERBOSE")) return;
        fflush(stderr);
        saved = dup(2);
        int nul = open("/dev/null", O_WRONLY);
        if (nul >= 0) { dup2(nul, 2); close(nul); }
    }
    ~stderr_mute() {
        if (saved >= 0) { fflush(stderr); dup2(saved, 2); close(saved); }
    }
};
}  // namespace

extern "C" {

void * mio_ref_load(const char * gguf_path) {
    stderr_mute m;
    return (void *) miocodec_load(std::string(gguf_path));
}

void mio_ref_free(void * ctx) { miocodec_free((miocodec_context *) ctx); }

int mio_ref_sample_rate(void * ctx)       { return miocodec_sample_rate((miocodec_context *) ctx); }
int mio_ref_n_fft(void * ctx)             { return miocodec_n_fft((miocodec_context *) ctx); }
int mio_ref_hop_length(void * ctx)        { return miocodec_hop_length((miocodec_context *) ctx); }
int mio_ref_samples_per_token(void * ctx) { return miocodec_samples_per_token((miocodec_context *) ctx); }

// Returns number of floats written (n_frames * 394) or -1. spec_out may be NULL to query.
long mio_ref_decode(void * ctx, const int * codes, int n_codes, const float * emb,
                    float * spec_out, long spec_cap, int * n_frames, double * seconds) {
    stderr_mute m;
    int nf = 0;
    auto t0 = std::chrono::steady_clock::now();
    std::vector<float> spec = miocodec_decode((miocodec_context *) ctx, codes, n_codes, emb, 0, &nf);
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    if (n_frames) *n_frames = nf;
    if (spec.empty()) return -1;
    if (spec_out) {
        if ((long) spec.size() > spec_cap) return -2;
        memcpy(spec_out, spec.data(), spec.size() * sizeof(float));
    }
    return (long) spec.size();
}

// istft() with the reference's own cache (win_length == n_fft as in
// src/test-to-speech.cpp:67). Returns samples written or -1.
long mio_ref_istft(const float * spec, int n_frames, int n_fft, int hop, float * out, long out_cap,
                   double * seconds) {
    static istft_cache * cache = nullptr;
    static int cache_n = 0;
    if (!cache || cache_n != n_fft) { delete cache; cache = new istft_cache(n_fft, n_fft); cache_n = n_fft; }
    auto t0 = std::chrono::steady_clock::now();
    std::vector<float> a = istft(spec, n_frames, hop, *cache);
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    if (a.empty()) return -1;
    if ((long) a.size() > out_cap) return -2;
    memcpy(out, a.data(), a.size() * sizeof(float));
    return (long) a.size();
}

// the tables of the reference's istft_cache (src/istft.cpp:7-41), copied out through its own getters.
// Returns n_mid; cos/sin need n_fft * n_mid floats, nyq n_fft, hann win_length.
int mio_ref_istft_tables(int n_fft, int win_length, float * cos_t, float * sin_t, float * nyq, float * hann, int * dims /*[4]*/) {
    istft_cache c(n_fft, win_length);
    if (dims) { dims[0] = c.n_fft(); dims[1] = c.win_length(); dims[2] = c.n_freq(); dims[3] = c.n_mid(); }
    if (cos_t) memcpy(cos_t, c.cos_table().data(), c.cos_table().size() * sizeof(float));
    if (sin_t) memcpy(sin_t, c.sin_table().data(), c.sin_table().size() * sizeof(float));
    if (nyq) memcpy(nyq, c.nyquist_sign().data(), c.nyquist_sign().size() * sizeof(float));
    if (hann) memcpy(hann, c.hann_window().data(), c.hann_window().size() * sizeof(float));
    return c.n_mid();
}

// codes -> waveform, exactly what decode_codes_to_audio times
// (src/test-to-speech.cpp:264-287), peak normalisation off.
long mio_ref_decode_wave(void * ctx, const int * codes, int n_codes, const float * emb,
                         float * out, long out_cap, double * codec_sec, double * istft_sec) {
    int nf = 0;
    std::vector<float> spec;
    {
        stderr_mute m;
        auto t0 = std::chrono::steady_clock::now();
        spec = miocodec_decode((miocodec_context *) ctx, codes, n_codes, emb, 0, &nf);
        auto t1 = std::chrono::steady_clock::now();
        if (codec_sec) *codec_sec = std::chrono::duration<double>(t1 - t0).count();
    }
    if (spec.empty()) return -1;
    miocodec_context * c = (miocodec_context *) ctx;
    return mio_ref_istft(spec.data(), nf, miocodec_n_fft(c), miocodec_hop_length(c), out, out_cap, istft_sec);
}

int mio_ref_load_voice(const char * path, float * out, int cap) {
    stderr_mute m;
    std::vector<float> e = load_voice_embedding(std::string(path));
    if (e.empty() || (int) e.size() > cap) return -1;
    memcpy(out, e.data(), e.size() * sizeof(float));
    return (int) e.size();
}

// parse_speech_tokens (/root/reference/src/token-parser.cpp:5-28). Returns the number of codes found; writes
// min(count, cap) of them.
long mio_ref_parse_speech_tokens(const char * text, long text_len, int * out, long cap) {
    std::vector<int> codes = parse_speech_tokens(std::string(text, (size_t) text_len));
    for (long i = 0; i < (long) codes.size() && i < cap; i++) out[i] = codes[i];
    return (long) codes.size();
}

// The reference's own streaming session: TestToSpeech::synthesize_stream_profiled, LLM branch
// (/root/reference/src/test-to-speech.cpp:430-625: maybe_emit :507-571, emit_range :367-417), with oracle/stub_llama.cpp
// emitting `codes` one per generated token.  skip_llm != 0 runs the skip_llm branch (:419-428) on "<|s_N|>" text instead.
// Writes the concatenated emitted audio to out, the size of every callback chunk to chunk_sizes, and
// profile[0..4] = decode_calls, decoded_codes, emitted_samples, llm_tokens, number of callbacks with is_last_chunk.
// Returns the number of emitted samples or < 0.
long mio_ref_stream_session(const char * codec_gguf, const char * voice_path, const int * codes, int n_codes, int skip_llm,
                            long chunk_samples, float * out, long out_cap, long * chunk_sizes, long chunk_cap, long * n_chunks,
                            long long * profile) {
    stderr_mute m;
    char script[] = "/tmp/mio_ref_scriptXXXXXX";
    std::string text;
    TestToSpeech::Config cfg;
    cfg.codec_path = codec_gguf;
    if (!skip_llm) {
        int fd = mkstemp(script);
        if (fd < 0) return -1;
        FILE * f = fdopen(fd, "w");
        for (int i = 0; i < n_codes; i++) fprintf(f, "%d\n", codes[i]);
        fclose(f);
        cfg.model_path = script;
        text = "scripted";
    } else {
        for (int i = 0; i < n_codes; i++) text += "<|s_" + std::to_string(codes[i]) + "|>";
    }
    long total = 0, nchunks = 0, lasts = 0;
    bool overflow = false;
    long rc = -2;
    {
        TestToSpeech tts(cfg);
        VoiceModel voice;
        if (tts.is_ready() && voice.load_from_file(voice_path)) {
            TestToSpeech::Options opt;
            opt.skip_llm = skip_llm != 0;
            opt.max_tokens = n_codes + 8 > 700 ? n_codes + 8 : 700;
            opt.apply_peak_normalization = false;
            TestToSpeech::StreamProfile prof;
            const bool ok = tts.synthesize_stream_profiled(
                voice, text,
                [&](const float * s, size_t n, int, bool is_last) -> bool {
                    if (total + (long) n > out_cap) overflow = true;
                    else if (n > 0) memcpy(out + total, s, n * sizeof(float));
                    total += (long) n;
                    if (nchunks < chunk_cap) chunk_sizes[nchunks] = (long) n;
                    nchunks++;
                    if (is_last) lasts++;
                    return true;
                },
                (size_t) chunk_samples, opt, prof);
            if (ok && !overflow) {
                profile[0] = prof.decode_calls; profile[1] = (long long) prof.decoded_codes;
                profile[2] = (long long) prof.emitted_samples; profile[3] = prof.llm_tokens; profile[4] = lasts;
                rc = total;
            } else rc = overflow ? -3 : -4;
        }
    }
    if (!skip_llm) unlink(script);
    if (n_chunks) *n_chunks = nchunks;
    return rc;
}

// TestToSpeech::synthesize_to_vector / synthesize_to_file with skip_llm (/root/reference/src/test-to-speech.cpp:248-345):
// `text` is raw "<|s_N|>" token text (anything else in it is skipped by the reference's own parser).  wav_path != NULL
// writes the file through the reference's writer; out (may be NULL) receives the samples.  Returns the sample count,
// -1 when the reference refuses the request (e.g. no codes parse), < -1 on set-up errors.
long mio_ref_tts_skip_llm(const char * codec_gguf, const char * voice_path, const char * text, int peak_normalize,
                          const char * wav_path, float * out, long out_cap) {
    stderr_mute m;
    TestToSpeech::Config cfg;
    cfg.codec_path = codec_gguf;
    TestToSpeech tts(cfg);
    VoiceModel voice;
    if (!tts.is_ready() || !voice.load_from_file(voice_path)) return -2;
    TestToSpeech::Options opt;
    opt.skip_llm = true;
    opt.apply_peak_normalization = peak_normalize != 0;
    std::vector<float> audio;
    if (!tts.synthesize_to_vector(voice, text, audio, opt)) return -1;
    if (wav_path && !tts.synthesize_to_file(voice, text, wav_path, opt)) return -1;
    if (out) {
        if ((long) audio.size() > out_cap) return -3;
        memcpy(out, audio.data(), audio.size() * sizeof(float));
    }
    return (long) audio.size();
}

// wav_write (/root/reference/src/wav-writer.cpp:24-45): 44-byte RIFF header + PCM16 (clamp(s*32767) truncated).
int mio_ref_wav_write(const char * path, const float * samples, long n, int sample_rate) {
    return wav_write(std::string(path), std::vector<float>(samples, samples + n), sample_rate) ? 0 : -1;
}

}  // extern "C"
