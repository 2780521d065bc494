// BatchTestToSpeech (shim/batch-tts.h) on top of the C ABI.  Host logic only: parsing, batch assembly, callback
// delivery; every sample is produced by the device kernels behind mio_decode_*_batch / mio_streams_step.
#include "batch-tts.h"

#include <algorithm>
#include <cstdio>
#include <map>

#include "../../include/miocodec_b200.h"

struct BatchTestToSpeech::impl {
    mio_ctx * ctx = nullptr;
    int sample_rate = 0, spt = 0;
    struct live { mio_stream * s = nullptr; StreamCallback cb; size_t chunk = 4096; bool finished = false; };
    std::map<int, live> streams;
    int next_id = 1;
    std::vector<float> buf;

    // the assembled ragged batch of the utterances that passed the reference's checks
    struct batch {
        std::vector<int32_t> codes, cu{0}, vid;
        std::vector<float> vt;
        std::vector<int> index;                       // batch slot -> caller index
    };
    // mirrors decode_tokens_to_audio's checks (src/test-to-speech.cpp:205-214)
    bool assemble(const std::vector<std::vector<float>> & voices, const std::vector<std::string> & texts, batch & b,
                  std::vector<bool> & ok) {
        if (voices.size() != texts.size()) { fprintf(stderr, "BatchTestToSpeech: %zu voices for %zu texts\n", voices.size(), texts.size()); return false; }
        ok.assign(texts.size(), false);
        for (size_t i = 0; i < texts.size(); i++) {
            if (voices[i].size() != 128) { fprintf(stderr, "TestToSpeech: voice model is not ready\n"); continue; }
            const int64_t n = mio_parse_speech_tokens(texts[i].data(), (int64_t) texts[i].size(), nullptr, 0);
            if (n <= 0) { fprintf(stderr, "TestToSpeech: no speech codes parsed from text\n"); continue; }
            const size_t at = b.codes.size();
            b.codes.resize(at + (size_t) n);
            mio_parse_speech_tokens(texts[i].data(), (int64_t) texts[i].size(), b.codes.data() + at, n);
            b.cu.push_back((int32_t) b.codes.size());
            b.vid.push_back((int32_t) b.index.size());
            b.vt.insert(b.vt.end(), voices[i].begin(), voices[i].end());
            b.index.push_back((int) i);
        }
        return true;
    }
};

BatchTestToSpeech::BatchTestToSpeech(const std::string & codec_path, int device) : p_(new impl()) {
    if (mio_ctx_create(codec_path.c_str(), device, &p_->ctx) != MIO_OK) {
        fprintf(stderr, "TestToSpeech: failed to load codec: %s\n", mio_last_error());
        p_->ctx = nullptr;
        return;
    }
    p_->sample_rate = mio_ctx_sample_rate(p_->ctx);
    p_->spt = mio_ctx_samples_per_token(p_->ctx);
}

BatchTestToSpeech::~BatchTestToSpeech() {
    for (auto & kv : p_->streams) mio_stream_close(kv.second.s);
    if (p_->ctx) mio_ctx_destroy(p_->ctx);
}

bool BatchTestToSpeech::is_ready() const { return p_->ctx != nullptr; }
int BatchTestToSpeech::sample_rate() const { return p_->sample_rate; }

bool BatchTestToSpeech::synthesize_to_vectors(const std::vector<std::vector<float>> & voices, const std::vector<std::string> & texts,
                                              std::vector<std::vector<float>> & out, std::vector<bool> & ok, const Options & opt) {
    out.assign(texts.size(), {});
    impl::batch b;
    if (!is_ready() || !p_->assemble(voices, texts, b, ok)) { ok.assign(texts.size(), false); return false; }
    const int B = (int) b.index.size();
    if (B > 0) {
        std::vector<float> & wave = p_->buf;
        wave.resize(b.codes.size() * (size_t) p_->spt);
        if (mio_decode_wave_batch(p_->ctx, b.codes.data(), b.cu.data(), B, b.vt.data(), B, b.vid.data(),
                                  opt.apply_peak_normalization ? 1 : 0, wave.data(), (int64_t) wave.size()) != MIO_OK) {
            fprintf(stderr, "TestToSpeech: codec decode failed: %s\n", mio_last_error());
            return false;
        }
        for (int j = 0; j < B; j++) {
            out[b.index[j]].assign(wave.begin() + (size_t) b.cu[j] * p_->spt, wave.begin() + (size_t) b.cu[j + 1] * p_->spt);
            ok[b.index[j]] = true;
        }
    }
    return std::all_of(ok.begin(), ok.end(), [](bool v) { return v; });
}

bool BatchTestToSpeech::synthesize_to_files(const std::vector<std::vector<float>> & voices, const std::vector<std::string> & texts,
                                            const std::vector<std::string> & paths, std::vector<bool> & ok, const Options & opt) {
    impl::batch b;
    if (!is_ready() || paths.size() != texts.size() || !p_->assemble(voices, texts, b, ok)) { ok.assign(texts.size(), false); return false; }
    const int B = (int) b.index.size();
    if (B > 0) {
        std::vector<int16_t> pcm(b.codes.size() * (size_t) p_->spt);
        if (mio_decode_pcm16_batch(p_->ctx, b.codes.data(), b.cu.data(), B, b.vt.data(), B, b.vid.data(),
                                   opt.apply_peak_normalization ? 1 : 0, pcm.data(), (int64_t) pcm.size()) != MIO_OK) {
            fprintf(stderr, "TestToSpeech: codec decode failed: %s\n", mio_last_error());
            return false;
        }
        for (int j = 0; j < B; j++) {
            const size_t off = (size_t) b.cu[j] * p_->spt, n = (size_t) (b.cu[j + 1] - b.cu[j]) * p_->spt;
            if (mio_wav_write_pcm16(paths[b.index[j]].c_str(), pcm.data() + off, (int64_t) n, p_->sample_rate) == MIO_OK) ok[b.index[j]] = true;
            else fprintf(stderr, "TestToSpeech: failed to write %s\n", paths[b.index[j]].c_str());
        }
    }
    return std::all_of(ok.begin(), ok.end(), [](bool v) { return v; });
}

bool BatchTestToSpeech::synthesize_streams(const std::vector<std::vector<float>> & voices, const std::vector<std::string> & texts,
                                           const std::vector<StreamCallback> & callbacks, size_t chunk_samples, std::vector<bool> & ok) {
    if (chunk_samples == 0) chunk_samples = 4096;                      // src/test-to-speech.cpp:360-362
    std::vector<std::vector<float>> audio;
    Options opt;
    opt.apply_peak_normalization = false;                              // decode_tokens_to_audio(voice, text, audio, false), :420
    if (callbacks.size() != texts.size()) { ok.assign(texts.size(), false); return false; }
    synthesize_to_vectors(voices, texts, audio, ok, opt);
    for (size_t i = 0; i < texts.size(); i++) {
        if (!ok[i]) continue;
        if (!callbacks[i]) { ok[i] = false; continue; }                // :356-358
        // emit_range(audio, 0, n, true) with an empty crossfade tail (:367-417): plain chunking, last chunk flagged
        const std::vector<float> & a = audio[i];
        for (size_t off = 0; off < a.size(); off += chunk_samples) {
            const size_t n = std::min(chunk_samples, a.size() - off);
            if (!callbacks[i](a.data() + off, n, p_->sample_rate, off + n >= a.size())) { ok[i] = false; break; }
        }
    }
    return std::all_of(ok.begin(), ok.end(), [](bool v) { return v; });
}

int BatchTestToSpeech::open_stream(const std::vector<float> & voice, const StreamCallback & cb, size_t chunk_samples) {
    if (!is_ready() || voice.size() != 128 || !cb) return -1;
    mio_stream_config cfg;
    mio_stream_default_config(p_->ctx, &cfg);
    impl::live l;
    if (mio_stream_open(p_->ctx, voice.data(), &cfg, &l.s) != MIO_OK) { fprintf(stderr, "TestToSpeech: %s\n", mio_last_error()); return -1; }
    l.cb = cb;
    l.chunk = chunk_samples ? chunk_samples : 4096;
    const int id = p_->next_id++;
    p_->streams[id] = l;
    return id;
}

bool BatchTestToSpeech::push_code(int id, int code) {
    auto it = p_->streams.find(id);
    if (it == p_->streams.end() || it->second.finished) return false;
    const int32_t c = code;
    return mio_stream_push_codes(it->second.s, &c, 1) == MIO_OK;
}

bool BatchTestToSpeech::finish_stream(int id) {
    auto it = p_->streams.find(id);
    if (it == p_->streams.end()) return false;
    it->second.finished = true;
    return mio_stream_finish(it->second.s) == MIO_OK;
}

bool BatchTestToSpeech::poll() {
    if (p_->streams.empty()) return true;
    std::vector<mio_stream *> hs;
    for (auto & kv : p_->streams) hs.push_back(kv.second.s);
    if (mio_streams_step(p_->ctx, hs.data(), (int) hs.size()) != MIO_OK) { fprintf(stderr, "TestToSpeech: %s\n", mio_last_error()); return false; }
    std::vector<int> closed;
    for (auto & kv : p_->streams) {
        impl::live & l = kv.second;
        p_->buf.resize(l.chunk);
        for (;;) {
            int last = 0;
            const int64_t n = mio_stream_read(l.s, p_->buf.data(), (int64_t) l.chunk, &last);
            if (n <= 0 && !last) break;
            if (n > 0 || last) {
                if (!l.cb(p_->buf.data(), (size_t) std::max<int64_t>(n, 0), p_->sample_rate, last != 0)) { closed.push_back(kv.first); break; }
            }
            if (last) { closed.push_back(kv.first); break; }
        }
    }
    for (int id : closed) { mio_stream_close(p_->streams[id].s); p_->streams.erase(id); }
    return true;
}

size_t BatchTestToSpeech::live_streams() const { return p_->streams.size(); }

bool BatchTestToSpeech::stream_profile(int id, int & decode_calls, size_t & decoded_codes, size_t & emitted_samples) const {
    auto it = p_->streams.find(id);
    if (it == p_->streams.end()) return false;
    int64_t a = 0, b = 0, c = 0;
    if (mio_stream_stats(it->second.s, &a, &b, &c) != MIO_OK) return false;
    decode_calls = (int) a; decoded_codes = (size_t) b; emitted_samples = (size_t) c;
    return true;
}
