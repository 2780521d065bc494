// Test program for shim/batch-tts.h: drives the batched skip_llm calls and the live stream pool the
// way a TestToSpeech-style caller would and writes raw f32 results for tests/test_batch_tts.py.
//   batch_tts_demo <miocodec.gguf> <voice0.emb.gguf> <voice1.emb.gguf> <out_prefix>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "batch-tts.h"
#include "miocodec.h"

static bool write_f32(const std::string & path, const std::vector<float> & v) {
    FILE * f = fopen(path.c_str(), "wb");
    if (!f) return false;
    const bool ok = fwrite(v.data(), sizeof(float), v.size(), f) == v.size();
    fclose(f);
    return ok;
}

static std::vector<int> lcg_codes(int n, unsigned seed) {
    std::vector<int> c(n);
    unsigned s = seed;
    for (int & x : c) { s = s * 1664525u + 1013904223u; x = (int) ((s >> 8) % 12800u); }
    return c;
}

static std::string token_text(const std::vector<int> & codes) {
    std::string t = "noise ";
    for (int c : codes) t += "<|s_" + std::to_string(c) + "|>";
    return t + " tail";
}

#define CHECK(cond) do { if (!(cond)) { fprintf(stderr, "CHECK failed line %d: %s\n", __LINE__, #cond); return 1; } } while (0)

int main(int argc, char ** argv) {
    if (argc < 5) { fprintf(stderr, "usage: %s model.gguf voice0 voice1 out_prefix\n", argv[0]); return 2; }
    const std::string prefix = argv[4];
    BatchTestToSpeech tts(argv[1], getenv("MIO_DEVICE") ? atoi(getenv("MIO_DEVICE")) : 0);
    CHECK(tts.is_ready());
    CHECK(tts.sample_rate() == 44100);
    const std::vector<float> v0 = load_voice_embedding(argv[2]), v1 = load_voice_embedding(argv[3]);
    CHECK(v0.size() == 128 && v1.size() == 128);

    const std::vector<int> c0 = lcg_codes(20, 1), c1 = lcg_codes(33, 2);
    const std::vector<std::vector<float>> voices = {v0, v1, v0, {}};
    const std::vector<std::string> texts = {token_text(c0), token_text(c1), "no tokens here <|s_|>", token_text(c0)};

    // ---- to vectors: utterances 2 (nothing parses) and 3 (voice not ready) fail, 0 and 1 still run
    std::vector<std::vector<float>> audio;
    std::vector<bool> ok;
    BatchTestToSpeech::Options raw;
    raw.apply_peak_normalization = false;
    CHECK(!tts.synthesize_to_vectors(voices, texts, audio, ok, raw));
    CHECK(ok.size() == 4 && ok[0] && ok[1] && !ok[2] && !ok[3]);
    CHECK(audio[0].size() == 20u * 1764u && audio[1].size() == 33u * 1764u && audio[2].empty() && audio[3].empty());
    CHECK(write_f32(prefix + ".v0.f32", audio[0]) && write_f32(prefix + ".v1.f32", audio[1]));

    // ---- to files (peak normalisation on, the reference default)
    const std::vector<std::vector<float>> voices2 = {v0, v1};
    const std::vector<std::string> texts2 = {texts[0], texts[1]};
    CHECK(tts.synthesize_to_files(voices2, texts2, {prefix + ".0.wav", prefix + ".1.wav"}, ok, BatchTestToSpeech::Options()));
    CHECK(ok[0] && ok[1]);

    // ---- skip_llm streams: chunks of at most 4096 samples, exactly one last-chunk flag, samples == the vector form
    std::vector<std::vector<float>> got(2);
    std::vector<int> lasts(2, 0), calls(2, 0);
    bool sizes_ok = true;
    std::vector<BatchTestToSpeech::StreamCallback> cbs;
    for (int i = 0; i < 2; i++)
        cbs.push_back([&, i](const float * s, size_t n, int sr, bool last) {
            if (n == 0 || n > 4096 || sr != 44100 || lasts[i]) sizes_ok = false;
            got[i].insert(got[i].end(), s, s + n);
            lasts[i] += last ? 1 : 0;
            calls[i]++;
            return true;
        });
    CHECK(tts.synthesize_streams(voices2, texts2, cbs, 4096, ok));
    CHECK(sizes_ok && lasts[0] == 1 && lasts[1] == 1);
    CHECK(got[0].size() == audio[0].size() && got[1].size() == audio[1].size());
    CHECK(calls[0] == (int) ((audio[0].size() + 4095) / 4096));
    CHECK(write_f32(prefix + ".s0.f32", got[0]));

    // ---- live streams: three generations in lock-step, one code per tick, one poll per tick
    const int T = 100;
    const std::vector<std::vector<int>> gen = {lcg_codes(T, 11), lcg_codes(T, 12), lcg_codes(T, 13)};
    std::vector<std::vector<float>> live(3);
    std::vector<int> live_last(3, 0);
    std::vector<int> ids;
    for (int i = 0; i < 3; i++)
        ids.push_back(tts.open_stream(i == 1 ? v1 : v0, [&, i](const float * s, size_t n, int, bool last) {
            live[i].insert(live[i].end(), s, s + n);
            live_last[i] += last ? 1 : 0;
            return true;
        }, 4096));
    CHECK(ids[0] > 0 && ids[1] > 0 && ids[2] > 0 && tts.live_streams() == 3);
    int dc = 0; size_t dcodes = 0, emitted = 0;
    for (int t = 0; t < T; t++) {
        for (int i = 0; i < 3; i++) CHECK(tts.push_code(ids[i], gen[i][t]));
        CHECK(tts.poll());
    }
    CHECK(tts.stream_profile(ids[0], dc, dcodes, emitted));
    printf("before flush: decode_calls=%d decoded_codes=%zu emitted_samples=%zu\n", dc, dcodes, emitted);
    for (int i = 0; i < 3; i++) CHECK(tts.finish_stream(ids[i]));
    CHECK(tts.poll());
    CHECK(tts.live_streams() == 0);
    for (int i = 0; i < 3; i++) {
        CHECK(live_last[i] == 1);
        CHECK(live[i].size() == (size_t) T * 1764u);
        CHECK(write_f32(prefix + ".l" + std::to_string(i) + ".f32", live[i]));
    }
    printf("ok\n");
    return 0;
}
