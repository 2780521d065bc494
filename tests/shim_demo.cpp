// Test program for the C++ drop-in layer (shim/miocodec.h, shim/istft.h): does what the reference's
// decode_codes_to_audio does (src/test-to-speech.cpp:248-303) through the reference-named entry points and
// writes the results as raw little-endian f32 so tests/test_shim.py can compare them with the golden vectors.
//   shim_demo <miocodec.gguf> <voice.emb.gguf> <codes.i32> <out_prefix>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <string>
#include <vector>

#include "istft.h"
#include "miocodec.h"

static bool write_f32(const std::string & path, const std::vector<float> & v) {
    FILE * f = fopen(path.c_str(), "wb");
    if (!f) return false;
    const bool ok = fwrite(v.data(), sizeof(float), v.size(), f) == v.size();
    fclose(f);
    return ok;
}

int main(int argc, char ** argv) {
    // shim_demo --dump-tables <n_fft> <win_length> <out_prefix>: the istft_cache tables (host-side, needs no GPU), for the
    // value-by-value comparison with the reference's istft_cache (tests/test_shim.py)
    if (argc == 5 && std::string(argv[1]) == "--dump-tables") {
        istft_cache c(atoi(argv[2]), atoi(argv[3]));
        const std::string pre = argv[4];
        printf("n_fft=%d win_length=%d n_freq=%d n_mid=%d\n", c.n_fft(), c.win_length(), c.n_freq(), c.n_mid());
        return (write_f32(pre + ".cos.f32", c.cos_table()) && write_f32(pre + ".sin.f32", c.sin_table()) &&
                write_f32(pre + ".nyq.f32", c.nyquist_sign()) && write_f32(pre + ".hann.f32", c.hann_window())) ? 0 : 1;
    }
    if (argc < 5) { fprintf(stderr, "usage: %s model.gguf voice.emb.gguf codes.i32 out_prefix\n", argv[0]); return 2; }
    const std::string prefix = argv[4];
    // istft() must work before any codec is loaded (it is a free function in the reference)
    auto cache = std::make_unique<istft_cache>(392, 392);
    {
        std::vector<float> spec(50 * 394);
        unsigned s = 12345u;
        for (float & x : spec) { s = s * 1664525u + 1013904223u; x = (float) ((s >> 8) & 0xffff) / 32768.0f - 1.0f; }
        std::vector<float> w = istft(spec.data(), 50, 98, *cache);
        if (w.size() != 50u * 98u) { fprintf(stderr, "standalone istft failed (%zu samples)\n", w.size()); return 1; }
        if (!write_f32(prefix + ".istft_in.f32", spec) || !write_f32(prefix + ".istft_out.f32", w)) return 1;
    }
    miocodec_context * codec = miocodec_load(argv[1]);
    if (!codec) return 1;
    std::vector<float> emb = load_voice_embedding(argv[2]);
    if (emb.size() != 128) { fprintf(stderr, "bad voice embedding\n"); return 1; }
    std::vector<int> codes;
    {
        FILE * f = fopen(argv[3], "rb");
        if (!f) { fprintf(stderr, "cannot open %s\n", argv[3]); return 1; }
        int v;
        while (fread(&v, sizeof(int), 1, f) == 1) codes.push_back(v);
        fclose(f);
    }
    const int spt = miocodec_samples_per_token(codec), hop = miocodec_hop_length(codec);
    int n_frames = 0;
    std::vector<float> spec = miocodec_decode(codec, codes.data(), (int) codes.size(), emb.data(), (int) codes.size() * spt, &n_frames);
    if (spec.empty()) { fprintf(stderr, "codec decode failed\n"); return 1; }
    std::vector<float> audio = istft(spec.data(), n_frames, hop, *cache);
    if (audio.empty()) { fprintf(stderr, "iSTFT failed\n"); return 1; }
    printf("frames=%d samples=%zu sample_rate=%d n_fft=%d hop=%d spt=%d\n", n_frames, audio.size(), miocodec_sample_rate(codec),
           miocodec_n_fft(codec), hop, spt);
    if (!write_f32(prefix + ".spec.f32", spec) || !write_f32(prefix + ".wave.f32", audio)) return 1;
    // failure conventions: out-of-range code -> empty vector, context stays usable
    std::vector<int> bad = codes;
    bad[0] = 1 << 20;
    if (!miocodec_decode(codec, bad.data(), (int) bad.size(), emb.data(), 0, &n_frames).empty()) { fprintf(stderr, "bad code accepted\n"); return 1; }
    // batch extension
    std::vector<std::vector<float>> out;
    if (!miocodec_decode_audio_batch(codec, {codes, std::vector<int>(codes.begin(), codes.begin() + 1)}, {emb, emb}, false, out)) return 1;
    if (out.size() != 2 || out[0].size() != audio.size() || out[1].size() != (size_t) spt) { fprintf(stderr, "batch sizes wrong\n"); return 1; }
    if (!write_f32(prefix + ".batch0.f32", out[0])) return 1;
    miocodec_free(codec);
    if (miocodec_load("/nonexistent/model.gguf") != nullptr) return 1;
    if (!load_voice_embedding("/nonexistent/voice.emb.gguf").empty()) return 1;
    printf("ok\n");
    return 0;
}
