// Drop-in replacement of the reference's src/miocodec.h (espresso3389/MioTTS-llama.cpp, src/miocodec.h:1-43):
// same free functions, same argument meaning, same failure conventions (nullptr / empty vector + a line on
// stderr), implemented by miocodec_shim.cpp on top of the C ABI of libmiocodec_b200.so
// (include/miocodec_b200.h).  Put this directory in front of the reference's src/ on the include path and link
// libmiocodec_shim.so instead of compiling src/miocodec.cpp + src/istft.cpp: test-to-speech.cpp, main.cpp and
// the stream-* examples build unchanged.  All work runs on the B200; there is no CPU fallback.
#pragma once

#include <string>
#include <vector>

struct miocodec_context;   // opaque: owns one mio_ctx (device weights + workspace)

// miocodec_load (src/miocodec.cpp:426-504): nullptr on any failure.  Device ordinal: $MIO_DEVICE (default 0).
miocodec_context * miocodec_load(const std::string & model_path);

// miocodec_free (src/miocodec.cpp:506-512)
void miocodec_free(miocodec_context * ctx);

// miocodec_decode (src/miocodec.cpp:519-810): codes[n_codes] in [0, vocab) and a 128-float voice embedding ->
// complex spectrogram, out[t*(n_fft+2) + 2*f + {0: re, 1: im}], *out_n_frames = 18*n_codes.
// `audio_length` is accepted and ignored exactly as the reference does (it only logs it, :546-552).
// Failure (including an out-of-range code, where the reference aborts inside ggml) returns an empty vector.
std::vector<float> miocodec_decode(miocodec_context * ctx, const int * codes, int n_codes, const float * global_emb,
                                   int audio_length, int * out_n_frames);

// accessors (src/miocodec.cpp:514-517)
int miocodec_sample_rate(const miocodec_context * ctx);
int miocodec_n_fft(const miocodec_context * ctx);
int miocodec_hop_length(const miocodec_context * ctx);
int miocodec_samples_per_token(const miocodec_context * ctx);

// load_voice_embedding (src/miocodec.cpp:816-853): first tensor of the .emb.gguf, F32 only; empty on failure.
std::vector<float> load_voice_embedding(const std::string & path);

// miocodec_print_tensors (src/miocodec.cpp:859-877)
void miocodec_print_tensors(const std::string & path);

// ---- extension (not in the reference): the fused device path codes -> waveform for a batch of utterances,
// the call TestToSpeech would use to turn kernel throughput into product throughput.
// out_audio[b] receives samples_per_token * n_codes[b] samples.  Returns false on failure.
bool miocodec_decode_audio_batch(miocodec_context * ctx, const std::vector<std::vector<int>> & codes,
                                 const std::vector<std::vector<float>> & voice_embs, bool peak_normalize,
                                 std::vector<std::vector<float>> & out_audio);
