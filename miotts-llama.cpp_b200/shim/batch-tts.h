// the batched / multi-stream form of the reference's caller, TestToSpeech
// (src/test-to-speech.h:26-124), for the part of it that lies on this side of the LLM: Options.skip_llm == true,
// i.e. `text` is raw "<|s_N|>" token text (src/test-to-speech.h:52, src/test-to-speech.cpp:201-246, :418-427), and the
// live streaming policy of synthesize_stream_profiled (:430-625) fed with codes as an LLM produces them.
// The reference's calls are one-utterance-synchronous; these take B utterances / streams and run ONE ragged device
// batch per call, which is what turns kernel throughput into product throughput.  Same names, argument meaning,
// callback signature and failure conventions (false + a line on stderr, no exceptions) as the reference.
#pragma once

#include <cstddef>
#include <functional>
#include <memory>
#include <string>
#include <vector>

struct miocodec_context;

class BatchTestToSpeech {
public:
    struct Options {                                  // the fields of TestToSpeech::Options that apply without an LLM
        bool apply_peak_normalization = true;         // src/test-to-speech.h:53
    };
    // TestToSpeech::StreamCallback (src/test-to-speech.h:56-59): return false to stop that stream
    using StreamCallback = std::function<bool(const float * samples, size_t n_samples, int sample_rate, bool is_last_chunk)>;

    // codec_path: miocodec.gguf (TestToSpeech::Config::codec_path); device: CUDA ordinal
    explicit BatchTestToSpeech(const std::string & codec_path, int device = 0);
    ~BatchTestToSpeech();
    BatchTestToSpeech(const BatchTestToSpeech &) = delete;
    BatchTestToSpeech & operator=(const BatchTestToSpeech &) = delete;

    bool is_ready() const;
    int sample_rate() const;

    // B x synthesize_to_vector(voice, token_text, out, {skip_llm = true}) (src/test-to-speech.cpp:201-246) in one device
    // batch.  voices[b] is a 128-float embedding (VoiceModel::embedding()).  ok[b] is that utterance's return value: false
    // when the voice is not ready or no code parses, like the reference; the other utterances still run.
    // Returns true when every utterance succeeded.
    bool synthesize_to_vectors(const std::vector<std::vector<float>> & voices, const std::vector<std::string> & token_texts,
                               std::vector<std::vector<float>> & out_audio, std::vector<bool> & ok, const Options & options);
    bool synthesize_to_vectors(const std::vector<std::vector<float>> & voices, const std::vector<std::string> & token_texts,
                               std::vector<std::vector<float>> & out_audio, std::vector<bool> & ok) {
        return synthesize_to_vectors(voices, token_texts, out_audio, ok, Options());
    }

    // B x synthesize_to_file (src/test-to-speech.cpp:75-99 + wav_write): PCM16 is produced on the device.
    bool synthesize_to_files(const std::vector<std::vector<float>> & voices, const std::vector<std::string> & token_texts,
                             const std::vector<std::string> & output_paths, std::vector<bool> & ok, const Options & options);

    // B x synthesize_stream(voice, token_text, callback, chunk_samples, {skip_llm = true}) (:418-427): decode without peak
    // normalisation, then emit_range(audio, 0, n, true): callbacks of at most chunk_samples samples, the last one
    // flagged.  Callbacks run on the calling thread, utterance by utterance, after the single device batch.
    bool synthesize_streams(const std::vector<std::vector<float>> & voices, const std::vector<std::string> & token_texts,
                            const std::vector<StreamCallback> & callbacks, size_t chunk_samples, std::vector<bool> & ok);

    // ---- live streams: the LLM branch of synthesize_stream_profiled (:430-625) for many concurrent generations.
    // open_stream -> id; push_code as each "<|s_N|>" token is sampled; finish_stream at end of generation; poll()
    // advances ALL streams with one batched re-decode (cadence 20 / holdback 32 / min step 24 / 30 ms crossfade, the
    // reference's constants) and delivers their callbacks in chunks of chunk_samples.  A callback returning false
    // closes its stream (the reference aborts the synthesis, :404-407).
    int  open_stream(const std::vector<float> & voice, const StreamCallback & callback, size_t chunk_samples = 4096);
    bool push_code(int stream_id, int code);
    bool finish_stream(int stream_id);
    bool poll();                                       // false on a device error
    size_t live_streams() const;
    // StreamProfile counters of one stream (src/test-to-speech.h:28-38): decode_calls, decoded_codes, emitted_samples
    bool stream_profile(int stream_id, int & decode_calls, size_t & decoded_codes, size_t & emitted_samples) const;

private:
    struct impl;
    std::unique_ptr<impl> p_;
};
