// Host state of one audio stream (see stream.cu).
#pragma once
#include <deque>
#include <string>
#include <vector>

#include "../../include/miocodec_b200.h"
#include "engine.h"

namespace mio {

constexpr int MIO_TAIL_CAP = 4096;   // crossfade_samples is capped at 4096 by the reference (:364)

struct stream_job {
    size_t prefix = 0;            // codes decoded (the whole prefix at the time the cadence fired)
    size_t commit_begin = 0, commit_end = 0;
    bool is_final = false;
};

struct stream_state {
    engine * eng = nullptr;
    mio_stream_config cfg;
    std::vector<float> emb;
    std::vector<int32_t> codes;
    size_t committed = 0;
    std::vector<stream_job> jobs;
    size_t job_head = 0;
    bool finished = false, final_queued = false, done = false;
    bool failed = false;          // a device pass of this stream failed: its tail on the device is undefined, the stream is over
    float * d_tail = nullptr;
    int * d_tail_len = nullptr;
    // emitted samples not yet read: slices of the engine's pinned round blocks (blk != nullptr, p points into it) or, when
    // the pool is exhausted, an own copy
    struct emit_seg {
        engine::emit_block * blk = nullptr;
        std::vector<float> own;
        const float * p = nullptr;
        int64_t n = 0, used = 0;
    };
    std::deque<emit_seg> segs;
    int64_t decode_calls = 0, decoded_codes = 0, emitted_samples = 0;
    ~stream_state();
};

void stream_default_config(const engine & E, mio_stream_config * cfg);
std::string stream_open(engine & E, const float * emb, const mio_stream_config * cfg, stream_state ** out);
std::string stream_push(stream_state & s, const int32_t * codes, int n);
std::string stream_finish(stream_state & s);
std::string streams_step(engine & E, stream_state ** streams, int n_streams);
std::string streams_step_begin(engine & E, stream_state ** streams, int n_streams, bool * in_flight = nullptr);
std::string streams_step_end(engine & E, stream_state ** streams, int n_streams);
std::string streams_drain(engine & E);
void streams_detach_all(engine & E);   // before the engine is destroyed: open streams keep their unread samples, lose the engine
int64_t stream_read(stream_state & s, float * out, int64_t cap, int * is_last);

}  // namespace mio
