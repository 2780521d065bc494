// Streaming policy of TestToSpeech::synthesize_stream_profiled (src/test-to-speech.cpp:348-626 of the reference),
// re-implemented for many concurrent streams:
//   * cadence / holdback / min-commit decisions are integer logic on the host (maybe_emit, :507-571)
//   * every stream whose cadence fired is re-decoded over its full prefix in ONE ragged device batch
//   * the crossfade with the previously emitted tail, the tail update and the cut [begin, end) of the window
//     (emit_range, :367-417) run in a device kernel, so only the newly committed samples cross PCIe
#include "stream.h"
#include "ptx.cuh"

#include <chrono>

#include <algorithm>
#include <cmath>
#include <cstring>

namespace mio {

struct emit_job {                 // one firing stream in the current round
    int64_t wave_off;             // window audio of this job inside engine::d_wave
    int64_t begin, end;           // sample range of the window to emit
    int64_t emit_off;             // offset in the round's emit buffer
    float * tail;                 // device, capacity MIO_TAIL_CAP
    int * tail_len;               // device
    int chunk, xfade;
};

// one block per job. out[j] = window[begin+j], the first chunk cross-faded with the stored tail:
//   a = (j+1)/(xf+1), out = (1-a)*tail[j] + a*x   (products and sum rounded separately like the reference's
//   scalar loop, :388-395); new tail = last `xfade` samples of the last chunk, or the whole last chunk if shorter.
__global__ void stream_emit_kernel(const float * __restrict__ wave, float * __restrict__ emit, const emit_job * __restrict__ jobs) {
    pdl_sync();
    const emit_job J = jobs[blockIdx.x];
    const int64_t n = J.end - J.begin;
    if (n <= 0) return;
    const float * src = wave + J.wave_off + J.begin;
    float * dst = emit + J.emit_off;
    const int tl = *J.tail_len;
    const int64_t first_n = min((int64_t) J.chunk, n);
    const int xf = (int) min((int64_t) tl, first_n);
    for (int64_t j = threadIdx.x; j < n; j += blockDim.x) {
        float v = src[j];
        if (j < xf) {
            const float a = (float) (j + 1) / (float) (xf + 1);
            const float b = 1.0f - a;
            v = __fadd_rn(__fmul_rn(b, J.tail[j]), __fmul_rn(a, v));
        }
        dst[j] = v;
    }
    __syncthreads();   // (global writes of this block are visible to the block after the barrier)
    const int64_t last0 = ((n - 1) / J.chunk) * J.chunk;
    const int64_t last_n = n - last0;
    const int new_len = (int) (last_n >= J.xfade ? J.xfade : last_n);
    const int64_t t0 = n - new_len;
    __threadfence_block();
    for (int j = threadIdx.x; j < new_len; j += blockDim.x) J.tail[j] = dst[t0 + j];
    if (threadIdx.x == 0) *J.tail_len = new_len;
}

void stream_default_config(const engine & E, mio_stream_config * cfg) {
    cfg->check_interval = 20;
    cfg->holdback_codes = 32;
    cfg->min_commit_codes = 24;
    cfg->crossfade_samples = std::min(E.hp.sample_rate * 3 / 100, 4096);
    cfg->chunk_samples = 4096;
}

stream_state::~stream_state() {
    // eng == nullptr: the context went first (streams_detach_all) and the segments already own their samples
    for (emit_seg & g : segs) if (g.blk && eng) g.blk->refs--;
    if (eng) {
        auto & v = eng->open_streams;
        v.erase(std::remove(v.begin(), v.end(), (void *) this), v.end());
    }
    if (d_tail) cudaFree(d_tail);
    if (d_tail_len) cudaFree(d_tail_len);
}

std::string stream_open(engine & E, const float * emb, const mio_stream_config * cfg, stream_state ** out) {
    stream_state * s = new stream_state();
    s->eng = &E;
    stream_default_config(E, &s->cfg);
    if (cfg) {
        s->cfg = *cfg;
        if (s->cfg.crossfade_samples < 0) s->cfg.crossfade_samples = std::min(E.hp.sample_rate * 3 / 100, 4096);
        if (s->cfg.chunk_samples <= 0) s->cfg.chunk_samples = 4096;     // :360-362
    }
    if (s->cfg.check_interval <= 0 || s->cfg.holdback_codes < 0 || s->cfg.min_commit_codes < 0 ||
        s->cfg.crossfade_samples > MIO_TAIL_CAP) { delete s; return "bad stream configuration"; }
    s->emb.assign(emb, emb + E.hp.adaln_dim);
    cudaSetDevice(E.device);
    if (cudaMalloc(&s->d_tail, MIO_TAIL_CAP * sizeof(float)) != cudaSuccess || cudaMalloc(&s->d_tail_len, sizeof(int)) != cudaSuccess ||
        cudaMemset(s->d_tail_len, 0, sizeof(int)) != cudaSuccess) { delete s; return "cudaMalloc failed (stream tail)"; }
    E.open_streams.push_back(s);
    *out = s;
    return {};
}

// mio_ctx_destroy with streams still open (e.g. a Python caller closing the codec before its Stream objects are collected):
// the engine's pinned emit blocks are about to be freed, so every unread slice becomes the stream's own copy and the
// stream forgets the engine.  Such a stream can still be read, queried and closed; stepping it needs a context and fails
// in the caller's hands, not here.
void streams_detach_all(engine & E) {
    for (void * v : E.open_streams) {
        stream_state & s = *(stream_state *) v;
        for (stream_state::emit_seg & g : s.segs) {
            if (!g.blk) continue;
            g.own.assign(g.p, g.p + g.n);
            g.p = g.own.data();
            g.blk->refs--;
            g.blk = nullptr;
        }
        s.eng = nullptr;
    }
    E.open_streams.clear();
}

// maybe_emit(is_final) evaluated with the prefix length `n` (src/test-to-speech.cpp:507-571)
static void maybe_emit(stream_state & s, size_t n, bool is_final) {
    if (n == 0) { if (is_final) s.final_queued = true; return; }
    const size_t hold = (size_t) s.cfg.holdback_codes;
    const size_t target = is_final ? n : (n > hold ? n - hold : 0);
    if (target <= s.committed) { if (is_final) s.final_queued = true; return; }
    if (!is_final && (target - s.committed) < (size_t) s.cfg.min_commit_codes) return;
    stream_job j;
    j.prefix = n; j.commit_begin = s.committed; j.commit_end = target; j.is_final = is_final;
    s.jobs.push_back(j);
    s.committed = target;          // the reference commits right after a successful decode
    if (is_final) s.final_queued = true;
}

std::string stream_push(stream_state & s, const int32_t * codes, int n) {
    if (s.failed) return "stream failed in an earlier round (close it)";
    if (s.finished) return "stream already finished";
    for (int i = 0; i < n; i++) {
        s.codes.push_back(codes[i]);
        if ((s.codes.size() % (size_t) s.cfg.check_interval) == 0) maybe_emit(s, s.codes.size(), false);   // :608-613
    }
    return {};
}

std::string stream_finish(stream_state & s) {
    if (s.finished) return {};
    s.finished = true;
    maybe_emit(s, s.codes.size(), true);   // :616-618
    s.final_queued = true;
    return {};
}

// One pass = the oldest pending job of up to max_voices streams as ONE ragged device batch.  pass_begin enqueues the whole
// pass (upload, decode, cut + crossfade, the D2H copy of the emitted samples into a pinned block) and returns without
// waiting; pass_end waits and hands the samples to the streams.  Between the two the host is free (the LLM produces the
// next tokens, the consumer reads what earlier rounds emitted) while the GPU works.
static std::string pass_begin(engine & E, stream_state ** streams, int n_streams, bool * any) {
    *any = false;
    engine::stream_pass & P = E.pass;
    if (P.active) return "a streaming round is already in flight (mio_streams_step_end first)";
    const int spc = E.hp.samples_per_token;
    // the oldest pending job of every stream (jobs of one stream are ordered through its tail)
    std::vector<stream_state *> act;
    for (int i = 0; i < n_streams; i++)
        if (streams[i] && streams[i]->job_head < streams[i]->jobs.size()) act.push_back(streams[i]);
    if (act.empty()) return {};
    P.t0 = std::chrono::steady_clock::now();
    auto since = [&](const std::chrono::steady_clock::time_point & t) { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t).count(); };
    // a pass holds at most max_voices streams (one conditioning row each); the rest fire in the next pass of the same
    // step call, so nothing is dropped
    if ((int) act.size() > E.max_voices) act.resize(E.max_voices);
    std::vector<int32_t> codes, cu(1, 0), vid;
    std::vector<float> vt;
    for (size_t k = 0; k < act.size(); k++) {
        const stream_job & j = act[k]->jobs[act[k]->job_head];
        codes.insert(codes.end(), act[k]->codes.begin(), act[k]->codes.begin() + j.prefix);
        cu.push_back((int32_t) codes.size());
        vid.push_back((int32_t) k);
        vt.insert(vt.end(), act[k]->emb.begin(), act[k]->emb.end());
    }
    P.ms_prep = since(P.t0);
    auto t1 = std::chrono::steady_clock::now();
    std::string e = E.upload(codes.data(), cu.data(), (int) act.size(), vt.data(), (int) act.size(), vid.data());
    if (!e.empty()) return e;
    P.ms_upload = since(t1); t1 = std::chrono::steady_clock::now();
    e = E.run(false);      // streaming never peak-normalises (decode_codes_to_audio(..., false), :546)
    if (!e.empty()) return e;
    // from here on GPU work is queued: an error below waits for it before returning, so that the caller is never handed an
    // error with kernels still running on buffers it may free next; the streams' counters are booked in pass_end only
    auto bail = [&](const char * msg) { (void) E.sync(); return std::string(msg); };
    P.ms_run = since(t1); t1 = std::chrono::steady_clock::now();
    // cut + crossfade on the device
    std::vector<emit_job> jobs(act.size());
    P.begin.assign(act.size(), 0); P.end.assign(act.size(), 0); P.off.assign(act.size(), 0); P.prefix.assign(act.size(), 0);
    int64_t emit_total = 0;
    for (size_t k = 0; k < act.size(); k++) {
        stream_state & s = *act[k];
        const stream_job & j = s.jobs[s.job_head];
        const int64_t win = (int64_t) j.prefix * spc;                       // window_audio.size()
        const double spc_actual = (double) win / (double) j.prefix;          // :554-556
        const int64_t b = (int64_t) std::llround((double) j.commit_begin * spc_actual);
        const int64_t en = std::min<int64_t>((int64_t) std::llround((double) j.commit_end * spc_actual), win);
        emit_job & J = jobs[k];
        J.wave_off = (int64_t) cu[k] * spc; J.begin = b; J.end = std::max(b, en); J.emit_off = emit_total;
        J.tail = s.d_tail; J.tail_len = s.d_tail_len; J.chunk = s.cfg.chunk_samples; J.xfade = s.cfg.crossfade_samples;
        P.begin[k] = J.begin; P.end[k] = J.end; P.off[k] = J.emit_off;
        emit_total += J.end - J.begin;
        P.prefix[k] = (int64_t) j.prefix;
    }
    // the two buffers grow independently and never shrink (a reservation made by mio_ctx_reserve survives every round)
    static_assert(sizeof(emit_job) <= engine::EMIT_JOB_BYTES, "engine::reserve sizes the job table with EMIT_JOB_BYTES per entry");
    if (E.emit_cap < emit_total) {
        if (E.d_emit) cudaFree(E.d_emit);
        E.d_emit = nullptr; E.emit_cap = 0;          // a failed allocation below must not leave a freed pointer behind
        const int64_t want_emit = emit_total + emit_total / 4 + 4096;
        if (cudaMalloc(&E.d_emit, (size_t) want_emit * sizeof(float)) != cudaSuccess) { E.d_emit = nullptr; return bail("cudaMalloc failed (emit buffer)"); }
        E.emit_cap = want_emit;
    }
    if (E.emit_jobs_cap < (int64_t) jobs.size()) {
        if (E.d_emit_jobs) cudaFree(E.d_emit_jobs);
        E.d_emit_jobs = nullptr; E.emit_jobs_cap = 0;
        const int64_t want_jobs = (int64_t) jobs.size() + 64;
        if (cudaMalloc(&E.d_emit_jobs, (size_t) want_jobs * engine::EMIT_JOB_BYTES) != cudaSuccess) { E.d_emit_jobs = nullptr; return bail("cudaMalloc failed (emit job table)"); }
        E.emit_jobs_cap = want_jobs;
    }
    bool pooled = true;
    engine::emit_block * blk = emit_total > 0 ? E.acquire_emit_block(emit_total, &pooled) : nullptr;
    if (emit_total > 0 && !blk) return bail("cudaMallocHost failed (emit block)");
    // the job table is copied from a buffer that outlives this function (the copy is asynchronous only for pinned memory,
    // but the pass state keeps it alive either way)
    P.jobs_host.assign((const uint8_t *) jobs.data(), (const uint8_t *) jobs.data() + jobs.size() * sizeof(emit_job));
    if (cudaMemcpyAsync(E.d_emit_jobs, P.jobs_host.data(), P.jobs_host.size(), cudaMemcpyHostToDevice, E.stream) != cudaSuccess)
        return bail("cudaMemcpyAsync failed (emit jobs)");
    launch_k(stream_emit_kernel, dim3((unsigned) jobs.size()), dim3(256), 0, E.stream, E.d_wave, (float *) E.d_emit, (const emit_job *) E.d_emit_jobs);
    E.launches++;
    if (cudaGetLastError() != cudaSuccess) return bail("stream_emit_kernel launch failed");
    if (emit_total > 0 &&
        cudaMemcpyAsync(blk->host, E.d_emit, (size_t) emit_total * sizeof(float), cudaMemcpyDeviceToHost, E.stream) != cudaSuccess)
        return bail("cudaMemcpyAsync failed (emit D2H)");
    P.ms_emit = since(t1);
    P.act.assign(act.begin(), act.end());
    P.blk = blk; P.pooled = pooled; P.n_codes = (int64_t) codes.size();
    if (blk) blk->refs++;                             // held by the pass until pass_end has handed out the slices
    P.active = true;
    *any = true;
    return {};
}

static std::string pass_end(engine & E) {
    engine::stream_pass & P = E.pass;
    if (!P.active) return {};
    static const bool timing = getenv("MIO_STREAM_TIMING") != nullptr;      // per-round host timeline on stderr
    const auto t1 = std::chrono::steady_clock::now();
    std::string e = E.sync();
    P.active = false;
    engine::emit_block * blk = P.blk;
    if (blk) blk->refs--;
    if (!e.empty()) {
        // the emit kernel may or may not have rewritten these streams' device tails: a retry of the same job would crossfade
        // against an undefined tail, so the streams of the pass end here (readable, closable, no longer decodable)
        for (void * v : P.act) {
            stream_state & s = *(stream_state *) v;
            s.failed = true; s.done = true; s.final_queued = true;
            s.job_head = s.jobs.size();
        }
        P.act.clear();
        return e;
    }
    const double ms_wait = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count();
    if (timing) fprintf(stderr, "mio stream round: %zu streams, %lld codes: prep %.2f upload %.2f enqueue %.2f emit-setup %.2f wait %.2f ms\n",
                        P.act.size(), (long long) P.n_codes, P.ms_prep, P.ms_upload, P.ms_run, P.ms_emit, ms_wait);
    for (size_t k = 0; k < P.act.size(); k++) {
        stream_state & s = *(stream_state *) P.act[k];
        const int64_t n = P.end[k] - P.begin[k];
        if (n > 0) {
            const float * src = blk->host + P.off[k];
            s.segs.emplace_back();
            stream_state::emit_seg & g = s.segs.back();
            g.n = n;
            if (P.pooled) { g.blk = blk; g.p = src; blk->refs++; }       // no second host copy: read straight from the pinned block
            else { g.own.assign(src, src + n); g.p = g.own.data(); }
        }
        s.emitted_samples += n;
        s.decode_calls++; s.decoded_codes += P.prefix[k];
        s.job_head++;
    }
    E.stream_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - P.t0).count();
    E.stream_passes++;
    P.act.clear();
    return {};
}

static void mark_done(stream_state ** streams, int n_streams) {
    for (int i = 0; i < n_streams; i++) {
        stream_state * s = streams[i];
        if (s && s->final_queued && s->job_head >= s->jobs.size()) s->done = true;
    }
}

std::string streams_step(engine & E, stream_state ** streams, int n_streams) {
    cudaSetDevice(E.device);
    std::string e = pass_end(E);                      // a round left in flight by streams_step_begin
    if (!e.empty()) return e;
    for (;;) {
        bool any = false;
        if (!(e = pass_begin(E, streams, n_streams, &any)).empty()) return e;
        if (!any) break;
        if (!(e = pass_end(E)).empty()) return e;
    }
    mark_done(streams, n_streams);
    return {};
}

// The asynchronous form: begin enqueues the first pass and returns while the GPU works; end waits for it and runs whatever
// further passes the streams still need (more than max_voices streams firing at once).  The streams of a pass in flight
// may be pushed to and read from, not closed.
std::string streams_step_begin(engine & E, stream_state ** streams, int n_streams, bool * in_flight) {
    cudaSetDevice(E.device);
    bool any = false;
    std::string e = pass_begin(E, streams, n_streams, &any);
    if (in_flight) *in_flight = any;
    return e;
}

std::string streams_step_end(engine & E, stream_state ** streams, int n_streams) {
    return streams_step(E, streams, n_streams);
}

// waits for a pass in flight and hands out its samples (before a stream of that pass is closed)
std::string streams_drain(engine & E) { cudaSetDevice(E.device); return pass_end(E); }

int64_t stream_read(stream_state & s, float * out, int64_t cap, int * is_last) {
    int64_t done = 0;
    while (done < cap && !s.segs.empty()) {
        stream_state::emit_seg & g = s.segs.front();
        const int64_t n = std::min(cap - done, g.n - g.used);
        if (n > 0) memcpy(out + done, g.p + g.used, (size_t) n * sizeof(float));
        g.used += n; done += n;
        if (g.used == g.n) {
            if (g.blk) g.blk->refs--;
            s.segs.pop_front();
        }
    }
    if (is_last) *is_last = (s.done && s.segs.empty()) ? 1 : 0;
    return done;
}

}  // namespace mio
