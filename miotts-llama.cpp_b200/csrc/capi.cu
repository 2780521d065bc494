// extern "C" boundary (include/miocodec_b200.h).  No exceptions cross it; errors are codes + mio_last_error().
#include "../../include/miocodec_b200.h"

#include <cstdio>
#include <cstring>
#include <exception>
#include <new>
#include <string>

#include "engine.h"
#include "stream.h"

using namespace mio;

namespace mio {   // csrc/noalloc.cu
std::string noalloc_peak_normalize(int device, const float * in, float * out, const int64_t * offs, int n_utts, cudaStream_t st);
}

struct mio_ctx { engine E; };
struct mio_stream { stream_state * S = nullptr; };

static thread_local std::string g_err;

static int fail(int code, const std::string & msg) {
    g_err = msg;
    return code;
}
static int classify(const std::string & e) {
    if (e.rfind("RANGE:", 0) == 0) return MIO_ERR_RANGE;
    if (e.find("cannot open") != std::string::npos) return MIO_ERR_IO;
    if (e.find("GGUF") != std::string::npos || e.find("tensor") != std::string::npos || e.find("unsupported") != std::string::npos)
        return MIO_ERR_FORMAT;
    if (e.find("bad argument") != std::string::npos || e.find("capacity") != std::string::npos || e.find("out of range") != std::string::npos ||
        e.find("exceeds") != std::string::npos || e.find("too many") != std::string::npos)
        return MIO_ERR_ARG;
    if (e.find("before") != std::string::npos) return MIO_ERR_STATE;
    return MIO_ERR_CUDA;
}
// No exception may cross the C ABI: host allocations (std::vector growth on a huge batch, std::string) can throw.
template <class F>
static int guarded(F && body) {
    try {
        return body();
    } catch (const std::bad_alloc &) {
        return fail(MIO_ERR_ARG, "out of host memory");
    } catch (const std::exception & e) {
        return fail(MIO_ERR_ARG, std::string("internal error: ") + e.what());
    } catch (...) {
        return fail(MIO_ERR_ARG, "internal error");
    }
}
#define TRY(expr)                                          \
    do {                                                   \
        std::string _e = (expr);                           \
        if (!_e.empty()) return fail(classify(_e), _e);    \
    } while (0)

extern "C" {

const char * mio_last_error(void) { return g_err.c_str(); }
const char * mio_version(void) { return "miocodec_b200 0.1 (sm_100a)"; }

int mio_ctx_create(const char * gguf_path, int device, mio_ctx ** out_ctx) {
    return guarded([&]() -> int {
        if (!gguf_path || !out_ctx) return fail(MIO_ERR_ARG, "bad argument");
        *out_ctx = nullptr;
        {   // file problems are reported before any CUDA call so the error is meaningful on machines without a GPU
            gguf_file probe;
            std::string e = probe.open(gguf_path);
            if (!e.empty()) return fail(classify(e), e);
        }
        mio_ctx * c = new (std::nothrow) mio_ctx();
        if (!c) return fail(MIO_ERR_ARG, "out of host memory");
        std::string e = c->E.load(gguf_path, device);
        if (!e.empty()) {
            delete c;
            return fail(classify(e), e);
        }
        if (getenv("MIO_GEMM") && !strcmp(getenv("MIO_GEMM"), "simt")) c->E.use_simt = true;
        *out_ctx = c;
        return MIO_OK;
    });
}

int mio_istft_ctx_create(int device, mio_ctx ** out_ctx) {
    return guarded([&]() -> int {
        if (!out_ctx) return fail(MIO_ERR_ARG, "bad argument");
        *out_ctx = nullptr;
        mio_ctx * c = new (std::nothrow) mio_ctx();
        if (!c) return fail(MIO_ERR_ARG, "out of host memory");
        std::string e = c->E.init_istft_only(device);
        if (!e.empty()) {
            delete c;
            return fail(classify(e), e);
        }
        *out_ctx = c;
        return MIO_OK;
    });
}

void mio_ctx_destroy(mio_ctx * ctx) {
    if (!ctx) return;
    if (ctx->E.pass.active) (void) streams_drain(ctx->E);
    streams_detach_all(ctx->E);      // streams that outlive their context keep their unread samples and stay closable
    delete ctx;
}

int mio_ctx_sample_rate(const mio_ctx * c) { return c ? c->E.hp.sample_rate : 0; }
int mio_ctx_n_fft(const mio_ctx * c) { return c ? c->E.hp.n_fft : 0; }
int mio_ctx_hop_length(const mio_ctx * c) { return c ? c->E.hp.hop : 0; }
int mio_ctx_samples_per_token(const mio_ctx * c) { return c ? c->E.hp.samples_per_token : 0; }
int mio_ctx_frames_per_code(const mio_ctx * c) { return c ? c->E.hp.frames_per_code : 0; }
int mio_ctx_vocab(const mio_ctx * c) { return c ? c->E.hp.vocab : 0; }
int mio_ctx_device(const mio_ctx * c) { return c ? c->E.device : -1; }

int mio_ctx_set_option(mio_ctx * c, const char * key, const char * value) {
    if (!c || !key || !value) return fail(MIO_ERR_ARG, "bad argument");
    if (!strcmp(key, "gemm")) {
        if (!strcmp(value, "tc") || !strcmp(value, "tc3")) c->E.use_simt = false;
        else if (!strcmp(value, "simt")) c->E.use_simt = true;
        else return fail(MIO_ERR_ARG, "bad argument: gemm must be tc or simt");
        return MIO_OK;
    }
    if (!strcmp(key, "chunk_acc")) {
        if (!strcmp(value, "off")) c->E.chunk_mode = 0;
        else if (!strcmp(value, "prenet")) c->E.chunk_mode = 1;
        else if (!strcmp(value, "all")) c->E.chunk_mode = 2;
        else return fail(MIO_ERR_ARG, "bad argument: chunk_acc must be off, prenet or all");
        return MIO_OK;
    }
    if (!strcmp(key, "fuse_tail")) {
        c->E.fuse_resblock_tail = strcmp(value, "0") != 0 && strcmp(value, "off") != 0;
        return MIO_OK;
    }
    if (!strcmp(key, "max_chunk_codes")) {
        if (c->E.ws_ready) return fail(MIO_ERR_STATE, "max_chunk_codes must be set before the first decode");
        const long v = strtol(value, nullptr, 10);
        if (v < 64 || v > (1 << 20)) return fail(MIO_ERR_ARG, "bad argument: max_chunk_codes out of range");
        c->E.max_rows1 = (int) v + (int) v / 32 + 64;
        return MIO_OK;
    }
    if (!strcmp(key, "fetch_chunk_codes")) {
        const long v = strtol(value, nullptr, 10);
        if (v < 64 || v > (1 << 20)) return fail(MIO_ERR_ARG, "bad argument: fetch_chunk_codes out of range");
        c->E.fetch_chunk_codes = (int) v;
        return MIO_OK;
    }
    if (!strcmp(key, "fetch_tail_percent")) {
        const long v = strtol(value, nullptr, 10);
        if (v < -1 || v > 50) return fail(MIO_ERR_ARG, "bad argument: fetch_tail_percent out of range");
        c->E.fetch_tail_percent = (int) v;
        return MIO_OK;
    }
    if (!strcmp(key, "debug_snapshot_filter")) {   // comma-separated tap names; empty = all
        c->E.snap_filter = value[0] ? std::string(",") + value + "," : std::string();
        return MIO_OK;
    }
    if (!strcmp(key, "debug_snapshots")) {
        c->E.debug_snap = strcmp(value, "0") != 0;
        return MIO_OK;
    }
    return fail(MIO_ERR_ARG, std::string("bad argument: unknown option ") + key);
}

int mio_ctx_set_stream(mio_ctx * c, void * cuda_stream) {
    if (!c) return fail(MIO_ERR_ARG, "bad argument");
    cudaSetDevice(c->E.device);
    if (c->E.stream && c->E.own_stream) {
        cudaStreamSynchronize(c->E.stream);
        cudaStreamDestroy(c->E.stream);
    }
    c->E.stream = (cudaStream_t) cuda_stream;
    c->E.own_stream = false;
    return MIO_OK;
}

int mio_voice_load(const char * path, float * out, int cap) {
    return guarded([&]() -> int {
        if (!path || !out) return fail(MIO_ERR_ARG, "bad argument");
        gguf_file g;
        std::string e = g.open(path);
        if (!e.empty()) return fail(classify(e), e);
        if (g.tensors.empty()) return fail(MIO_ERR_FORMAT, std::string("no tensors in ") + path);
        const gguf_tensor & t = g.tensors[0];               // "the first tensor", src/miocodec.cpp:825
        if (t.dtype != GT_F32) return fail(MIO_ERR_FORMAT, "unsupported voice embedding type (F32 only)");   // :840-847
        const int64_t n = t.nelements();
        if (n > cap) return fail(MIO_ERR_ARG, "capacity too small");
        if (!g.read_raw(t, out, (size_t) n * sizeof(float))) return fail(MIO_ERR_IO, "short read");
        return (int) n;
    });
}

int mio_print_tensors(const char * path) {
    return guarded([&]() -> int {
        if (!path) return fail(MIO_ERR_ARG, "bad argument");
        gguf_file g;
        std::string e = g.open(path);
        if (!e.empty()) return fail(classify(e), e);
        printf("Tensors in %s: %lld\n", path, (long long) g.tensors.size());
        for (const gguf_tensor & t : g.tensors)
            printf("  %-60s [%5lld, %5lld, %5lld, %5lld] type=%d\n", t.name.c_str(), (long long) t.ne[0], (long long) t.ne[1],
                   (long long) t.ne[2], (long long) t.ne[3], (int) t.dtype);
        return MIO_OK;
    });
}

static int upload_one(mio_ctx * c, const int32_t * codes, int n_codes, const float * emb) {
    if (!c || !codes || !emb || n_codes <= 0) return fail(MIO_ERR_ARG, "bad argument");
    const int32_t cu[2] = {0, n_codes};
    const int32_t vid = 0;
    TRY(c->E.upload(codes, cu, 1, emb, 1, &vid));
    return MIO_OK;
}

int mio_decode_spec(mio_ctx * c, const int32_t * codes, int n_codes, const float * emb, float * spec_out, int64_t cap,
                    int * out_n_frames) {
    return guarded([&]() -> int {
        if (!c) return fail(MIO_ERR_ARG, "bad argument");
        const int nf = n_codes * c->E.hp.frames_per_code;
        if (out_n_frames) *out_n_frames = nf;
        if (!spec_out) return MIO_OK;
        if (cap < (int64_t) nf * 2 * c->E.hp.n_freq) return fail(MIO_ERR_ARG, "capacity too small");
        int r = upload_one(c, codes, n_codes, emb);
        if (r != MIO_OK) return r;
        TRY(c->E.run(false));
        TRY(c->E.fetch_spec(0, spec_out, cap));
        return MIO_OK;
    });
}

int mio_istft(mio_ctx * c, const float * spec, int n_frames, float * wave_out, int64_t cap, int64_t * out_n) {
    return guarded([&]() -> int {
        if (!c || !spec || !wave_out || n_frames <= 0) return fail(MIO_ERR_ARG, "bad argument");
        const int64_t n = (int64_t) n_frames * c->E.hp.hop;
        if (out_n) *out_n = n;
        if (cap < n) return fail(MIO_ERR_ARG, "capacity too small");
        TRY(c->E.run_istft_only(spec, n_frames));
        TRY(c->E.fetch_wave(wave_out, cap));
        return MIO_OK;
    });
}

int mio_decode_wave(mio_ctx * c, const int32_t * codes, int n_codes, const float * emb, int peak_normalize,
                    float * wave_out, int64_t cap, int64_t * out_n) {
    return guarded([&]() -> int {
        if (!c || !wave_out) return fail(MIO_ERR_ARG, "bad argument");
        if (n_codes == 0) { if (out_n) *out_n = 0; return MIO_OK; }     // decode_codes_to_audio: n_codes <= 0 -> empty, true
        const int64_t n = (int64_t) n_codes * c->E.hp.samples_per_token;
        if (out_n) *out_n = n;
        if (cap < n) return fail(MIO_ERR_ARG, "capacity too small");
        int r = upload_one(c, codes, n_codes, emb);
        if (r != MIO_OK) return r;
        TRY(c->E.run(peak_normalize != 0));
        TRY(c->E.fetch_wave(wave_out, cap));
        return MIO_OK;
    });
}

int mio_batch_upload(mio_ctx * c, const int32_t * codes, const int32_t * cu, int n_utts, const float * vt, int nv,
                     const int32_t * vids) {
    return guarded([&]() -> int {
        if (!c) return fail(MIO_ERR_ARG, "bad argument");
        TRY(c->E.upload(codes, cu, n_utts, vt, nv, vids));
        return MIO_OK;
    });
}
int mio_batch_run(mio_ctx * c, int peak_normalize) {
    return guarded([&]() -> int {
        if (!c) return fail(MIO_ERR_ARG, "bad argument");
        TRY(c->E.run(peak_normalize != 0));
        return MIO_OK;
    });
}
int mio_batch_sync(mio_ctx * c) {
    if (!c) return fail(MIO_ERR_ARG, "bad argument");
    TRY(c->E.sync());
    return MIO_OK;
}
int mio_batch_fetch(mio_ctx * c, float * out, int64_t cap) {
    return guarded([&]() -> int {
        if (!c || !out) return fail(MIO_ERR_ARG, "bad argument");
        TRY(c->E.fetch_wave(out, cap));
        return MIO_OK;
    });
}
int64_t mio_batch_kernel_launches(const mio_ctx * c) { return c ? c->E.launches : 0; }
void * mio_batch_wave_device(mio_ctx * c) { return c ? (void *) c->E.d_wave : nullptr; }

int mio_decode_wave_batch(mio_ctx * c, const int32_t * codes, const int32_t * cu, int n_utts, const float * vt, int nv,
                          const int32_t * vids, int peak_normalize, float * out, int64_t cap) {
    return guarded([&]() -> int {
        if (!c || !out) return fail(MIO_ERR_ARG, "bad argument");
        // pipelined delivery (engine::upload: chunks as large as the workspace allows + one tail split); chunk i's D2H copy runs on a
        // second stream while chunk i+1 computes
        TRY(c->E.upload(codes, cu, n_utts, vt, nv, vids, c->E.fetch_chunk_codes + c->E.fetch_chunk_codes / 16));
        if (cap < c->E.wave_samples) return fail(MIO_ERR_ARG, "capacity too small");
        TRY(c->E.run(peak_normalize != 0, out));
        return MIO_OK;
    });
}

int mio_decode_pcm16_batch(mio_ctx * c, const int32_t * codes, const int32_t * cu, int n_utts, const float * vt, int nv,
                           const int32_t * vids, int peak_normalize, int16_t * out, int64_t cap) {
    return guarded([&]() -> int {
        if (!c || !out) return fail(MIO_ERR_ARG, "bad argument");
        TRY(c->E.upload(codes, cu, n_utts, vt, nv, vids, c->E.fetch_chunk_codes + c->E.fetch_chunk_codes / 16));
        if (cap < c->E.wave_samples) return fail(MIO_ERR_ARG, "capacity too small");
        TRY(c->E.run(peak_normalize != 0, nullptr, out));
        return MIO_OK;
    });
}
int mio_batch_fetch_pcm16(mio_ctx * c, int16_t * out, int64_t cap) {
    return guarded([&]() -> int {
        if (!c || !out) return fail(MIO_ERR_ARG, "bad argument");
        TRY(c->E.fetch_pcm16(out, cap));
        return MIO_OK;
    });
}

int mio_ctx_reserve(mio_ctx * c, int64_t max_batch_codes, int max_utts, int64_t max_emit_samples) {
    return guarded([&]() -> int {
        if (!c) return fail(MIO_ERR_ARG, "bad argument");
        TRY(c->E.reserve(max_batch_codes, max_utts, max_emit_samples));
        return MIO_OK;
    });
}

void * mio_batch_pcm16_device(mio_ctx * c) {
    if (!c) { fail(MIO_ERR_ARG, "bad argument"); return nullptr; }
    int16_t * p = nullptr;
    const std::string e = c->E.pcm16_device(&p);
    if (!e.empty()) { fail(MIO_ERR_CUDA, e); return nullptr; }
    return p;
}

int mio_batch_pcm16_to_device(mio_ctx * c, void * dst_device, int64_t cap) {
    return guarded([&]() -> int {
        if (!c) return fail(MIO_ERR_ARG, "bad argument");
        TRY(c->E.pcm16_to_device((int16_t *) dst_device, cap));
        return MIO_OK;
    });
}

// ---- device buffers shared between the processes of one box (the P2P form of the gather, sharding.P2PGather)
void * mio_device_alloc(int device, size_t bytes) {
    void * p = nullptr;
    if (cudaSetDevice(device) != cudaSuccess || cudaMalloc(&p, bytes ? bytes : 1) != cudaSuccess) {
        fail(MIO_ERR_CUDA, "cudaMalloc failed");
        return nullptr;
    }
    return p;
}
void mio_device_free(void * p) { if (p) cudaFree(p); }
int mio_ipc_export(const void * device_ptr, uint8_t out_handle[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    if (!device_ptr || !out_handle) return fail(MIO_ERR_ARG, "bad argument");
    cudaIpcMemHandle_t h;
    const cudaError_t e = cudaIpcGetMemHandle(&h, const_cast<void *>(device_ptr));
    if (e != cudaSuccess) return fail(MIO_ERR_CUDA, std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e));
    memcpy(out_handle, &h, 64);
    return MIO_OK;
}
void * mio_ipc_open(const uint8_t handle[64], int device) {
    if (!handle) { fail(MIO_ERR_ARG, "bad argument"); return nullptr; }
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    void * p = nullptr;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);   // peer access device -> owner comes with it
    if (e != cudaSuccess) { fail(MIO_ERR_CUDA, std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e)); return nullptr; }
    return p;
}
int mio_ipc_close(void * p) {
    if (!p) return MIO_OK;
    return cudaIpcCloseMemHandle(p) == cudaSuccess ? MIO_OK : fail(MIO_ERR_CUDA, "cudaIpcCloseMemHandle failed");
}

// ---- host-side neighbours of the path
// mio_parse_speech_tokens is a RESTATEMENT of a 20-line reference function (parse_speech_tokens, src/token-parser.cpp:5-28),
// not a design of its own: the contract to reproduce bit for bit is libc's strstr/strtol behaviour on the token text, so the
// loop has the same shape by necessity.  tests/test_ingest_epilogue.py holds it to the compiled reference function.
int64_t mio_parse_speech_tokens(const char * text, int64_t text_len, int32_t * out, int64_t cap) {
    if (!text || text_len < 0 || cap < 0 || (!out && cap > 0)) return fail(MIO_ERR_ARG, "bad argument");
    // A private NUL-terminated copy: the reference scans text.c_str() with strstr/strtol (src/token-parser.cpp:9-16),
    // which is where "stops at an embedded NUL" and strtol's whitespace/sign/overflow behaviour come from.
    const std::string buf(text, (size_t) text_len);
    const char * p = buf.c_str();
    const char * const end = p + buf.size();
    int64_t n = 0;
    while (p < end) {
        const char * tag = strstr(p, "<|s_");
        if (!tag) break;
        const char * digits = tag + 4;
        char * after = nullptr;
        const long v = strtol(digits, &after, 10);
        const bool closed = after && after > digits && after + 1 < end && after[0] == '|' && after[1] == '>';
        if (closed) {
            if (n < cap) out[n] = (int32_t) (int) v;
            n++;
            p = after + 2;
        } else {
            p = tag + 1;
        }
    }
    return n;
}

// the one entry point that allocates nothing (csrc/noalloc.cu: scratch in a file-scope __device__ array)
int mio_peak_normalize_device_noalloc(int device, const float * wave_in, float * wave_out, const int64_t * sample_offsets, int n_utts,
                                      void * cuda_stream) {
    return guarded([&]() -> int {
        TRY(noalloc_peak_normalize(device, wave_in, wave_out, sample_offsets, n_utts, (cudaStream_t) cuda_stream));
        return MIO_OK;
    });
}

int mio_wav_header(int sample_rate, int64_t n_samples, uint8_t * h) {
    if (!h || sample_rate <= 0 || n_samples < 0) return fail(MIO_ERR_ARG, "bad argument");
    auto u32 = [&](int at, uint32_t v) { for (int i = 0; i < 4; i++) h[at + i] = (uint8_t) (v >> (8 * i)); };
    auto u16 = [&](int at, uint16_t v) { h[at] = (uint8_t) v; h[at + 1] = (uint8_t) (v >> 8); };
    const uint32_t data = (uint32_t) ((uint64_t) n_samples * 2u);      // the reference's uint32 field wraps the same way
    memcpy(h + 0, "RIFF", 4); u32(4, 36u + data); memcpy(h + 8, "WAVE", 4); memcpy(h + 12, "fmt ", 4);
    u32(16, 16); u16(20, 1); u16(22, 1); u32(24, (uint32_t) sample_rate); u32(28, (uint32_t) sample_rate * 2u);
    u16(32, 2); u16(34, 16); memcpy(h + 36, "data", 4); u32(40, data);
    return MIO_OK;
}

int mio_wav_write_pcm16(const char * path, const int16_t * pcm, int64_t n, int sample_rate) {
    if (!path || (!pcm && n > 0) || n < 0) return fail(MIO_ERR_ARG, "bad argument");
    uint8_t h[44];
    if (mio_wav_header(sample_rate, n, h) != MIO_OK) return MIO_ERR_ARG;
    FILE * f = fopen(path, "wb");
    if (!f) return fail(MIO_ERR_IO, std::string("cannot open ") + path + " for writing");
    bool ok = fwrite(h, 1, 44, f) == 44;
    if (ok && n > 0) ok = fwrite(pcm, sizeof(int16_t), (size_t) n, f) == (size_t) n;
    ok = (fclose(f) == 0) && ok;
    return ok ? MIO_OK : fail(MIO_ERR_IO, std::string("short write to ") + path);
}

double mio_ctx_stream_seconds(const mio_ctx * c, int64_t * out_passes) {
    if (out_passes) *out_passes = c ? c->E.stream_passes : 0;
    return c ? c->E.stream_seconds : 0.0;
}

// ---- streaming
void mio_stream_default_config(const mio_ctx * c, mio_stream_config * cfg) {
    if (c && cfg) stream_default_config(c->E, cfg);
}
int mio_stream_open(mio_ctx * c, const float * emb, const mio_stream_config * cfg, mio_stream ** out) {
    return guarded([&]() -> int {
        if (!c || !emb || !out) return fail(MIO_ERR_ARG, "bad argument");
        *out = nullptr;
        stream_state * s = nullptr;
        std::string e = stream_open(c->E, emb, cfg, &s);
        if (!e.empty()) return fail(classify(e), e);
        mio_stream * h = new mio_stream();
        h->S = s;
        *out = h;
        return MIO_OK;
    });
}
void mio_stream_close(mio_stream * s) {
    if (!s) return;
    if (s->S && s->S->eng && s->S->eng->pass.active) (void) streams_drain(*s->S->eng);   // a round in flight may reference this stream
    delete s->S;
    delete s;
}
int mio_stream_push_codes(mio_stream * s, const int32_t * codes, int n) {
    return guarded([&]() -> int {
        if (!s || (!codes && n > 0) || n < 0) return fail(MIO_ERR_ARG, "bad argument");
        TRY(stream_push(*s->S, codes, n));
        return MIO_OK;
    });
}
int mio_stream_finish(mio_stream * s) {
    if (!s) return fail(MIO_ERR_ARG, "bad argument");
    TRY(stream_finish(*s->S));
    return MIO_OK;
}
int mio_streams_step(mio_ctx * c, mio_stream ** streams, int n) {
    return guarded([&]() -> int {
        if (!c || (!streams && n > 0) || n < 0) return fail(MIO_ERR_ARG, "bad argument");
        std::vector<stream_state *> v(n);
        for (int i = 0; i < n; i++) v[i] = streams[i] ? streams[i]->S : nullptr;
        TRY(streams_step(c->E, v.data(), n));
        return MIO_OK;
    });
}
int mio_streams_step_begin(mio_ctx * c, mio_stream ** streams, int n) {
    return guarded([&]() -> int {
        if (!c || (!streams && n > 0) || n < 0) return fail(MIO_ERR_ARG, "bad argument");
        std::vector<stream_state *> v(n);
        for (int i = 0; i < n; i++) v[i] = streams[i] ? streams[i]->S : nullptr;
        bool in_flight = false;
        TRY(streams_step_begin(c->E, v.data(), n, &in_flight));
        return in_flight ? 1 : 0;
    });
}
int mio_streams_step_end(mio_ctx * c, mio_stream ** streams, int n) {
    return guarded([&]() -> int {
        if (!c || (!streams && n > 0) || n < 0) return fail(MIO_ERR_ARG, "bad argument");
        std::vector<stream_state *> v(n);
        for (int i = 0; i < n; i++) v[i] = streams[i] ? streams[i]->S : nullptr;
        TRY(streams_step_end(c->E, v.data(), n));
        return MIO_OK;
    });
}
int64_t mio_stream_read(mio_stream * s, float * out, int64_t cap, int * is_last) {
    if (!s || (!out && cap > 0)) return fail(MIO_ERR_ARG, "bad argument");
    return stream_read(*s->S, out, cap, is_last);
}
int mio_stream_stats(const mio_stream * s, int64_t * dc, int64_t * codes, int64_t * emitted) {
    if (!s) return fail(MIO_ERR_ARG, "bad argument");
    if (dc) *dc = s->S->decode_calls;
    if (codes) *codes = s->S->decoded_codes;
    if (emitted) *emitted = s->S->emitted_samples;
    return MIO_OK;
}

// ---- pinned host memory for callers that want asynchronous / full-rate PCIe transfers
void * mio_host_alloc(size_t bytes) {
    void * p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { g_err = "cudaMallocHost failed"; return nullptr; }
    return p;
}
void mio_host_free(void * p) { if (p) cudaFreeHost(p); }

// ---- per-category device timings accumulated since the last reset (CUDA events around every launch)
int mio_profile_enable(mio_ctx * c, int on) {
    if (!c) return fail(MIO_ERR_ARG, "bad argument");
    c->E.profiling = on != 0;
    for (auto & t : c->E.prof_tot) t = prof_total();
    return MIO_OK;
}
int mio_profile_categories(void) { return CAT_COUNT; }
const char * mio_profile_name(int cat) {
    static const char * names[CAT_COUNT] = {"gemm_split3", "gemm_f16", "attention", "layernorm", "gn_stats", "gn_apply",
                                            "split", "gather", "cond", "istft", "peak"};
    return (cat >= 0 && cat < CAT_COUNT) ? names[cat] : "";
}
int mio_profile_read(mio_ctx * c, int cat, double * ms, double * flops, double * bytes, int64_t * launches) {
    if (!c || cat < 0 || cat >= CAT_COUNT) return fail(MIO_ERR_ARG, "bad argument");
    const prof_total & t = c->E.prof_tot[cat];
    if (ms) *ms = t.ms;
    if (flops) *flops = t.flops;
    if (bytes) *bytes = t.bytes;
    if (launches) *launches = t.launches;
    return MIO_OK;
}
int64_t mio_batch_h2d_bytes(const mio_ctx * c) { return c ? (int64_t) c->E.h2d_bytes : 0; }

int mio_debug_write(mio_ctx * c, const char * name, const float * data, int64_t n) {
    return guarded([&]() -> int {
        if (!c || !name || n < 0 || (!data && n > 0)) return fail(MIO_ERR_ARG, "bad argument");
        if (n == 0) c->E.overrides.erase(name);
        else c->E.overrides[name].assign(data, data + n);
        return MIO_OK;
    });
}
int mio_debug_clear(mio_ctx * c) {
    if (!c) return fail(MIO_ERR_ARG, "bad argument");
    c->E.overrides.clear();
    return MIO_OK;
}

int mio_debug_gemm(mio_ctx * c, int M, int N, int K, const float * A, const float * B, int npass, float * out) {
    return guarded([&]() -> int {
        if (!c) return fail(MIO_ERR_ARG, "bad argument");
        TRY(c->E.debug_gemm(M, N, K, A, B, npass, out));
        return MIO_OK;
    });
}

int64_t mio_debug_read(mio_ctx * c, const char * name, float * out, int64_t cap) {
    if (!c || !name) return fail(MIO_ERR_ARG, "bad argument");
    auto it = c->E.snaps.find(name);
    if (it == c->E.snaps.end()) return fail(MIO_ERR_ARG, std::string("bad argument: no snapshot named ") + name);
    const int64_t n = (int64_t) it->second.size();
    if (!out) return n;
    if (cap < n) return fail(MIO_ERR_ARG, "capacity too small");
    memcpy(out, it->second.data(), (size_t) n * sizeof(float));
    return n;
}

}  // extern "C"
