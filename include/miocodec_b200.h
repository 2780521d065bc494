/*
 * miocodec_b200 — C ABI of the B200-native codec back half (codes -> spectrogram -> waveform).
 *
 * This is the drop-in boundary for the hot path of espresso3389/MioTTS-llama.cpp.  Every entry
 * point below cites the reference interface it replaces (paths relative to the reference root).
 * Plain pointers and sizes only; all buffers are HOST memory unless the name says `_device`.
 * Every function returns MIO_OK (0) or a negative error code; mio_last_error() gives the text.
 * There is no CPU fallback: without a CUDA device mio_ctx_create fails with MIO_ERR_CUDA.
 *
 * Row/sample conventions (identical to the reference):
 *   spec  : [n_frames][n_fft/2+1][2] float, spec[t*394 + 2*f + {0:re,1:im}], n_frames = 18*n_codes
 *   wave  : float mono at sample_rate, exactly hop_length*n_frames = 1764*n_codes samples
 */
#ifndef MIOCODEC_B200_H
#define MIOCODEC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MIO_OK             0
#define MIO_ERR_ARG       -1   /* bad argument (null pointer, negative size, capacity too small) */
#define MIO_ERR_IO        -2   /* file missing / unreadable */
#define MIO_ERR_FORMAT    -3   /* not a GGUF v2/v3 file, missing tensor, unsupported dtype */
#define MIO_ERR_CUDA      -4   /* no device, CUDA error, kernel fault */
#define MIO_ERR_RANGE     -5   /* a code index outside [0, vocab) (the reference aborts in ggml) */
#define MIO_ERR_STATE     -6   /* call sequence error (e.g. fetch before run) */

typedef struct mio_ctx mio_ctx;
typedef struct mio_stream mio_stream;

/* ---- context: replaces miocodec_load / miocodec_free (src/miocodec.h:10-16, src/miocodec.cpp:426-512)
 * Parses the GGUF, reads the 22 KV hyper-parameters with the reference's defaults, uploads all weights
 * to HBM once (pre-split for the tensor-core kernels) and builds the ISTFT tables (istft_cache,
 * src/istft.cpp:7-32).  `device` is a CUDA ordinal. */
int  mio_ctx_create(const char * gguf_path, int device, mio_ctx ** out_ctx);
/* Waits for a streaming round still in flight.  Streams of this context that are still open are detached, not destroyed:
 * they keep the samples they have not handed out yet (mio_stream_read, mio_stream_stats and mio_stream_close keep working)
 * but can no longer be stepped.  Closing the streams first is the tidy order; the other order is safe. */
void mio_ctx_destroy(mio_ctx * ctx);
/* A context that can only run mio_istft: replaces constructing an istft_cache (src/istft.h:6-30,
 * src/istft.cpp:7-32) when no codec is loaded — the reference's istft() is a free function that needs no
 * miocodec_context.  Geometry is the reference's fixed n_fft = win_length = 392, hop = 98. */
int  mio_istft_ctx_create(int device, mio_ctx ** out_ctx);

/* accessors: miocodec_sample_rate / n_fft / hop_length / samples_per_token (src/miocodec.h:34-37) */
int  mio_ctx_sample_rate(const mio_ctx * ctx);
int  mio_ctx_n_fft(const mio_ctx * ctx);
int  mio_ctx_hop_length(const mio_ctx * ctx);
int  mio_ctx_samples_per_token(const mio_ctx * ctx);
int  mio_ctx_frames_per_code(const mio_ctx * ctx);   /* 2 * prod(upsampler factors) = 18 */
int  mio_ctx_vocab(const mio_ctx * ctx);
int  mio_ctx_device(const mio_ctx * ctx);

/* options: "gemm" = "tc" (tcgen05, default) | "simt" (fp32 CUDA-core validation path, still on the GPU);
 *          "max_chunk_codes" = integer, workspace rows at code resolution (default 65536, ~8.5 GB of HBM); must be
 *          set before the first decode;
 *          "fetch_tail_percent" = integer -1..50 (default -1 = auto): mio_decode_wave_batch / mio_decode_pcm16_batch copy chunk
 *          i to the host on a second stream while chunk i+1 computes; the batch is cut where the workspace forces it
 *          and the final chunk once more into (100-p) % and p %, so that the small tail computes while the big piece is
 *          copied and only the tail's own copy is exposed (p ~ D2H time / (compute + D2H time)); 0 = no split; -1 = start
 *          at 12 % and re-estimate p after every call from the measured compute and copy times of the big piece (the
 *          copy is slower when several GPUs of one host deliver at once);
 *          "fetch_chunk_codes" = integer: optional extra cap on the chunk size of that pipeline (default: none);
 *          "debug_snapshots" = "1" | "0": keep a host copy of every tap of the next decodes (mio_debug_read);
 *          "debug_snapshot_filter" = comma-separated tap names (empty = all). */
int  mio_ctx_set_option(mio_ctx * ctx, const char * key, const char * value);
/* run all kernels of this context on an existing CUDA stream (cudaStream_t as void*), e.g. torch's */
int  mio_ctx_set_stream(mio_ctx * ctx, void * cuda_stream);
/* Pre-size the context for batches of up to max_batch_codes codes in max_utts utterances and for streaming rounds that
 * emit up to max_emit_samples samples (0 = skip that part): device output / layout buffers, the PCM16 image, the pinned
 * staging and emit blocks.  Optional -- everything grows on demand -- but a serving process should call it at start-up:
 * growing by cudaMalloc / cudaMallocHost costs a first streaming session ~2 of its ~7 s. */
int  mio_ctx_reserve(mio_ctx * ctx, int64_t max_batch_codes, int max_utts, int64_t max_emit_samples);

/* ---- voice embedding: replaces load_voice_embedding (src/miocodec.h:41, src/miocodec.cpp:816-853).
 * Reads the FIRST tensor of the file, F32 only. Returns the number of floats (128) or <0. */
int  mio_voice_load(const char * emb_gguf_path, float * out, int cap);

/* ---- debug listing: replaces miocodec_print_tensors (src/miocodec.h:43, src/miocodec.cpp:859-877) */
int  mio_print_tensors(const char * gguf_path);

/* ---- single utterance: replaces miocodec_decode (src/miocodec.h:26-32, src/miocodec.cpp:519-810).
 * spec_out may be NULL to query sizes. *out_n_frames = 18*n_codes. */
int  mio_decode_spec(mio_ctx * ctx, const int32_t * codes, int n_codes, const float * global_emb,
                     float * spec_out, int64_t spec_cap_floats, int * out_n_frames);

/* ---- replaces istft() (src/istft.h:38-42, src/istft.cpp:68-108) with win_length == n_fft as the
 * reference always constructs it (src/test-to-speech.cpp:67).  Writes hop*n_frames samples. */
int  mio_istft(mio_ctx * ctx, const float * spec, int n_frames, float * wave_out, int64_t wave_cap,
               int64_t * out_n_samples);

/* ---- fused codes -> waveform: replaces decode_codes_to_audio (src/test-to-speech.cpp:248-303).
 * peak_normalize != 0 applies the reference's 0.95/max|s| gain (src/test-to-speech.cpp:289-300). */
int  mio_decode_wave(mio_ctx * ctx, const int32_t * codes, int n_codes, const float * global_emb,
                     int peak_normalize, float * wave_out, int64_t wave_cap, int64_t * out_n_samples);

/* ---- batch of independent utterances (the data-parallel form of the same call).
 * codes_concat[cu_seqlens[b] .. cu_seqlens[b+1]) are utterance b's codes; voice_table is
 * [n_voices][128]; voice_ids[b] selects a row.  Utterance b's samples are written at
 * wave_out + samples_per_token*cu_seqlens[b].  wave_cap >= samples_per_token*cu_seqlens[B]. */
int  mio_decode_wave_batch(mio_ctx * ctx, const int32_t * codes_concat, const int32_t * cu_seqlens, int n_utts,
                           const float * voice_table, int n_voices, const int32_t * voice_ids,
                           int peak_normalize, float * wave_out, int64_t wave_cap);

/* Same work split in three so a caller can time the device part alone (bench.py `value`):
 *   upload: H2D of codes / layout metadata / voice table into context-owned device buffers
 *   run   : kernels only, asynchronous on the context stream, result stays in HBM
 *   fetch : D2H of the waveform (synchronises the stream) */
int  mio_batch_upload(mio_ctx * ctx, const int32_t * codes_concat, const int32_t * cu_seqlens, int n_utts,
                      const float * voice_table, int n_voices, const int32_t * voice_ids);
int  mio_batch_run(mio_ctx * ctx, int peak_normalize);
int  mio_batch_fetch(mio_ctx * ctx, float * wave_out, int64_t wave_cap);
int  mio_batch_sync(mio_ctx * ctx);                       /* cudaStreamSynchronize + device error flag check */
int64_t mio_batch_kernel_launches(const mio_ctx * ctx);   /* kernels launched by the last mio_batch_run */
/* device pointer of the resident waveform of the last batch (float, samples_per_token*sum(T)) */
void * mio_batch_wave_device(mio_ctx * ctx);

/* ---- the steps either side of the path --------------------------------------
 * F2, pre-path ingest: replaces parse_speech_tokens (src/token-parser.h:8, src/token-parser.cpp:5-28): every
 * "<|s_N|>" in the text yields int(strtol(N)) with the reference's exact acceptance rule (optional whitespace / sign as
 * strtol takes them, at least one digit, "|>" must follow and must not end at the last byte + 1 of the text; the scan
 * stops at an embedded NUL like strstr on c_str()).  Returns the number of codes found (which may exceed cap) and
 * writes the first min(count, cap) of them; <0 on a bad argument. */
int64_t mio_parse_speech_tokens(const char * text, int64_t text_len, int32_t * codes_out, int64_t cap);

/* F1, post-path epilogue: peak normalisation (src/test-to-speech.cpp:289-300) + float -> PCM16 conversion
 * int16(clamp(s * 32767, -32768, 32767)) (truncation; src/wav-writer.cpp:39-42) run on the device behind the ISTFT,
 * so 2 bytes per sample cross PCIe instead of 4 and the two host passes of the reference disappear.
 * mio_decode_pcm16_batch = mio_decode_wave_batch delivering PCM16 (same chunked, overlapped D2H);
 * mio_batch_fetch_pcm16  = mio_batch_fetch delivering PCM16 of the resident waveform. */
int  mio_decode_pcm16_batch(mio_ctx * ctx, const int32_t * codes_concat, const int32_t * cu_seqlens, int n_utts,
                            const float * voice_table, int n_voices, const int32_t * voice_ids,
                            int peak_normalize, int16_t * pcm_out, int64_t pcm_cap);
int  mio_batch_fetch_pcm16(mio_ctx * ctx, int16_t * pcm_out, int64_t pcm_cap);
/* PCM16 image of the resident waveform, converted on the context's stream and LEFT ON THE DEVICE (int16,
 * samples_per_token*sum(T); valid until the next upload).  The sharded sweep's one exchange step (the final
 * gather of variable-length waveforms to the collecting rank) sends this buffer peer to peer.  NULL on error. */
void * mio_batch_pcm16_device(mio_ctx * ctx);
/* The same conversion stored straight into a caller-supplied, 16-byte aligned DEVICE buffer of at least
 * samples_per_token*sum(T) int16 (asynchronous on the context's stream).  The buffer may be ANOTHER GPU's memory mapped into this process (CUDA IPC /
 * peer access): the stores then cross NVLink inside the kernel and the gather needs no separate transfer step. */
int  mio_batch_pcm16_to_device(mio_ctx * ctx, void * dst_device, int64_t cap);
/* Device buffers shared between the processes of one box (one process per GPU): the collecting rank allocates, exports a
 * 64-byte handle (cudaIpcGetMemHandle), every other rank opens it on ITS device (cudaIpcOpenMemHandle with lazy peer access)
 * and passes the mapped pointer (+ its slice offset) to mio_batch_pcm16_to_device. */
void * mio_device_alloc(int device, size_t bytes);
void   mio_device_free(void * p);
int    mio_ipc_export(const void * device_ptr, uint8_t out_handle[64]);
void * mio_ipc_open(const uint8_t handle[64], int device);
int    mio_ipc_close(void * p);
/* The 44-byte RIFF/WAVE header wav_write emits (src/wav-writer.cpp:7-37): PCM, mono, 16 bit. */
int  mio_wav_header(int sample_rate, int64_t n_samples, uint8_t out44[44]);
/* header + samples to a file: byte-identical to wav_write(path, samples, sample_rate) when pcm came from the
 * conversion above (src/wav-writer.cpp:24-45). */
int  mio_wav_write_pcm16(const char * path, const int16_t * pcm, int64_t n_samples, int sample_rate);

/* ---- peak normalisation without any allocation.  This entry point has no counterpart in the reference.
 * It exists for callers that must not allocate device memory while they run: a plug-in inside a serving process whose pool
 * took all of HBM at start-up (our cudaMalloc would fail or spend another tenant's budget), or a real-time audio thread
 * that may not wait on the driver's allocator lock.  Everything else in this header works on a context whose workspace
 * grows by cudaMalloc; this call takes no context and calls neither cudaMalloc / cudaFree nor any routine that does.
 *
 * Applies the per-utterance gain 0.95 / max|s| (skipped when the peak is <= 1e-8; src/test-to-speech.cpp:289-300 does this
 * on the host) to a waveform that is already in device memory -- exactly what peak_normalize != 0 does inside
 * mio_decode_wave_batch, bit for bit.  ALL pointers are DEVICE memory on `device`, supplied by the caller:
 *   wave_in_device / wave_out_device : float samples; the two may be the same buffer (in place) but must not partly overlap
 *   sample_offsets_device            : int64[n_utts + 1], utterance u owns samples [offsets[u], offsets[u+1])
 *                                      (= samples_per_token * cu_seqlens for the output of mio_batch_run)
 *   cuda_stream                      : cudaStream_t as void* (NULL = the default stream); the call only enqueues work
 * The one float of scratch per utterance lives in a fixed array inside the library image, so n_utts may not exceed
 * MIO_NOALLOC_MAX_UTTS (more is refused with MIO_ERR_ARG: split the batch).  Each GPU the library is used on holds those
 * 16 KiB until the process exits.  Every call on a GPU uses the SAME array: two calls whose kernels can overlap on the GPU
 * (different streams, different threads) corrupt each other's maxima -- order them on one stream or with events.  The
 * array keeps the last call's maxima; each call clears the slots it needs itself.  The first call on each device looks up
 * the array's device address (cudaGetSymbolAddress, which also loads this library's kernels onto that GPU): make that first
 * call during start-up, before any stream capture begins.  Leaves `device` current, like the other calls of this library. */
#define MIO_NOALLOC_MAX_UTTS 4096
int  mio_peak_normalize_device_noalloc(int device, const float * wave_in_device, float * wave_out_device,
                                       const int64_t * sample_offsets_device, int n_utts, void * cuda_stream);

/* ---- streaming policy on top of the decoder: replaces the LLM-branch of
 * TestToSpeech::synthesize_stream_profiled (src/test-to-speech.cpp:430-625): commit holdback (32 codes),
 * min commit step (24), check cadence (every 20 pushed codes), full-prefix re-decode, llround() sample
 * indexing, 1323-sample linear crossfade and chunked emission (emit_range, :367-417).
 * Many streams are advanced together: mio_streams_step re-decodes, in ONE batched device pass, every
 * stream whose cadence fired since the last step. */
typedef struct mio_stream_config {
    int check_interval;      /* 20   (src/test-to-speech.cpp:499) */
    int holdback_codes;      /* 32   (:500) */
    int min_commit_codes;    /* 24   (:501) */
    int crossfade_samples;   /* min(sample_rate*3/100, 4096) = 1323 (:364); <0 = derive */
    int chunk_samples;       /* 4096 (src/test-to-speech.h:94) */
} mio_stream_config;
void mio_stream_default_config(const mio_ctx * ctx, mio_stream_config * cfg);
int  mio_stream_open(mio_ctx * ctx, const float * global_emb, const mio_stream_config * cfg, mio_stream ** out);
void mio_stream_close(mio_stream * s);
/* append generated codes (one per LLM token in the reference); cadence is evaluated per code */
int  mio_stream_push_codes(mio_stream * s, const int32_t * codes, int n);
/* mark end of generation: the next step flushes everything (maybe_emit(true), :616-618) */
int  mio_stream_finish(mio_stream * s);
/* advance all given streams: one batched decode for those with a pending emit */
int  mio_streams_step(mio_ctx * ctx, mio_stream ** streams, int n_streams);
/* The same round split in two so that the host works while the GPU decodes: _begin enqueues the round (upload, decode, cut +
 * crossfade, the copy of the emitted samples into pinned memory) and returns at once; _end waits for it, hands the samples to
 * the streams and runs whatever further passes they still need.  Between the two the caller may push codes to and read from
 * any stream (reads return what earlier rounds emitted) -- e.g. generate the next tokens -- but not close a stream.
 * mio_streams_step == _begin + _end.  _begin returns 1 when a round is now in flight, 0 when no stream had work, < 0 on error. */
int  mio_streams_step_begin(mio_ctx * ctx, mio_stream ** streams, int n_streams);
int  mio_streams_step_end(mio_ctx * ctx, mio_stream ** streams, int n_streams);
/* pop emitted audio (already crossfaded, in emission order). Returns samples copied. *is_last is set when
 * the stream has finished and its queue is drained. */
int64_t mio_stream_read(mio_stream * s, float * out, int64_t cap, int * is_last);
/* counters mirroring StreamProfile (src/test-to-speech.h:28-38): decode_calls, decoded_codes, emitted_samples */
int  mio_stream_stats(const mio_stream * s, int64_t * decode_calls, int64_t * decoded_codes, int64_t * emitted_samples);
/* the timers of StreamProfile (codec_sec + istft_sec, src/test-to-speech.h:31-32, accumulated at :548-552): wall seconds
 * this context has spent inside mio_streams_step (decoder + ISTFT + cut/crossfade + D2H are one fused device pass here, and
 * one pass serves every stream that fired, so the time belongs to the context, not to a stream) and the number of passes */
double mio_ctx_stream_seconds(const mio_ctx * ctx, int64_t * out_passes);

/* ---- pinned host buffers (cudaMallocHost) so callers can hand the library transfer-friendly memory */
void * mio_host_alloc(size_t bytes);
void   mio_host_free(void * p);

/* ---- per-kernel-category device timing: CUDA events are recorded around every launch while enabled;
 * totals accumulate over mio_batch_run calls until the next mio_profile_enable.  Read after mio_batch_sync. */
int  mio_profile_enable(mio_ctx * ctx, int on);
int  mio_profile_categories(void);
const char * mio_profile_name(int cat);
int  mio_profile_read(mio_ctx * ctx, int cat, double * ms, double * flops, double * bytes, int64_t * launches);
int64_t mio_batch_h2d_bytes(const mio_ctx * ctx);         /* bytes copied host->device by the last upload */

/* ---- introspection for parity tests: copy a named intermediate of the LAST decode to the host (option
 * "debug_snapshots" = "1" must be set before the decode).  Returns number of floats written or <0.  Names: the
 * snap_* / tap() calls of engine::run_chunk in csrc/engine.cu (e.g. "prenet_out", "prior0_c1", "dec3_out", "spec"). */
int64_t mio_debug_read(mio_ctx * ctx, const char * name, float * out, int64_t cap);
/* teacher forcing: the next decodes REPLACE the leading n floats of the named tap with `data` when the kernel sequence
 * reaches it and continue from there (n = 0 removes one override, mio_debug_clear all).  Lets a test feed each stage the
 * oracle's input for that stage — the stages of miocodec_decode, src/miocodec.cpp:598-740 — and compare its output at
 * fp32 accuracy, free of the error the earlier stages accumulated. */
int  mio_debug_write(mio_ctx * ctx, const char * name, const float * data, int64_t n);
int  mio_debug_clear(mio_ctx * ctx);
/* one bare GEMM through the product tensor-core kernel: out[M][N] = A[M][K] * B[N][K]^T, operands split (npass = 3: fp16
 * hi/lo pairs, the decoder's fp32-class layers) or rounded (npass = 1: fp16, its conv1d layers) exactly as the decoder
 * does; K % 64 == 0.  The per-op parity test of the kernel (cf. the linear layers of src/miocodec.cpp:205-209). */
int  mio_debug_gemm(mio_ctx * ctx, int M, int N, int K, const float * A, const float * B, int npass, float * out);

const char * mio_last_error(void);
const char * mio_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MIOCODEC_B200_H */
