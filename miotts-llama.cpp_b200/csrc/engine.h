// The decoder engine: owns the device weights, the workspace and the per-batch layout, and enqueues the kernel
// sequence of one ragged batch.  Host-side replacement of miocodec_decode()'s graph build + compute
// (src/miocodec.cpp:519-810 of the reference) followed by istft() (src/istft.cpp:68-108).
#pragma once
#include <map>
#include <chrono>
#include <string>
#include <vector>

#include "kernels.cuh"
#include "model.h"

namespace mio {

struct a_view {                  // an fp16 operand buffer viewed as [rows][K] with its TMA maps
    __half * hi = nullptr; __half * lo = nullptr;
    int K = 0, rows = 0;
    CUtensorMap tm_hi, tm_lo;
};

enum prof_cat : int { CAT_GEMM_SPLIT = 0, CAT_GEMM_F16, CAT_ATTN, CAT_LN, CAT_GN_STATS, CAT_GN_APPLY, CAT_SPLIT, CAT_GATHER,
                      CAT_COND, CAT_ISTFT, CAT_PEAK, CAT_COUNT };
struct prof_rec { int cat; double flops, bytes; size_t ev; };
struct prof_total { double ms = 0, flops = 0, bytes = 0; int64_t launches = 0; };

struct chunk_info {              // a run of whole utterances that fits the workspace
    int utt0 = 0, n_utts = 0;    // utterance range
    int row0 = 0, rows1 = 0;     // rows (code resolution) in the concatenated device arrays
    int cu0 = 0;                 // offset of this chunk's ISTFT block prefix in d_chunk_cu
    int istft_blocks = 0;
    int scu0 = 0, n_segs = 0;    // offset of this chunk's GroupNorm segment prefix in d_seg_cu, segments in the chunk
};

struct out_map_key {
    const void * base; int esize, rows, cols; size_t ld; int box_c, box_r;
    bool operator<(const out_map_key & o) const {
        if (base != o.base) return base < o.base;
        if (esize != o.esize) return esize < o.esize;
        if (rows != o.rows) return rows < o.rows;
        if (cols != o.cols) return cols < o.cols;
        if (ld != o.ld) return ld < o.ld;
        if (box_c != o.box_c) return box_c < o.box_c;
        return box_r < o.box_r;
    }
};

struct engine {
    int device = 0;
    hparams hp;
    weights w;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    bool use_simt = false;
    // serpentine row order: consecutive kernels of a chunk walk the rows in opposite directions, so each starts on the
    // rows its predecessor wrote last (still in the 126 MB L2).  MIO_SERPENTINE=0 disables (A/B measurements).
    bool serpentine = true, dir_flip = false;
    bool next_dir() { if (!serpentine) return false; dir_flip = !dir_flip; return dir_flip; }
    // chunked tensor-core accumulation (gemm3.cu): 0 = off, 1 = the fp32-class layers in front of the first conv1d
    // (default: the only place where the tensor core's truncating adder is audible, profiles/r2a_parity_attribution.md),
    // 2 = every fp32-class layer.  MIO_CHUNK_ACC=off|prenet|all, option "chunk_acc".
    bool exp_skip_norms = false; int64_t exp_runs = 0; int exp_slot_align = 0;   // measurement-only switches (MIO_EXP_*), see engine.cu
    int chunk_mode = 1;
    int chunk_kb = 3;                // k-blocks (64 of K each) per accumulation chunk: 3 costs no step time and gives 6e-7 (profiles/r2c/r2c_chunk_sweep.log; 2: 3.8e-7 at +4.6 %, 4: 8.4e-7); MIO_CHUNK_KB
    bool fuse_resblock_tail = true;  // MIO_FUSE_TAIL=0 / option "fuse_tail": A/B switch of the residual + split epilogue (engine::run_resblock)
    int num_sms = 148;               // multiProcessorCount of the device (tile-width choice of the GEMM dispatch)
    int max_rows1 = 65536 + 4096;    // workspace rows at code resolution (codes + one pad row per utterance), ~8.5 GB
    bool ws_ready = false;
    bool istft_only = false;         // created by mio_istft_ctx_create: no decoder weights, ISTFT workspace only
    bool debug_snap = false;
    std::map<std::string, std::vector<float>> snaps;
    std::string err;
    int64_t launches = 0;

    // ---- workspace (device)
    float *x768 = nullptr, *x512 = nullptr, *y512 = nullptr, *xu0 = nullptr, *yu0 = nullptr,
          *xu1 = nullptr, *yu1 = nullptr, *spec = nullptr, *mod = nullptr, *peak = nullptr;
    __half *aA_hi = nullptr, *aA_lo = nullptr, *aFF_hi = nullptr, *aFF_lo = nullptr, *c16 = nullptr,
           *hH_hi = nullptr, *hH_lo = nullptr, *qkv_hi = nullptr, *qkv_lo = nullptr;
    bool attn_simt = false;           // MIO_ATTN=simt: CUDA-core validation kernel
    double2 * gn_part = nullptr;      // [segments][32] (sum, sum of squares) of the GroupNorm being computed
    float * voice_silu = nullptr;     // [max_voices][128] silu(voice embedding), input of cond_kernel
    float2 * gn_mr = nullptr;         // [max_utts][32] (mean, rstd) of the GroupNorm being applied
    int2 * attn_b1 = nullptr;         // [max_rows1]   valid key range per prenet row (attn_bounds_kernel)
    int2 * attn_b2 = nullptr;         // [2*max_rows1] valid key range per decoder row
    int * err_flag = nullptr;
    a_view vA768, vA512, vFF2048, vFF1536, vFFu0, vFFu1, vFF512, vC512, vCu0, vCu1, vH;
    bool vFF512_ready = false;
    std::vector<void *> w_allocs;     // every weight allocation (freed in the destructor)
    int max_utts = 4096;
    int max_voices = 1024;

    // ---- current batch (device copies of the layout)
    batch_layout lay;
    std::vector<chunk_info> chunks;
    int cur_row0 = 0, cur_utt0 = 0, cur_rows1 = 0, cur_n_utts = 0, cur_scu0 = 0, cur_n_segs = 0;
    double cur_live_frac = 1.0;      // live codes / slot rows of the current chunk: the profile books FLOPs of live rows only
    size_t cap_rows = 0, cap_rows_ru = 0, cap_utts = 0, cap_ccu = 0, cap_scu = 0, cap_vt = 0;
    size_t h2d_bytes = 0;
    int32_t *d_codes = nullptr, *d_row_utt = nullptr, *d_start = nullptr, *d_len = nullptr, *d_voice = nullptr,
            *d_chunk_cu = nullptr, *d_seg_cu = nullptr;
    int64_t * d_out_off = nullptr;
    float * d_voice_table = nullptr;
    int n_voices = 0;
    float * d_wave = nullptr; int64_t wave_cap = 0;       // resident output of the last batch
    int64_t wave_samples = 0;
    int istft_chunks = 0;
    double stream_seconds = 0.0; int64_t stream_passes = 0;   // wall time / device passes inside streams_step
    // streaming emit buffers (stream.cu)
    void * d_emit = nullptr; void * d_emit_jobs = nullptr;
    int64_t emit_cap = 0, emit_jobs_cap = 0;         // samples / job-table entries; each grows on its own and never shrinks
    static constexpr size_t EMIT_JOB_BYTES = 64;     // allocation size of one job-table entry (>= sizeof(emit_job), stream.cu)
    // Pinned host blocks the emitted samples of a streaming round land in (one D2H copy per round).  The streams' queues
    // REFERENCE slices of the block (refs counts unread slices) instead of copying them again; a block returns to the pool
    // when its last slice has been read.  At most EMIT_POOL_MAX blocks are kept referenced: beyond that a round's samples
    // are copied into the streams (a consumer that never reads must not pin host memory without bound).
    struct emit_block { float * host = nullptr; int64_t cap = 0; int64_t refs = 0; };
    static constexpr int EMIT_POOL_MAX = 8;
    std::vector<emit_block *> emit_pool;
    emit_block * acquire_emit_block(int64_t samples, bool * pooled);
    // every stream_state opened on this context and not yet closed (stream.cu): mio_ctx_destroy detaches them so that a
    // stream closed or read AFTER its context never touches freed engine state
    std::vector<void *> open_streams;
    // a streaming pass in flight between streams_step_begin and streams_step_end (stream.cu)
    struct stream_pass {
        bool active = false, pooled = true;
        std::vector<void *> act;                      // stream_state * of the pass
        std::vector<int64_t> begin, end, off;         // cut range and offset in the emit block per stream
        std::vector<int64_t> prefix;                  // codes decoded per stream (booked into its counters when the pass ends well)
        std::vector<uint8_t> jobs_host;               // the emit job table, alive until the pass ends
        emit_block * blk = nullptr;
        int64_t n_codes = 0;
        std::chrono::steady_clock::time_point t0;
        double ms_prep = 0, ms_upload = 0, ms_run = 0, ms_emit = 0;
    } pass;
    // pinned staging
    void * h_stage = nullptr; size_t h_stage_bytes = 0;
    bool batch_uploaded = false, batch_ran = false;

    ~engine();
    std::string load(const std::string & gguf_path, int device);
    std::string init_istft_only(int device);
    std::string ensure_workspace();
    std::string ensure_rope(int max_pos);                 // grows the RoPE tables to cover positions [0, max_pos)
    // layout + H2D. codes_concat/cu_seqlens in the caller's (unpadded) indexing.
    std::string upload(const int32_t * codes, const int32_t * cu, int n_utts, const float * voice_table, int n_voices,
                       const int32_t * voice_ids, int chunk_rows_limit = 0);
    // enqueue all kernels of the uploaded batch; host_dst != nullptr: also copy each chunk's waveform to the host on a
    // second stream as soon as it is complete (overlaps with the next chunk) and return when everything has landed
    // host_pcm != nullptr: the same pipeline delivering PCM16 (converted on the device, half the D2H bytes)
    std::string run(bool peak_normalize, float * host_dst = nullptr, int16_t * host_pcm = nullptr);
    std::string run_impl(bool peak_normalize, float * host_dst, int16_t * host_pcm);
    std::string fetch_pcm16(int16_t * out, int64_t cap);
    std::string reserve(int64_t total_codes, int n_utts, int64_t emit_samples);
    std::string pcm16_device(int16_t ** out);
    std::string pcm16_to_device(int16_t * dst, int64_t cap);
    int16_t * d_pcm = nullptr; int64_t pcm_cap = 0;       // PCM16 image of d_wave (grown on demand)
    std::string ensure_pcm();
    cudaStream_t copy_stream = nullptr; cudaEvent_t chunk_done = nullptr;
    int fetch_tail_percent = -1;     // size of the last chunk of the host-delivery pipeline, % of the final workspace chunk;
                                     // < 0: auto = tail_auto_pct, re-estimated after every delivery from the measured
                                     // compute and copy times of the big piece (copy / (compute + copy))
    float tail_auto_pct = 12.f;
    int big_piece_chunk = -1;        // index of the chunk in front of the split tail (-1: no split in this batch)
    cudaEvent_t tune_ev[4] = {nullptr, nullptr, nullptr, nullptr};
    int fetch_chunk_codes = 1 << 20; // optional extra cap on the chunk size of the host-delivery pipeline
    std::string run_istft_only(const float * spec_host, int n_frames);   // mio_istft
    float * istft_spec = nullptr; size_t istft_spec_rows = 0;            // its staging buffer (grown on demand)
    std::string sync();                                   // stream sync + error flag
    std::string fetch_wave(float * out, int64_t cap);
    std::string fetch_spec(int utt, float * out, int64_t cap);
    row_map rmap() const;
    std::string run_chunk(const chunk_info & c);
    // split_out != nullptr (and no taps): the block's result is written as the (hi, lo) operand of the next product only
    // (conv2 epilogue adds the residual itself and splits); x is then NOT updated
    std::string run_resblock(const resblock_w & R, float * x, float * y, const a_view & vc, int rows, int row_div,
                             const std::string & tag, const a_view * split_out = nullptr);
    std::string run_istft(const chunk_info & c, const float * spec_dev);
    std::string view512_ff(a_view & v);
    cudaError_t gemm(const gemm_weight & W, const a_view & A, int M, gemm_params p);
    std::string debug_gemm(int M, int N, int K, const float * A, const float * B, int npass, float * out);
    std::map<out_map_key, CUtensorMap> out_maps;          // TMA maps of GEMM outputs, keyed by (buffer, shape)
    // box {box_c, box_r}; swizzle = box_c * esize (128-byte or 64-byte box rows)
    bool out_map(CUtensorMap * out, const void * base, int esize, int rows, int cols, size_t ld_bytes, int box_c = 32, int box_r = 32);
    gemm_params qkv_params(const row_map & rm, int D, int row_div);
    std::string run_attention(int rows, int D, int n_head, int window, int row_div, const a_view & out, const int2 * bounds);
    // profiling
    bool profiling = false;
    std::vector<cudaEvent_t> prof_events; size_t prof_used = 0;
    std::vector<prof_rec> prof_recs;
    prof_total prof_tot[CAT_COUNT];
    void prof_begin(int cat, double flops, double bytes);
    void prof_end();
    void prof_collect();
    // named taps of the kernel sequence: snapshot (debug_snapshots) and teacher-forced overrides (mio_debug_write)
    std::map<std::string, std::vector<float>> overrides;
    std::string snap_filter;          // option "debug_snapshot_filter": ",name1,name2," — empty = every tap
    bool taps_on() const { return debug_snap || !overrides.empty(); }
    bool snap_wanted(const std::string & name) const {
        return debug_snap && (snap_filter.empty() || snap_filter.find("," + name + ",") != std::string::npos);
    }
    void tap_f32(const std::string & name, float * d, size_t rows, size_t cols, size_t ld = 0);
    void tap_f16(const std::string & name, __half * hi, __half * lo, size_t n);
};

}  // namespace mio
