// Fused ISTFT: batched shared-memory 392-point inverse real FFT (14x14 decomposition, radix-7/2 butterflies in
// registers) + Hann window + 4-frame overlap-add + window-envelope normalisation + centre trim.
// Replaces istft() / irfft() of the reference (src/istft.cpp:43-108).  Each spectrogram frame is read once from
// HBM (plus a 4-frame halo per block) and each output sample is written exactly once.
//
// Block = 384 threads.  A block produces ISTFT_SEG_PER_BLOCK output hops (98 samples each) of one utterance and
// therefore needs the frames g0-2 .. g0+G+1 (G+4 = 48 frame slots in shared memory, 1680 B each; 9 % of the FFTs are
// halo recomputation, two blocks = 24 warps per SM).
//   phase 0: one bulk async copy (cp.async.bulk, 1600 B = one padded spectrogram row) per frame, issued by the lanes
//            of warp 0 behind one mbarrier: no per-element load instructions, all rows in flight at once
//   phase 1: every 16-lane group (14 active lanes) transforms one frame in place, 24 frames at a time
//   phase 2: each output sample sums its <= 4 windowed contributions in ascending frame order (the order of the
//            reference's accumulation loop, src/istft.cpp:84-93), divides by the window-square sum and is stored.
// The kernel is instruction-bound (about 350 warp instructions per frame for the FFT), not HBM-bound: see DESIGN.md.
#include "kernels.cuh"
#include "ptx.cuh"
#include "istft_math.h"

namespace mio {

constexpr int IST_NFFT = 392, IST_HOP = 98, IST_NFREQ = 197, IST_PAD = 147;
constexpr int IST_G = ISTFT_SEG_PER_BLOCK;
constexpr int IST_SLOTS = IST_G + 4;                       // 48
constexpr int IST_SLOT_FLOATS = 2 * IFFT_SLOT_F2;          // 420
constexpr int IST_THREADS = 384;                          // 24 groups of 16 lanes: the 48 slots in two passes
constexpr size_t IST_SMEM = (size_t) (IST_SLOTS * IST_SLOT_FLOATS + 2 * 196 * 2 + IST_NFFT + 100) * sizeof(float) + 16;

__device__ __forceinline__ void bulk_load_1d(void * dst, const void * src, uint32_t bytes, uint64_t * bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

__global__ void __launch_bounds__(IST_THREADS, 2) istft_kernel(istft_params p, const float * __restrict__ tw392g,
                                                            const float * __restrict__ tw196g) {
    pdl_sync();
    extern __shared__ __align__(16) float ist_smem[];
    float * slots = ist_smem;
    float * tw392 = slots + IST_SLOTS * IST_SLOT_FLOATS;
    float * tw196 = tw392 + 2 * 196;
    float * hann = tw196 + 2 * 196;
    float * wsum4 = hann + IST_NFFT;                       // [98] envelope of interior samples
    uint64_t * bar = (uint64_t *) (wsum4 + 100);

    // block -> (utterance, chunk)
    int lo = 0, hi = p.n_utts;                             // largest u with chunk_cu[u] <= blockIdx.x
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (p.chunk_cu[mid] <= (int) blockIdx.x) lo = mid; else hi = mid;
    }
    const int u = lo;
    const int chunk = blockIdx.x - p.chunk_cu[u];
    const int F = p.len[u] * p.frames_per_code;            // frames of this utterance
    const int g0 = chunk * IST_G;                          // first output hop of this block
    const int g1 = min(F, g0 + IST_G);
    const int f_lo = max(0, g0 - 2), f_hi = min(F, g1 + 2);   // frames needed [f_lo, f_hi)
    const int nfr = f_hi - f_lo;
    const float * spec = p.spec + ((size_t) p.start[u] * p.frames_per_code + f_lo) * p.spec_ld;
    const uint32_t row_bytes = (uint32_t) p.spec_ld * 4u;  // 1600: whole padded row, multiple of 16

    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(bar, (uint32_t) nfr * row_bytes);
    }
    __syncthreads();
    if (threadIdx.x < nfr) bulk_load_1d(slots + threadIdx.x * IST_SLOT_FLOATS, spec + (size_t) threadIdx.x * p.spec_ld, row_bytes, bar);
    // tables while the rows fly
    for (int i = threadIdx.x; i < 2 * 196; i += IST_THREADS) { tw392[i] = tw392g[i]; tw196[i] = tw196g[i]; }
    for (int i = threadIdx.x; i < IST_NFFT; i += IST_THREADS) hann[i] = p.hann[i];
    __syncthreads();
    if (threadIdx.x < IST_HOP) {
        // interior envelope: the four overlapping frames in ascending frame order == descending window index
        const int r = threadIdx.x;
        const int j0 = (r < 49) ? r + 343 : r + 245;
        float ws = 0.f;
#pragma unroll
        for (int k = 0; k < 4; k++) { const float w = hann[j0 - 98 * k]; ws += w * w; }
        wsum4[r] = ws;
    }
    mbar_wait(bar, 0);

    // phase 1
    {
        const int grp = threadIdx.x >> 4, t14 = threadIdx.x & 15;   // IST_THREADS / 16 groups of 16 lanes
#pragma unroll 1
        for (int base = 0; base < IST_SLOTS; base += IST_THREADS / 16) {
            const int fr = base + grp;
            const bool active = (t14 < 14) && (fr < nfr);
            float * slot = slots + fr * IST_SLOT_FLOATS;
            cpx a[14], b[14];
            if (active) ifft_load_z(slot, t14, tw392, a);
            __syncwarp();
            if (active) { dft14_inv(a, b); ifft_store_y(b, t14, tw196, slot); }
            __syncwarp();
            if (active) { ifft_load_y(slot, t14, a); dft14_inv(a, b); }
            __syncwarp();
            if (active) ifft_store_time(b, t14, hann, 1.0f / (float) IST_NFFT, slot);
        }
    }
    __syncthreads();

    // phase 2: trimmed output sample s = 98 H + r sits at untrimmed position P = s + 147 = 98 (H + 1) + (r + 49):
    // frames H-2 .. H+1 (r < 49) or H-1 .. H+2 (r >= 49) cover it with window indices j0, j0-98, j0-196, j0-294.
    // A thread owns one r (so the frame offset, the window index and the envelope are loop invariants) and walks
    // every third hop of the block: 294 of the 384 threads, ~10 instructions per sample.
    constexpr int P2 = IST_THREADS / IST_HOP;             // hop phases: 3
    if (threadIdx.x < P2 * IST_HOP) {
        float * out = p.wave + p.out_off[u];
        const int r = threadIdx.x % IST_HOP, par = threadIdx.x / IST_HOP;
        const int up = (r >= 49) ? 1 : 0;
        const int j0 = r + 343 - 98 * up;
        const float ws4 = wsum4[r];
        const float * col = slots + j0;                   // + (frame - f_lo) * IST_SLOT_FLOATS - 98 * k
        for (int H = g0 + par; H < g1; H += P2) {
            const int t0 = H - 2 + up;
            float acc = 0.f, wsum;
            if (t0 >= 0 && t0 + 3 < F) {
                const float * sl = col + (t0 - f_lo) * IST_SLOT_FLOATS;
#pragma unroll
                for (int k = 0; k < 4; k++) acc += sl[k * (IST_SLOT_FLOATS - IST_HOP)];
                wsum = ws4;
            } else {
                wsum = 0.f;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const int t = t0 + k, j = j0 - IST_HOP * k;
                    if (t >= 0 && t < F) {
                        acc += slots[(t - f_lo) * IST_SLOT_FLOATS + j];
                        const float w = hann[j];
                        wsum += w * w;
                    }
                }
            }
            if (wsum > 1e-8f) acc /= wsum;
            out[(size_t) H * IST_HOP + r] = acc;
        }
    }
}

// twiddles of the two DFT-14 passes: tw392 (196 complex) followed by tw196 (196 complex); built by the engine once per
// context (engine::load / init_istft_only) so that contexts on different devices never share device pointers
void istft_twiddles(float * h /*[784]*/) {
    for (int k = 0; k < 196; k++) {
        h[2 * k] = (float) cos(2.0 * M_PI * k / 392.0);
        h[2 * k + 1] = (float) sin(2.0 * M_PI * k / 392.0);
        h[392 + 2 * k] = (float) cos(2.0 * M_PI * k / 196.0);
        h[392 + 2 * k + 1] = (float) sin(2.0 * M_PI * k / 196.0);
    }
}

cudaError_t istft_init_device() {
    return cudaFuncSetAttribute(istft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) IST_SMEM);
}

cudaError_t launch_istft(const istft_params & p, cudaStream_t st) {
    if (p.total_chunks <= 0) return cudaSuccess;
    if (p.spec_ld % 4 != 0 || p.spec_ld < 2 * IST_NFREQ) return cudaErrorInvalidValue;   // 16-byte rows for the bulk copies
    if (!p.tw) return cudaErrorInvalidValue;
    launch_k(istft_kernel, dim3(p.total_chunks), dim3(IST_THREADS), IST_SMEM, st, p, p.tw, p.tw + 392);
    return cudaGetLastError();
}

}  // namespace mio
