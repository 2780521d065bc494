// Fused ISTFT: batched shared-memory 392-point inverse real FFT (14x14 decomposition, radix-7/2 butterflies in
// registers) + Hann window + 4-frame overlap-add + window-envelope normalisation + centre trim.
// Replaces istft() / irfft() of the reference (src/istft.cpp:43-108).  Each spectrogram frame is read once from
// HBM (plus a 4-frame halo per block) and each output sample is written exactly once.
//
// Block = 128 threads.  A block produces ISTFT_SEG_PER_BLOCK output hops (98 samples each) of one utterance and
// therefore needs the frames g0-2 .. g0+G+1 (G+4 = 32 frame slots in shared memory, 1680 B each).
//   phase 0: cooperative coalesced load of the spectra into the slots
//   phase 1: every 14-lane group transforms one frame in place (2 frames per warp at a time)
//   phase 2: each output sample sums its <= 4 windowed contributions in ascending frame order (the order of the
//            reference's accumulation loop, src/istft.cpp:84-93), divides by the window-square sum and is stored.
#include "kernels.cuh"
#include "istft_math.h"

namespace mio {

constexpr int IST_NFFT = 392, IST_HOP = 98, IST_NFREQ = 197, IST_PAD = 147;
constexpr int IST_G = ISTFT_SEG_PER_BLOCK;
constexpr int IST_SLOTS = IST_G + 4;                       // 32
constexpr int IST_SLOT_FLOATS = 2 * IFFT_SLOT_F2;          // 420
constexpr size_t IST_SMEM = (size_t) (IST_SLOTS * IST_SLOT_FLOATS + 2 * 196 * 2 + IST_NFFT) * sizeof(float);

__global__ void __launch_bounds__(128) istft_kernel(istft_params p, const float * __restrict__ tw392g,
                                                    const float * __restrict__ tw196g) {
    extern __shared__ float ist_smem[];
    float * slots = ist_smem;
    float * tw392 = slots + IST_SLOTS * IST_SLOT_FLOATS;
    float * tw196 = tw392 + 2 * 196;
    float * hann = tw196 + 2 * 196;

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

    for (int i = threadIdx.x; i < 2 * 196; i += blockDim.x) { tw392[i] = tw392g[i]; tw196[i] = tw196g[i]; }
    for (int i = threadIdx.x; i < IST_NFFT; i += blockDim.x) hann[i] = p.hann[i];
    // phase 0: frame rows are spec_ld floats apart (8-byte aligned) -> float2 stream
    {
        const float2 * src = (const float2 *) spec;
        const int total = nfr * IST_NFREQ;
        const int ld2 = p.spec_ld >> 1;
        for (int i = threadIdx.x; i < total; i += blockDim.x) {
            const int fr = i / IST_NFREQ, k = i - fr * IST_NFREQ;
            const float2 v = __ldg(src + (size_t) fr * ld2 + k);
            *(float2 *) (slots + fr * IST_SLOT_FLOATS + 2 * k) = v;
        }
    }
    __syncthreads();

    // phase 1
    {
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        const int sub = lane >> 4, t14 = lane & 15;        // two 14-lane groups per warp
        for (int base = 0; base < IST_SLOTS; base += 8) {
            const int fr = base + warp * 2 + sub;
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

    // phase 2: trimmed output sample s (0 <= s < 98 F) sits at untrimmed position P = s + 147
    {
        float * out = p.wave + p.out_off[u];
        const int s0 = g0 * IST_HOP, s1 = g1 * IST_HOP;
        for (int s = s0 + threadIdx.x; s < s1; s += blockDim.x) {
            const int P = s + IST_PAD;
            int t_hi = P / IST_HOP;                         // last frame covering P
            int t_lo = (P - (IST_NFFT - 1) + IST_HOP - 1) / IST_HOP;   // ceil((P-391)/98), P-391 may be negative
            if (P - (IST_NFFT - 1) < 0) t_lo = 0;
            t_hi = min(t_hi, F - 1);
            float acc = 0.f, wsum = 0.f;
            for (int t = t_lo; t <= t_hi; t++) {
                const int j = P - t * IST_HOP;
                acc += slots[(t - f_lo) * IST_SLOT_FLOATS + j];
                const float w = hann[j];
                wsum += w * w;
            }
            if (wsum > 1e-8f) acc /= wsum;
            out[s] = acc;
        }
    }
}

static float * g_tw = nullptr;   // device: tw392 (392 floats) followed by tw196 (392 floats)
static int g_tw_dev = -1;

static cudaError_t ensure_tables() {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (g_tw && g_tw_dev == dev) return cudaSuccess;
    float h[4 * 196];
    for (int k = 0; k < 196; k++) {
        h[2 * k] = (float) cos(2.0 * M_PI * k / 392.0);
        h[2 * k + 1] = (float) sin(2.0 * M_PI * k / 392.0);
        h[392 + 2 * k] = (float) cos(2.0 * M_PI * k / 196.0);
        h[392 + 2 * k + 1] = (float) sin(2.0 * M_PI * k / 196.0);
    }
    e = cudaMalloc(&g_tw, sizeof(h));
    if (e != cudaSuccess) return e;
    e = cudaMemcpy(g_tw, h, sizeof(h), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return e;
    g_tw_dev = dev;
    return cudaFuncSetAttribute(istft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) IST_SMEM);
}

cudaError_t launch_istft(const istft_params & p, cudaStream_t st) {
    if (p.total_chunks <= 0) return cudaSuccess;
    cudaError_t e = ensure_tables();
    if (e != cudaSuccess) return e;
    istft_kernel<<<p.total_chunks, 128, IST_SMEM, st>>>(p, g_tw, g_tw + 392);
    return cudaGetLastError();
}

}  // namespace mio
