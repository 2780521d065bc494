// mio_peak_normalize_device_noalloc: the per-utterance peak normalisation (0.95 / max|s|, the arithmetic of
// peak_max_kernel / peak_scale_kernel in elementwise.cu) for callers that are not allowed to allocate device memory.
//
// Who that is: a plug-in living inside a serving process that handed the whole of HBM to its own pool at start-up (a
// cudaMalloc from us would fail, or succeed and eat into someone else's budget), and a real-time audio thread that must
// never wait on the driver's allocator (cudaMalloc / cudaFree take a device-wide lock and may synchronise with all work
// in flight).  Every other entry point of this library owns a context whose workspace grows by cudaMalloc; this one owns
// nothing.  Its only scratch, one float per utterance holding that utterance's running maximum, lives in the
// file-scope array below, which is part of the library image: each GPU that loads the library carries the 16 KiB from
// the first use of the library on that GPU until the process exits, whether or not this entry point is ever called.
//
// This entry point has no counterpart in the reference (which normalises on the host, src/test-to-speech.cpp:289-300).
#include "../../include/miocodec_b200.h"

#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>
#include <string>

namespace mio {

// One slot per utterance.  Shared by every call on a GPU and never cleared behind the caller's back: a call zeroes the
// slots it is about to use (on its stream, ahead of its kernels) and leaves its maxima behind when it is done.
static __device__ __align__(256) float g_noalloc_peak[MIO_NOALLOC_MAX_UTTS];

constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mu;
static float * g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};   // device address of g_noalloc_peak on each GPU (0 = not looked up yet)

// In host code `g_noalloc_peak` is only a handle the runtime uses to find the symbol; it is NOT a device address and must
// never be given to a kernel or to cudaMemset.  The address differs per GPU (each has its own copy of the array), so it is
// asked for once per device, with that device current, and remembered.
static cudaError_t noalloc_scratch(int device, float ** out) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        float * p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **) &p, g_noalloc_peak);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}

// offs[u] .. offs[u+1] are utterance u's samples; blockIdx.y = utterance, blockIdx.x strides over its samples
__global__ void noalloc_peak_max_kernel(const float * __restrict__ wave, const int64_t * __restrict__ offs,
                                        float * __restrict__ peak) {
    const int u = blockIdx.y;
    const int64_t b = offs[u], n = offs[u + 1] - b;
    const float * w = wave + b;
    float m = 0.f;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) m = fmaxf(m, fabsf(w[i]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax((int *) (peak + u), __float_as_int(m));   // m >= 0: int order == float order
}

// in and out may be the same buffer, hence no __restrict__ on them
__global__ void noalloc_peak_scale_kernel(const float * in, float * out,
                                          const int64_t * __restrict__ offs, const float * __restrict__ peak) {
    const int u = blockIdx.y;
    const float pk = peak[u];
    const bool scale = pk > 1e-8f;                          // a silent utterance is copied unchanged, as in the reference
    if (!scale && in == out) return;
    const float gain = scale ? 0.95f / pk : 1.0f;
    const int64_t b = offs[u], n = offs[u + 1] - b;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const float s = in[b + i];
        out[b + i] = scale ? s * gain : s;
    }
}

// returns "" or an error text (capi.cu turns it into a code + mio_last_error)
std::string noalloc_peak_normalize(int device, const float * in, float * out, const int64_t * offs, int n_utts, cudaStream_t st) {
    if (!in || !out || !offs || n_utts < 0) return "bad argument";
    if (n_utts > MIO_NOALLOC_MAX_UTTS)
        return "bad argument: n_utts = " + std::to_string(n_utts) + " exceeds MIO_NOALLOC_MAX_UTTS = " + std::to_string(MIO_NOALLOC_MAX_UTTS) +
               ", the size of the built-in scratch of mio_peak_normalize_device_noalloc (split the batch)";
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return "bad argument: device ordinal out of range";
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return std::string("cudaSetDevice: ") + cudaGetErrorString(e);
    if (n_utts == 0) return "";
    float * peak = nullptr;
    if ((e = noalloc_scratch(device, &peak)) != cudaSuccess) return std::string("cudaGetSymbolAddress: ") + cudaGetErrorString(e);
    // the previous call's maxima are still in the slots: start this call's from zero
    if ((e = cudaMemsetAsync(peak, 0, sizeof(float) * (size_t) n_utts, st)) != cudaSuccess) return std::string("cudaMemsetAsync: ") + cudaGetErrorString(e);
    const dim3 grid(32, n_utts);
    noalloc_peak_max_kernel<<<grid, 256, 0, st>>>(in, offs, peak);
    noalloc_peak_scale_kernel<<<grid, 256, 0, st>>>(in, out, offs, peak);
    if ((e = cudaGetLastError()) != cudaSuccess) return std::string("kernel launch: ") + cudaGetErrorString(e);
    return "";
}

}  // namespace mio
