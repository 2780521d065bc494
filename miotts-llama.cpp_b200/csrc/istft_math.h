// Arithmetic of the 392-point inverse real FFT used by the ISTFT kernel, written as host/device functions so
// the exact same code is unit-tested on the CPU (tests/test_istft_math.py via csrc/istft_math_test.cpp).
//
// Replaces irfft() of the reference (src/istft.cpp:43-66), an O(N^2) table DFT:
//   x[n] = ( X0.re + (-1)^n X196.re + 2*sum_{k=1}^{195} (re_k cos(2 pi k n/392) - im_k sin(2 pi k n/392)) ) / 392
// i.e. the inverse DFT of the Hermitian extension with Im(X0) = Im(X196) = 0.
//
// Algorithm: pack the Hermitian spectrum into a 196-point complex sequence
//   Z[k] = (X[k] + conj(X[196-k])) + i * (X[k] - conj(X[196-k])) * e^{+2 pi i k/392},  k = 0..195
// whose 196-point inverse DFT z[m] = x[2m] + i x[2m+1].  196 = 14 x 14: fourteen cooperating threads each do a
// 14-point DFT (2 x radix-7 + radix-2) over registers, exchange through shared memory with the 196-th root
// twiddles applied, and do a second 14-point DFT.
#pragma once

#if defined(__CUDACC__)
#define MIO_HD __host__ __device__ __forceinline__
#else
#define MIO_HD inline
#endif

namespace mio {

struct alignas(8) cpx { float re, im; };   // 8-byte aligned: shared-memory accesses compile to 64-bit LDS/STS

MIO_HD cpx cmul(cpx a, cpx b) { return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }

// out[m] = sum_a in[a] e^{+2 pi i a m / 7}
MIO_HD void dft7_inv(const cpx * in, cpx * out) {
    const float c1 = 0.62348980185873353f, c2 = -0.22252093395631440f, c3 = -0.90096886790241913f;
    const float s1 = 0.78183148246802981f, s2 = 0.97492791218182361f, s3 = 0.43388373911755812f;
    const cpx p1 = {in[1].re + in[6].re, in[1].im + in[6].im}, q1 = {in[1].re - in[6].re, in[1].im - in[6].im};
    const cpx p2 = {in[2].re + in[5].re, in[2].im + in[5].im}, q2 = {in[2].re - in[5].re, in[2].im - in[5].im};
    const cpx p3 = {in[3].re + in[4].re, in[3].im + in[4].im}, q3 = {in[3].re - in[4].re, in[3].im - in[4].im};
    out[0] = {in[0].re + (p1.re + p2.re + p3.re), in[0].im + (p1.im + p2.im + p3.im)};
    // m = 1: cos (c1,c2,c3) sin (s1,s2,s3)
    {
        const cpx r = {in[0].re + c1 * p1.re + c2 * p2.re + c3 * p3.re, in[0].im + c1 * p1.im + c2 * p2.im + c3 * p3.im};
        const cpx s = {s1 * q1.re + s2 * q2.re + s3 * q3.re, s1 * q1.im + s2 * q2.im + s3 * q3.im};
        out[1] = {r.re - s.im, r.im + s.re};
        out[6] = {r.re + s.im, r.im - s.re};
    }
    // m = 2: cos (c2,c3,c1) sin (s2,-s3,-s1)
    {
        const cpx r = {in[0].re + c2 * p1.re + c3 * p2.re + c1 * p3.re, in[0].im + c2 * p1.im + c3 * p2.im + c1 * p3.im};
        const cpx s = {s2 * q1.re - s3 * q2.re - s1 * q3.re, s2 * q1.im - s3 * q2.im - s1 * q3.im};
        out[2] = {r.re - s.im, r.im + s.re};
        out[5] = {r.re + s.im, r.im - s.re};
    }
    // m = 3: cos (c3,c1,c2) sin (s3,-s1,s2)
    {
        const cpx r = {in[0].re + c3 * p1.re + c1 * p2.re + c2 * p3.re, in[0].im + c3 * p1.im + c1 * p2.im + c2 * p3.im};
        const cpx s = {s3 * q1.re - s1 * q2.re + s2 * q3.re, s3 * q1.im - s1 * q2.im + s2 * q3.im};
        out[3] = {r.re - s.im, r.im + s.re};
        out[4] = {r.re + s.im, r.im - s.re};
    }
}

// out[m] = sum_j in[j] e^{+2 pi i j m / 14}
MIO_HD void dft14_inv(const cpx * in, cpx * out) {
    cpx e[7], o[7], E[7], O[7];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int a = 0; a < 7; a++) { e[a] = in[2 * a]; o[a] = in[2 * a + 1]; }
    dft7_inv(e, E);
    dft7_inv(o, O);
    // W^m = e^{+2 pi i m/14}
    const float wr[7] = {1.0f, 0.90096886790241913f, 0.62348980185873353f, 0.22252093395631440f,
                         -0.22252093395631440f, -0.62348980185873353f, -0.90096886790241913f};
    const float wi[7] = {0.0f, 0.43388373911755812f, 0.78183148246802981f, 0.97492791218182361f,
                         0.97492791218182361f, 0.78183148246802981f, 0.43388373911755812f};
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int m = 0; m < 7; m++) {
        const cpx t = {O[m].re * wr[m] - O[m].im * wi[m], O[m].re * wi[m] + O[m].im * wr[m]};
        out[m] = {E[m].re + t.re, E[m].im + t.im};
        out[m + 7] = {E[m].re - t.re, E[m].im - t.im};
    }
}

constexpr int IFFT_SLOT_F2 = 210;   // float2 per frame slot: 14 rows x 15 (padded) >= 197 inputs, 196 outputs

// thread k1 (0..13): gather its 14 packed inputs Z[k1 + 14*k2] from the frame's spectrum X[0..196] (re,im pairs)
MIO_HD void ifft_load_z(const float * X, int k1, const float * tw392, cpx * z) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int k2 = 0; k2 < 14; k2++) {
        const int k = k1 + 14 * k2;
        cpx a = *(const cpx *) (X + 2 * k);
        cpx b = *(const cpx *) (X + 2 * (196 - k));
        b.im = -b.im;                                           // conj(X[196-k])
        if (k == 0) { a.im = 0.f; b.im = 0.f; }                 // Im X0, Im X196 ignored by the reference
        const cpx xe = {a.re + b.re, a.im + b.im};
        const cpx d = {a.re - b.re, a.im - b.im};
        const cpx w = *(const cpx *) (tw392 + 2 * k);
        const cpx xo = cmul(d, w);
        z[k2] = {xe.re - xo.im, xe.im + xo.re};
    }
}

// thread k1: y[m2] (result of the first DFT14) * e^{+2 pi i k1 m2/196} -> slot[k1*15 + m2]
MIO_HD void ifft_store_y(const cpx * y, int k1, const float * tw196, float * slot) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int m2 = 0; m2 < 14; m2++) {
        const int t = k1 * m2;
        const cpx w = *(const cpx *) (tw196 + 2 * t);
        *(cpx *) (slot + 2 * (k1 * 15 + m2)) = cmul(y[m2], w);
    }
}

// thread m2: read column m2
MIO_HD void ifft_load_y(const float * slot, int m2, cpx * y) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int k1 = 0; k1 < 14; k1++) y[k1] = *(const cpx *) (slot + 2 * (k1 * 15 + m2));
}

// thread m2: z[m1] = x[2(14 m1 + m2)] + i x[2(14 m1 + m2)+1] (unnormalised). Store x/392 * hann in place.
// The reference scales by 1/n_fft first (src/istft.cpp:64) and multiplies by the window in the overlap-add loop
// (src/istft.cpp:90), the same two roundings as here.
MIO_HD void ifft_store_time(const cpx * z, int m2, const float * hann, float inv_n, float * slot) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int m1 = 0; m1 < 14; m1++) {
        const int n = 2 * (14 * m1 + m2);
        const cpx h = *(const cpx *) (hann + n);
        const cpx o = {(z[m1].re * inv_n) * h.re, (z[m1].im * inv_n) * h.im};
        *(cpx *) (slot + n) = o;
    }
}

}  // namespace mio
