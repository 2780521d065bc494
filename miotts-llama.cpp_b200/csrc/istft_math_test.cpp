// CPU unit test of csrc/istft_math.h: emulates the 14-thread cooperation of the CUDA kernel with the very same
// host/device functions and compares against the reference's irfft formula (src/istft.cpp:43-66) in double.
// Build: g++ -O2 -std=c++17 istft_math_test.cpp -o istft_math_test ; prints max-abs error, exit 0 if < 2e-6.
#include "istft_math.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
using namespace mio;
int main(int argc, char ** argv) {
    const int N = 392, NF = 197;
    unsigned seed = argc > 1 ? (unsigned) atoi(argv[1]) : 1u;
    srand(seed);
    std::vector<float> tw392(2 * 196), tw196(2 * 196), hann(N, 1.0f);
    for (int k = 0; k < 196; k++) {
        tw392[2 * k] = (float) cos(2.0 * M_PI * k / 392.0); tw392[2 * k + 1] = (float) sin(2.0 * M_PI * k / 392.0);
        tw196[2 * k] = (float) cos(2.0 * M_PI * k / 196.0); tw196[2 * k + 1] = (float) sin(2.0 * M_PI * k / 196.0);
    }
    double worst = 0.0, peak = 0.0;
    for (int trial = 0; trial < 20; trial++) {
        std::vector<float> X(2 * NF);
        for (auto & v : X) v = (float) ((rand() / (double) RAND_MAX - 0.5) * 4.0);
        std::vector<double> ref(N);
        for (int n = 0; n < N; n++) {
            double s = X[0] + ((n & 1) ? -1.0 : 1.0) * X[2 * 196];
            for (int k = 1; k <= 195; k++) s += 2.0 * (X[2 * k] * cos(2.0 * M_PI * k * n / N) - X[2 * k + 1] * sin(2.0 * M_PI * k * n / N));
            ref[n] = s / N;
        }
        std::vector<float> slot(2 * IFFT_SLOT_F2, 0.f);
        for (int i = 0; i < 2 * NF; i++) slot[i] = X[i];
        cpx regs[14][14], y[14];
        for (int k1 = 0; k1 < 14; k1++) ifft_load_z(slot.data(), k1, tw392.data(), regs[k1]);
        for (int k1 = 0; k1 < 14; k1++) { dft14_inv(regs[k1], y); ifft_store_y(y, k1, tw196.data(), slot.data()); }
        for (int m2 = 0; m2 < 14; m2++) ifft_load_y(slot.data(), m2, regs[m2]);
        for (int m2 = 0; m2 < 14; m2++) { dft14_inv(regs[m2], y); ifft_store_time(y, m2, hann.data(), 1.0f / N, slot.data()); }
        for (int n = 0; n < N; n++) { worst = fmax(worst, fabs(slot[n] - ref[n])); peak = fmax(peak, fabs(ref[n])); }
    }
    printf("max_abs_err %.3e peak %.3f\n", worst, peak);
    return worst < 2e-6 * fmax(1.0, peak) ? 0 : 1;
}
