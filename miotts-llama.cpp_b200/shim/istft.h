// Drop-in replacement of the reference's src/istft.h (src/istft.h:1-42): the istft_cache class keeps its
// constructor and its eight getters (tables are built on the host with the reference's f32 formulae so code
// that inspects them sees the same numbers), while istft() itself runs the fused FFT + window + overlap-add +
// envelope-normalisation kernel on the B200 through mio_istft (include/miocodec_b200.h).
#pragma once

#include <vector>

class istft_cache {
public:
    istft_cache(int n_fft, int win_length);

    int n_fft() const;
    int win_length() const;
    int n_freq() const;                                // n_fft/2 + 1
    int n_mid() const;                                 // bins strictly between DC and Nyquist
    const std::vector<float> & cos_table() const;      // [n_fft][n_mid]
    const std::vector<float> & sin_table() const;      // [n_fft][n_mid]
    const std::vector<float> & nyquist_sign() const;   // [n_fft]  (+1, -1, +1, ...)
    const std::vector<float> & hann_window() const;    // [win_length] periodic Hann

private:
    int n_fft_ = 0;
    int win_length_ = 0;
    int n_freq_ = 0;
    int n_mid_ = 0;
    std::vector<float> cos_table_;
    std::vector<float> sin_table_;
    std::vector<float> nyquist_sign_;
    std::vector<float> hann_window_;
};

// istft (src/istft.cpp:68-108): spec[frame*n_freq*2 + bin*2 + {0,1}] -> hop_length*n_frames samples
// ((win_length - hop_length)/2 trimmed from each end).  Empty vector on failure; the device kernels are built for
// the reference's fixed geometry n_fft = win_length = 392, hop = 98 and any other geometry fails loudly.
std::vector<float> istft(const float * spec, int n_frames, int hop_length, const istft_cache & cache);
