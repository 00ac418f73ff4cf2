// fft_kernels.h -- stand-alone FFT / detector launches (kissfft.hh:100-185, LoRaDetector.hpp:39-74).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace lora_b200 {

struct DetectOut { // == lora_b200_detect_result
    uint32_t index;
    float power, power_avg, f_index;
};

struct FftArgs {
    const float2* in;   // n_windows x N samples
    size_t n_windows;
    int inverse;        // kissfft plan built with inverse = true
    const float2* tw;   // N forward twiddles, host-built (kissfft.hh:24-29)
    float power_scale;  // float(20*log10(N)), LoRaDetector.hpp:29
    float2* out;        // optional: n_windows x N bins, natural order
    DetectOut* det;     // optional: one record per window
};

cudaError_t launch_fft_detect(int sf, const FftArgs& a, int sm_count, cudaStream_t st);

} // namespace lora_b200
