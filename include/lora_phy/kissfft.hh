// lora_phy/kissfft.hh -- B200 build of the `lora_phy` API: plan handle only.
//
// The reference embeds two ~33 KB kissfft plans (twiddle table + radix list, reference
// include/lora_phy/kissfft.hh:44-55) in every workspace and rebuilds them in init().  In this build the FFT runs on
// the GPU (csrc/fft_blocked.cuh), the tables live in device memory inside a shared lora_b200_ctx, and a "plan" is
// just the transform length -- which phy.cpp:13-18 of the reference also reads back to recover sf -- plus the
// context handle.  Host-side FFTs (kissfft<T>::transform) are not part of this build: the product has no CPU path.
#pragma once

#include <cstddef>

namespace kissfft_utils {
// hard limit of the reference's static plans (reference kissfft.hh:34); SF12 is the largest LoRa symbol
static const std::size_t KISSFFT_MAX_N = 4096;
} // namespace kissfft_utils

template <typename T>
struct kissfft_plan {
    std::size_t nfft = 0;  // transform length, 1 << sf
    bool inverse = false;
    void* device = nullptr; // lora_b200_ctx shared by every workspace with the same (sf, bw, osr, window, GPU)
};
