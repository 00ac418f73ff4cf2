// lora_phy/LoRaDetector.hpp -- B200 build of the `lora_phy` API: the FFT peak detector as a device call.
//
// Same constructor and methods as the reference's class (include/lora_phy/LoRaDetector.hpp:16-82): the caller owns
// fft_in / fft_out, feed() stores one sample, detect() transforms the window and returns the first strict maximum of
// |X|^2 with power, powerAvg (dB, minus 20*log10(N)) and the parabolic fIndex.  detect() forwards to
// lora_b200_fft_detect_host (csrc/fft_kernels.cu): the spectrum lands in fft_out (or the buffer passed in) exactly as
// the reference leaves it there, and every returned value carries the CPU's bits.  Inside the receive kernels the
// same detector is fused after the FFT (csrc/window_common.cuh); this class exists for callers that use it directly.
#pragma once

#include <complex>
#include <cstddef>

#include <lora_phy/kissfft.hh>

template <typename Type>
class LoRaDetector {
public:
    LoRaDetector(const std::size_t N, std::complex<Type>* fft_in, std::complex<Type>* fft_out, kissfft<Type>& fft)
        : N(N), fft_in(fft_in), fft_out(fft_out), _fft(fft)
    {
    }

    //! feed simply sets an input sample (reference :33-36)
    void feed(const std::size_t i, const std::complex<Type>& samp) { fft_in[i] = samp; }

    //! argmax(abs(fft(input))) (reference :39-74)
    std::size_t detect(Type& power, Type& powerAvg, Type& fIndex, std::complex<Type>* fftOutput = nullptr)
    {
        static_assert(sizeof(Type) == sizeof(float) && Type(0.5) != Type(0), "this build detects on complex<float> only");
        if (fftOutput == nullptr) fftOutput = fft_out;
        std::size_t index = 0;
        float p = 0, pa = 0, fi = 0;
        const int rc = lora_phy::detail::device_detect(N, reinterpret_cast<const std::complex<float>*>(fft_in),
                                                       reinterpret_cast<std::complex<float>*>(fftOutput), &index, &p, &pa, &fi);
        if (rc) { index = 0; p = pa = fi = 0; }
        power = p; powerAvg = pa; fIndex = fi;
        return index;
    }

private:
    const std::size_t N;
    std::complex<Type>* fft_in;
    std::complex<Type>* fft_out;
    kissfft<Type>& _fft;
};
