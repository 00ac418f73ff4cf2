// lora_phy/kissfft.hh -- B200 build of the `lora_phy` API: the FFT plan and transform as device calls.
//
// The reference embeds two ~33 KB kissfft plans (twiddle table + radix list, reference
// include/lora_phy/kissfft.hh:44-55) in every workspace and runs the transform on the caller's core.  In this build
// the tables live in device memory inside a shared lora_b200_ctx and a "plan" is the transform length -- which
// phy.cpp:13-18 of the reference also reads back to recover sf -- plus that context handle.
// `kissfft<float>::transform` keeps the reference's signature (kissfft.hh:100-103) and forwards to the stand-alone
// FFT kernel (csrc/fft_kernels.cu, lora_b200_fft_detect_host): same butterfly graph, same twiddle bits, bit-identical
// output.  There is no host FFT in this build: without a CUDA device, or for a length that is not 2^2..2^12
// (the reference's own static limit is 4096, kissfft.hh:34), the output is zero-filled and last_error() is non-zero.
#pragma once

#include <complex>
#include <cstddef>

namespace kissfft_utils {
// hard limit of the reference's static plans (reference kissfft.hh:34); SF12 is the largest LoRa symbol
static const std::size_t KISSFFT_MAX_N = 4096;
} // namespace kissfft_utils

namespace lora_phy {
namespace detail {
// liblora_phy.so (host/phy.cpp): transform / detect one window of nfft = 2^sf samples on the device.  0 on success.
int device_fft(std::size_t nfft, bool inverse, const std::complex<float>* src, std::complex<float>* dst);
int device_detect(std::size_t nfft, const std::complex<float>* src, std::complex<float>* spectrum, std::size_t* index,
                  float* power, float* power_avg, float* f_index);
} // namespace detail
} // namespace lora_phy

template <typename T>
struct kissfft_plan {
    std::size_t nfft = 0;  // transform length, 1 << sf
    bool inverse = false;
    void* device = nullptr; // lora_b200_ctx shared by every workspace with the same (sf, bw, osr, window, GPU)
};

template <typename T>
class kissfft {
public:
    typedef std::complex<T> cpx_type;
    typedef kissfft_plan<T> plan_type;

    explicit kissfft(plan_type& plan) : _p(plan), _err(0) {}

    // reference kissfft.hh:71-98: twiddles and the radix list belong to the device context; only the shape is kept here
    static void init(plan_type& plan, int nfft, bool inverse)
    {
        plan.nfft = nfft > 0 ? static_cast<std::size_t>(nfft) : 0;
        plan.inverse = inverse;
    }

    // reference kissfft.hh:100-103: unnormalised DFT of src into dst (out of place), plan.nfft points
    void transform(const cpx_type* src, cpx_type* dst) const
    {
        static_assert(sizeof(T) == sizeof(float) && T(0.5) != T(0), "this build transforms complex<float> only");
        _err = lora_phy::detail::device_fft(_p.nfft, _p.inverse, reinterpret_cast<const std::complex<float>*>(src),
                                            reinterpret_cast<std::complex<float>*>(dst));
        if (_err)
            for (std::size_t i = 0; i < _p.nfft; ++i) dst[i] = cpx_type();
    }

    int last_error() const { return _err; }
    const plan_type& plan() const { return _p; }

private:
    plan_type& _p;
    mutable int _err;
};
