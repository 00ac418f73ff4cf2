// fft_kernels.cu -- the reference's FFT and detector as stand-alone device entry points.
//
//   fft_kernel      kissfft<float>::transform (include/lora_phy/kissfft.hh:100-185), forward or inverse, one
//                   transform per CTA, spectrum written in natural bin order -- bit for bit the CPU's.
//   detect_kernel   LoRaDetector<float>::detect (include/lora_phy/LoRaDetector.hpp:39-74): FFT, first-strict-maximum
//                   argmax over |X|^2, power / powerAvg in dB, parabolic fIndex.
//
// These are the caller-facing forms of what the receive kernels fuse (demod_fast.cu never writes a spectrum); they
// back kissfft<T>::transform and LoRaDetector<T>::detect of include/lora_phy/*.hpp, which reference callers and
// tests use directly.  One window per CTA on the shared-memory schedule of faithful_fft.cuh (the same butterfly
// graph as the reference, every operation individually rounded), N/4 threads, all SF 2..12.  Not a throughput path.
//
// The inverse transform runs the same graph with the inverse plan's two differences (kissfft.hh:24-29,176-181):
// conjugated twiddles (sin(-x) = -sin(x) exactly, so the forward table conjugated IS the inverse table) and the radix-4
// rotation by +j.  (conj(FFT(conj x)) gives the same numbers but the wrong sign on exact zeros.)
#include "fft_kernels.h"

#include "demod_kernels.h" // g_launches
#include "exact_math.h"
#include "faithful_fft.cuh"

namespace lora_b200 {

namespace {

template <int SF>
__global__ void __launch_bounds__((FftShape<SF>::TPW > 32 ? FftShape<SF>::TPW : 32)) fft_detect_kernel(const FftArgs a)
{
    constexpr int N = 1 << SF;
    constexpr int TPW = FftShape<SF>::TPW;
    constexpr int CTA = TPW > 32 ? TPW : 32;
    __shared__ float2 F[N];
    __shared__ Peak red[32];
    const int t = threadIdx.x;
    const bool active = t < TPW;
    for (size_t w = blockIdx.x; w < a.n_windows; w += gridDim.x) {
        const float2* x = a.in + w * (size_t)N;
        for (int i = t; i < N; i += CTA) { // kissfft's leaf copy (kissfft.hh:114-120): digit-reversed scatter
            F[digit_reverse<SF>(i)] = __ldg(x + i);
        }
        __syncthreads();
        if (a.inverse) fft_passes<SF, true>(F, a.tw, t, active, [] { __syncthreads(); });
        else fft_passes<SF, false>(F, a.tw, t, active, [] { __syncthreads(); });
        if (a.out) {
            float2* y = a.out + w * (size_t)N;
            for (int i = t; i < N; i += CTA) y[i] = F[i];
        }
        if (a.det) { // LoRaDetector.hpp:43-71
            Peak best{0.0f, 0};
            for (int k = t; k < N; k += CTA) { // ascending k per thread: strict '>' keeps the first maximum
                const float2 b = F[k];
                const float m2 = __fadd_rn(__fmul_rn(b.x, b.x), __fmul_rn(b.y, b.y));
                if (m2 > best.v) { best.v = m2; best.i = k; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                Peak other;
                other.v = __shfl_xor_sync(0xffffffffu, best.v, o);
                other.i = __shfl_xor_sync(0xffffffffu, best.i, o);
                best = peak_merge(best, other);
            }
            if ((t & 31) == 0) red[t >> 5] = best;
            __syncthreads();
            if (t == 0) {
                for (int q = 1; q < CTA / 32; ++q) best = peak_merge(best, red[q]);
                // `total` is a double that receives the float |X|^2 of every bin IN BIN ORDER (:45-51).  When the terms
                // span more than ~2^18 the sum is inexact and its last bits depend on the order, so it is formed
                // sequentially here, exactly as the reference forms it.
                double total = 0.0;
                for (int k = 0; k < N; ++k) {
                    const float2 b = F[k];
                    total = __dadd_rn(total, (double)__fadd_rn(__fmul_rn(b.x, b.x), __fmul_rn(b.y, b.y)));
                }
                const float noise = __fsqrt_rn((float)__dsub_rn(total, (double)best.v));
                const float fundamental = __fsqrt_rn(best.v);
                DetectOut r;
                r.index = (unsigned)best.i;
                r.power_avg = __fsub_rn(__fmul_rn(20.0f, log10f(noise)), a.power_scale);
                r.power = __fsub_rn(__fmul_rn(20.0f, log10f(fundamental)), a.power_scale);
                const float2 Lb = F[best.i > 0 ? best.i - 1 : N - 1];
                const float2 Rb = F[best.i < N - 1 ? best.i + 1 : 0];
                const float left = xm_hypotf(Lb.x, Lb.y), right = xm_hypotf(Rb.x, Rb.y);
                const double demon = __dsub_rn(__dsub_rn(__dmul_rn(2.0, (double)fundamental), (double)right), (double)left);
                r.f_index = (demon == 0.0) ? 0.0f : (float)__ddiv_rn(__dmul_rn(0.5, (double)__fsub_rn(right, left)), demon);
                a.det[w] = r;
            }
        }
        __syncthreads(); // F and red are rewritten by the next window
    }
}

template <int SF>
cudaError_t launch_sf(const FftArgs& a, int sm_count, cudaStream_t st)
{
    constexpr int TPW = FftShape<SF>::TPW;
    constexpr int CTA = TPW > 32 ? TPW : 32;
    size_t grid = a.n_windows;
    const size_t cap = (size_t)sm_count * 8;
    if (grid > cap) grid = cap;
    if (grid == 0) return cudaSuccess;
    fft_detect_kernel<SF><<<(unsigned)grid, CTA, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

} // namespace

cudaError_t launch_fft_detect(int sf, const FftArgs& a, int sm_count, cudaStream_t st)
{
    switch (sf) {
    case 2: return launch_sf<2>(a, sm_count, st);
    case 3: return launch_sf<3>(a, sm_count, st);
    case 4: return launch_sf<4>(a, sm_count, st);
    case 5: return launch_sf<5>(a, sm_count, st);
    case 6: return launch_sf<6>(a, sm_count, st);
    case 7: return launch_sf<7>(a, sm_count, st);
    case 8: return launch_sf<8>(a, sm_count, st);
    case 9: return launch_sf<9>(a, sm_count, st);
    case 10: return launch_sf<10>(a, sm_count, st);
    case 11: return launch_sf<11>(a, sm_count, st);
    case 12: return launch_sf<12>(a, sm_count, st);
    default: return cudaErrorInvalidValue;
    }
}

} // namespace lora_b200
