// faithful_fft.cuh -- the reference's FFT + peak search as a GPU schedule of the SAME butterfly graph.
//
// Why "faithful": the reference picks symbols with a strict-'>' argmax over float32 |X[k]|^2
// (include/lora_phy/LoRaDetector.hpp:46-58).  On noisy windows the top two bins are routinely within 1e-6
// relative of each other, so an FFT that is merely accurate flips symbols.  kissfft for
// N = 2^SF is a decimation-in-time recursion with radix list 4,4,...[,2] (kissfft.hh:80-97); every
// butterfly (kf_bfly4 :164-185, kf_bfly2 :155-162) is a pure function of its operands, so any schedule
// that performs the same butterflies on the same operands with un-fused IEEE float ops yields the same
// bits.  Here: mixed-radix digit-reversed scatter into shared memory (the recursion's m==1 leaf,
// :114-120), then one pass per stage from the innermost outwards, N/4 (or N/2) butterflies in parallel.
//
// All arithmetic goes through __fmul_rn/__fadd_rn/__fsub_rn so nvcc cannot contract it into FMAs.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace lora_b200 {

// std::complex<float> operator* as g++ -O2 emits it for finite operands: (ac - bd, ad + bc).
__device__ __forceinline__ float2 cmul_rn(float2 a, float2 b)
{
    return make_float2(__fsub_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)),
                       __fadd_rn(__fmul_rn(a.x, b.y), __fmul_rn(a.y, b.x)));
}
__device__ __forceinline__ float2 cadd_rn(float2 a, float2 b) { return make_float2(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y)); }
__device__ __forceinline__ float2 csub_rn(float2 a, float2 b) { return make_float2(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y)); }

template <int SF>
struct FftShape {
    static constexpr int N = 1 << SF;
    static constexpr int R4 = SF / 2;          // number of radix-4 stages (outermost first)
    static constexpr bool HAS_R2 = (SF & 1);   // a trailing radix-2 stage when SF is odd
    // threads cooperating on one window: one radix-4 butterfly each
    static constexpr int TPW = (N / 4) > 0 ? (N / 4) : 1;
};

// Output position of input index i: digits q_s of i in the mixed radix (F_s = 4^s) land at weight
// m_s = N / (r_0..r_s).  For odd SF the top bit of i is the radix-2 digit and lands at weight 1.
template <int SF>
__device__ __forceinline__ int digit_reverse(int i)
{
    constexpr int N = 1 << SF;
    int pos = 0;
#pragma unroll
    for (int s = 0; s < SF / 2; ++s) pos += ((i >> (2 * s)) & 3) * (N >> (2 * s + 2));
    if (SF & 1) pos += (i >> (SF - 1)) & 1;
    return pos;
}

// Runs all butterfly passes on one window held in shared memory at `F` (digit-reversed order on entry,
// natural bin order on exit).  `t` is this thread's index among the TPW threads of the window; `sync()`
// is a barrier every calling thread reaches.  Threads with active == false do no butterflies (the passes
// are in place, so a second thread repeating a butterfly would corrupt it) but still take every barrier.  `tw` is the N-entry twiddle table exp(-2*pi*j*i/N) built on the
// host the way kissfft.hh:24-29 builds it.
// INV = the plan built with inverse = true (kissfft.hh:24-29,176-181): conjugated twiddles (sin(-x) = -sin(x) exactly, so
// conjugating the forward table gives the inverse table's bits) and the radix-4 rotation by +j instead of -j.
template <int SF, bool INV = false, typename Sync>
__device__ __forceinline__ void fft_passes(float2* __restrict__ F, const float2* __restrict__ tw, int t, bool active, Sync sync)
{
    auto twd = [&](int i) { float2 w = tw[i]; if (INV) w.y = -w.y; return w; };
    constexpr int N = 1 << SF;
    constexpr int TPW = FftShape<SF>::TPW;
    if (SF & 1) { // innermost radix-2 stage: m = 1, fstride = N/2, twiddle index k*fstride = 0
        const float2 w0 = twd(0);
#pragma unroll
        for (int j = active ? t : N; j < N / 2; j += TPW) {
            const float2 a = F[2 * j], b = F[2 * j + 1];
            const float2 x = cmul_rn(b, w0);
            F[2 * j + 1] = csub_rn(a, x);
            F[2 * j] = cadd_rn(a, x);
        }
        sync();
    }
#pragma unroll
    for (int s = SF / 2 - 1; s >= 0; --s) { // radix-4 stages, innermost first
        const int m = (N >> (2 * s + 2));   // stageRemainder
        const int fstride = 1 << (2 * s);   // product of the outer radices
        for (int j = active ? t : N; j < N / 4; j += TPW) {
            const int k = j & (m - 1);
            const int base = ((j - k) << 2) + k; // block*4m + k
            const float2 f0 = F[base];
            const float2 s0 = cmul_rn(F[base + m], twd(k * fstride));
            const float2 s1 = cmul_rn(F[base + 2 * m], twd(k * fstride * 2));
            const float2 s2 = cmul_rn(F[base + 3 * m], twd(k * fstride * 3));
            const float2 s5 = csub_rn(f0, s1);
            const float2 g0 = cadd_rn(f0, s1);
            const float2 s3 = cadd_rn(s0, s2);
            const float2 d = csub_rn(s0, s2);
            // forward transform: (im, -re), kissfft.hh:176-177; inverse: (-im, re), :179-180
            const float2 s4 = INV ? make_float2(-d.y, d.x) : make_float2(d.y, -d.x);
            F[base + 2 * m] = csub_rn(g0, s3);
            F[base] = cadd_rn(g0, s3);
            F[base + m] = cadd_rn(s5, s4);
            F[base + 3 * m] = csub_rn(s5, s4);
        }
        sync();
    }
}

// (value, index) pair ordering of LoRaDetector.hpp:46-58: larger |X|^2 wins, equal values keep the lower
// index, and a value that is not > 0 (zero, NaN) never wins -- such a lane reports (0, 0).
struct Peak {
    float v;
    int i;
};
__device__ __forceinline__ Peak peak_merge(Peak a, Peak b)
{
    if (b.v > a.v || (b.v == a.v && b.i < a.i)) return b;
    return a;
}

} // namespace lora_b200
