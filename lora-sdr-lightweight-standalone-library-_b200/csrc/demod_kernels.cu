// demod_kernels.cu -- bit-faithful receive kernels: normalisation prepass, offset estimator, and the fused
// dechirp * CFO-rotate * window -> FFT -> argmax kernel, following src/phy/phy.cpp:78-239 and
// src/phy/LoRaDemod.cpp:49-195 operation for operation (see faithful_fft.cuh for why bit-faithful).
#include "demod_kernels.h"
#include "exact_math.h"
#include "faithful_fft.cuh"

namespace lora_b200 {

static __device__ __forceinline__ float pi_f() { return 3.14159274101257324f; } // float(M_PI)

// What one FFT input sample goes through before the transform, in the reference's order:
//   demodulate():        (iq * down[i]) * rot, then *= window[i]            phy.cpp:216-225
//   lora_demodulate():   (sample * scale) * rot, then *= window[i]          LoRaDemod.cpp:73-75,153-160
//   fused caller-dechirp: ((iq * down[i]) * scale) * rot, *= window[i]      performance_test.cpp:109-114 first
struct Prep {
    const float2* down;   // nullptr = no dechirp
    int down_ofs;         // table index of in-window sample 0, and the index mask (see below)
    int down_mask;
    const float* window;  // nullptr = no window multiply
    float scale;          // 1.0f = no normalisation (x * 1.0f is exact)
    bool rotate;
    float rate, start;
    __device__ __forceinline__ float2 operator()(float2 v, int i) const
    {
        // demodulate() multiplies in-window sample i by down[i] (phy.cpp:221).  The callers of
        // lora_demodulate() dechirp the whole buffer first, sample n by down[n mod N]
        // (performance_test.cpp:109-114), and only then does lora_demodulate shift its windows by t_off
        // (LoRaDemod.cpp:141-151) -- so in fused caller-dechirp mode the table index follows the sample's
        // ABSOLUTE position: down_ofs = window base mod N.
        if (down) v = cmul_rn(v, __ldg(down + ((down_ofs + i) & down_mask)));
        v = make_float2(__fmul_rn(v.x, scale), __fmul_rn(v.y, scale));
        if (rotate) {
            const float ph = __fadd_rn(start, __fmul_rn(rate, (float)i));
            float sn, cs;
            xm_sincosf(ph, &sn, &cs); // glibc cosf/sinf bits, exact_math.h
            v = cmul_rn(v, make_float2(cs, sn));
        }
        if (window) {
            const float w = __ldg(window + i);
            v = make_float2(__fmul_rn(v.x, w), __fmul_rn(v.y, w));
        }
        return v;
    }
};

// Group-wide argmax over the natural-order spectrum in F.  Every thread of the CTA must call it.
// `red` is shared scratch of (CTA/32) entries.  Returns the window's peak in every thread of the window
// for TPW <= 32, and in thread t == 0 otherwise.
template <int SF>
__device__ __forceinline__ Peak window_argmax(const float2* F, int t, int win_in_cta, Peak* red)
{
    constexpr int N = 1 << SF;
    constexpr int TPW = FftShape<SF>::TPW;
    Peak best{0.0f, 0};
#pragma unroll 4
    for (int k = t; k < N; k += TPW) { // ascending k per thread: strict '>' keeps the first maximum
        const float2 b = F[k];
        const float mag2 = __fadd_rn(__fmul_rn(b.x, b.x), __fmul_rn(b.y, b.y));
        if (mag2 > best.v) { best.v = mag2; best.i = k; }
    }
    constexpr int W = TPW < 32 ? TPW : 32;
#pragma unroll
    for (int o = W / 2; o > 0; o >>= 1) {
        Peak other;
        other.v = __shfl_xor_sync(0xffffffffu, best.v, o);
        other.i = __shfl_xor_sync(0xffffffffu, best.i, o);
        best = peak_merge(best, other);
    }
    if (TPW > 32) {
        constexpr int WARPS = TPW / 32;
        const int warp_in_win = t >> 5;
        __syncthreads();
        if ((t & 31) == 0) red[win_in_cta * WARPS + warp_in_win] = best;
        __syncthreads();
        if (t == 0) {
            best = red[win_in_cta * WARPS];
            for (int w = 1; w < WARPS; ++w) best = peak_merge(best, red[win_in_cta * WARPS + w]);
        }
    }
    return best;
}

// ---- normalisation prepass (LoRaDemod.cpp:59-77) --------------------------------------------------------
// One CTA per packet: max over all samples of max(|re|,|im|), with the reference's comparisons so a NaN
// never becomes the maximum.  In fused-dechirp mode the values examined are iq * down[n mod N].
__global__ void __launch_bounds__(256) maxabs_kernel(MaxAbsArgs a)
{
    const int p = blockIdx.x;
    const float2* x = a.iq + (size_t)p * a.packet_stride;
    float mx = 0.0f;
    for (size_t n = threadIdx.x; n < a.count; n += blockDim.x) {
        float2 v = __ldg(x + n);
        if (a.down) v = cmul_rn(v, __ldg(a.down + (n & (size_t)(a.N - 1))));
        const float r = fabsf(v.x), im = fabsf(v.y);
        const float m = (r < im) ? im : r; // std::max(r, im)
        if (m > mx) mx = m;
    }
    __shared__ float red[8];
    for (int o = 16; o > 0; o >>= 1) { const float y = __shfl_xor_sync(0xffffffffu, mx, o); if (y > mx) mx = y; }
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) if (red[w] > mx) mx = red[w];
        float4 e = a.est[p];
        e.w = mx;
        e.z = (mx > 1.0f) ? __fdiv_rn(1.0f, mx) : 1.0f;
        a.est[p] = e;
    }
}

// ---- offset estimator (phy.cpp:78-145 and its twin LoRaDemod.cpp:79-135) -------------------------------------
// One CTA per packet; the S = min(total, 2) symbols x osr phases are transformed one after another and
// thread 0 carries the scalar bookkeeping exactly as the reference's loop does.
template <int SF>
__global__ void __launch_bounds__((FftShape<SF>::TPW > 32 ? FftShape<SF>::TPW : 32))
estimate_kernel(EstimateArgs a)
{
    constexpr int N = 1 << SF;
    constexpr int TPW = FftShape<SF>::TPW;
    __shared__ float2 F[N];
    __shared__ Peak red[32];
    const int p = blockIdx.x;
    const int t = threadIdx.x;
    const bool active = t < TPW;
    const float2* x = a.iq + (size_t)p * a.packet_stride;
    float4 e = a.est[p];
    Prep prep{a.dechirp ? a.down : nullptr, 0, N - 1, a.window, a.use_scale ? e.z : 1.0f, false, 0.0f, 0.0f};

    float sum_index = 0.0f, phase_diff = 0.0f, prev_phase = 0.0f;
    bool have_prev = false;
    unsigned sum_t = 0;
    const size_t step = (size_t)N * a.osr;
    for (int s = 0; s < a.symbols; ++s) {
        float best_p = -1e30f, best_f = 0.0f;
        int best_idx = 0;
        unsigned best_t = 0;
        float2 best_bin = make_float2(0.0f, 0.0f);
        for (unsigned ph = 0; ph < (unsigned)a.osr; ++ph) {
            if (active)
                for (int i = t; i < N; i += TPW)
                    F[digit_reverse<SF>(i)] = prep(__ldg(x + s * step + ph + (size_t)i * a.osr), i);
            __syncthreads();
            fft_passes<SF>(F, a.tw, t, active, [] { __syncthreads(); });
            // threads >= TPW (they exist only for SF < 7, to fill a warp) scan thread 0's bins again: reads
            // only, and merging a duplicate (value, index) changes nothing
            const Peak pk = window_argmax<SF>(F, active ? t : 0, 0, red);
            if (t == 0) { // LoRaDetector.hpp:60-71
                const float fundamental = __fsqrt_rn(pk.v);
                const float pw = __fsub_rn(__fmul_rn(20.0f, log10f(fundamental)), a.power_scale);
                const float2 L = F[pk.i > 0 ? pk.i - 1 : N - 1];
                const float2 R = F[pk.i < N - 1 ? pk.i + 1 : 0];
                const float left = xm_hypotf(L.x, L.y), right = xm_hypotf(R.x, R.y);
                const double demon = __dsub_rn(__dsub_rn(__dmul_rn(2.0, (double)fundamental), (double)right), (double)left);
                const float fi = (demon == 0.0) ? 0.0f
                                                : (float)__ddiv_rn(__dmul_rn(0.5, (double)__fsub_rn(right, left)), demon);
                if (pw > best_p || (a.legacy_tie && pw == best_p && pk.i < best_idx)) {
                    best_p = pw; best_idx = pk.i; best_f = fi; best_t = ph; best_bin = F[pk.i];
                }
            }
            __syncthreads();
        }
        if (t == 0) {
            sum_t += best_t;
            sum_index = __fadd_rn(sum_index, __fadd_rn((float)best_idx, best_f));
            const float phase = xm_atan2f(best_bin.y, best_bin.x); // std::arg
            if (have_prev) {
                float d = __fsub_rn(phase, prev_phase);
                const float two_pi = __fmul_rn(2.0f, pi_f());
                while (d > pi_f()) d = __fsub_rn(d, two_pi);
                while (d < -pi_f()) d = __fadd_rn(d, two_pi);
                phase_diff = __fadd_rn(phase_diff, d);
            }
            prev_phase = phase;
            have_prev = true;
        }
    }
    if (t == 0) {
        const float S = (float)a.symbols;
        const float avg_index = __fdiv_rn(sum_index, S);
        const float cfo_coarse = __fdiv_rn(avg_index, (float)N);
        float cfo_fine = 0.0f;
        if (a.symbols > 1)
            cfo_fine = __fdiv_rn(__fdiv_rn(phase_diff, (float)(a.symbols - 1)),
                                 __fmul_rn(__fmul_rn(2.0f, pi_f()), (float)N));
        e.x = __fadd_rn(cfo_coarse, cfo_fine);
        const float frac = __fsub_rn(avg_index, floorf(__fadd_rn(avg_index, 0.5f)));
        const float avg_t = __fdiv_rn((float)sum_t, S);
        e.y = __fsub_rn(avg_t, __fmul_rn(__fmul_rn(frac, (float)N), (float)a.osr));
        a.est[p] = e;
    }
}

// ---- the hot kernel: dechirp * rotate * window -> FFT -> argmax (phy.cpp:200-234, LoRaDemod.cpp:141-175) ----
// A CTA holds WPC windows in shared memory, TPW threads each; CTAs stride over the window list.
template <int SF>
struct DemodShape {
    static constexpr int TPW = FftShape<SF>::TPW;
    static constexpr int CTA = TPW > 128 ? TPW : 128;
    static constexpr int WPC = CTA / TPW;
};

template <int SF>
__global__ void __launch_bounds__(DemodShape<SF>::CTA) demod_kernel(DemodArgs a)
{
    constexpr int N = 1 << SF;
    constexpr int TPW = DemodShape<SF>::TPW;
    constexpr int WPC = DemodShape<SF>::WPC;
    __shared__ float2 smem[WPC * N];
    __shared__ Peak red[32];
    const int win_in_cta = threadIdx.x / TPW;
    const int t = threadIdx.x % TPW;
    float2* F = smem + win_in_cta * N;
    const size_t step = (size_t)N * a.osr;
    const long long groups = (a.n_windows + WPC - 1) / WPC;

    for (long long g = blockIdx.x; g < groups; g += gridDim.x) {
        const long long w = g * WPC + win_in_cta;
        const bool active = w < a.n_windows;
        int p = 0, s = 0;
        if (active) {
            p = (int)(w / a.total);
            s = (int)(w - (long long)p * a.total);
            const float4 e = a.est[p];
            const int t_off = (int)roundf(e.y);                                   // phy.cpp:197
            const float rate = __fdiv_rn(__fmul_rn(-2.0f * 3.14159274101257324f, e.x), (float)N); // :198
            size_t base = (size_t)s * step;                                       // :205-212
            if (t_off > 0) {
                if (base + (size_t)t_off + step <= a.count) base += (size_t)t_off;
            } else if (t_off < 0) {
                const size_t off = (size_t)(-(long long)t_off);
                if (off <= base) base -= off;
            }
            const float start = __fmul_rn(rate, __fadd_rn((float)((size_t)s * N),
                                                          __fdiv_rn((float)t_off, (float)a.osr))); // :213-214
            const int down_ofs = a.dechirp_absolute ? (int)(base & (size_t)(N - 1)) : 0;
            const Prep prep{a.dechirp ? a.down : nullptr, down_ofs, N - 1, a.window, a.use_scale ? e.z : 1.0f, true, rate, start};
            const float2* x = a.iq + (size_t)p * a.packet_stride + base;
#pragma unroll 4
            for (int i = t; i < N; i += TPW) F[digit_reverse<SF>(i)] = prep(__ldg(x + (size_t)i * a.osr), i);
        }
        __syncthreads();
        fft_passes<SF>(F, a.tw, t, active, [] { __syncthreads(); });
        const Peak pk = window_argmax<SF>(F, t, win_in_cta, red);
        if (active && t == 0) {
            if (!a.have_sync) a.symbols[(size_t)p * a.symbol_stride + s] = (uint16_t)pk.i;
            else if (s >= 2) a.symbols[(size_t)p * a.symbol_stride + (s - 2)] = (uint16_t)pk.i;
            else a.sync_raw[2 * (size_t)p + s] = (uint16_t)pk.i;
            if (a.peaks) a.peaks[w] = pk.v;
        }
        __syncthreads(); // F is rewritten by the next iteration's loads
    }
}

__global__ void finalize_kernel(FinalizeArgs a)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n_packets) return;
    const float4 e = a.est[p];
    lora_b200_pkt_metrics m;
    m.cfo = e.x;
    m.time_offset = e.y;
    m.max_amp = a.legacy ? e.w : 0.0f;
    m.normalised = (a.legacy && e.w > 1.0f) ? 1 : 0;
    m.n_symbols = (uint16_t)a.n_symbols;
    m.sync_word = 0;
    if (a.have_sync) { // phy.cpp:235-237, LoRaDemod.cpp:177-190
        const unsigned shift = a.sf > 4 ? a.sf - 4 : 0;
        const unsigned hi = ((unsigned)a.sync_raw[2 * p] >> shift) & 0x0f;
        const unsigned lo = ((unsigned)a.sync_raw[2 * p + 1] >> shift) & 0x0f;
        m.sync_word = (uint8_t)((hi << 4) | lo);
    }
    a.metrics[p] = m;
}

__global__ void fill_est_kernel(float4* est, int n)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) est[p] = make_float4(0.0f, 0.0f, 1.0f, 0.0f);
}

std::atomic<unsigned long long> g_launches{0};

// ---- compensate_offsets (phy.cpp:147-176) ---------------------------------------------------------------------------
// The reference rotates in place and then moves the samples; out[n] is therefore rotated(in[n - offset]) where that
// index exists and zero elsewhere.  cos/sin carry glibc's bits (exact_math.h), the product is std::complex's.
__global__ void compensate_kernel(CompensateArgs a)
{
    const size_t n = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= a.count) return;
    const long long m = (long long)n - a.offset;
    float2 v = make_float2(0.0f, 0.0f);
    if (m >= 0 && (size_t)m < a.count) {
        const float ph = __fmul_rn(a.rate, (float)(size_t)m);
        float sn, cs;
        xm_sincosf(ph, &sn, &cs);
        v = cmul_rn(a.in[m], make_float2(cs, sn));
    }
    a.out[n] = v;
}

cudaError_t launch_compensate(const CompensateArgs& a, cudaStream_t st)
{
    if (a.count == 0) return cudaSuccess;
    compensate_kernel<<<(unsigned)((a.count + 255) / 256), 256, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

// ---- launchers ---------------------------------------------------------------------------------------------

cudaError_t launch_fill_est(float4* est, int n, cudaStream_t st)
{
    fill_est_kernel<<<(n + 255) / 256, 256, 0, st>>>(est, n);
    ++g_launches;
    return cudaGetLastError();
}

cudaError_t launch_maxabs(const MaxAbsArgs& a, int n_packets, cudaStream_t st)
{
    maxabs_kernel<<<n_packets, 256, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

template <int SF>
static cudaError_t launch_estimate_sf(const EstimateArgs& a, int n_packets, cudaStream_t st)
{
    constexpr int T = FftShape<SF>::TPW > 32 ? FftShape<SF>::TPW : 32;
    estimate_kernel<SF><<<n_packets, T, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

template <int SF>
static cudaError_t launch_demod_sf(const DemodArgs& a, int sm_count, cudaStream_t st)
{
    constexpr int WPC = DemodShape<SF>::WPC;
    const long long groups = (a.n_windows + WPC - 1) / WPC;
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, demod_kernel<SF>, DemodShape<SF>::CTA, 0);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    long long grid = (long long)sm_count * per_sm; // whole waves of resident CTAs, striding over the windows
    if (grid > groups) grid = groups;
    if (grid < 1) grid = 1;
    demod_kernel<SF><<<(unsigned)grid, DemodShape<SF>::CTA, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

#define LORA_SF_SWITCH(fn, ...)                                   \
    switch (sf) {                                                 \
    case 2: return fn<2>(__VA_ARGS__);                            \
    case 3: return fn<3>(__VA_ARGS__);                            \
    case 4: return fn<4>(__VA_ARGS__);                            \
    case 5: return fn<5>(__VA_ARGS__);                            \
    case 6: return fn<6>(__VA_ARGS__);                            \
    case 7: return fn<7>(__VA_ARGS__);                            \
    case 8: return fn<8>(__VA_ARGS__);                            \
    case 9: return fn<9>(__VA_ARGS__);                            \
    case 10: return fn<10>(__VA_ARGS__);                          \
    case 11: return fn<11>(__VA_ARGS__);                          \
    case 12: return fn<12>(__VA_ARGS__);                          \
    default: return cudaErrorInvalidValue;                        \
    }

cudaError_t launch_estimate(int sf, const EstimateArgs& a, int n_packets, cudaStream_t st)
{
    LORA_SF_SWITCH(launch_estimate_sf, a, n_packets, st)
}

cudaError_t launch_demod(int sf, const DemodArgs& a, int sm_count, cudaStream_t st)
{
    LORA_SF_SWITCH(launch_demod_sf, a, sm_count, st)
}

cudaError_t launch_finalize(const FinalizeArgs& a, cudaStream_t st)
{
    finalize_kernel<<<(a.n_packets + 255) / 256, 256, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) convert_i16_kernel(const short2* __restrict__ in, float2* __restrict__ out, size_t n, float scale)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t n4 = ((reinterpret_cast<uintptr_t>(in) & 15u) == 0 && (reinterpret_cast<uintptr_t>(out) & 15u) == 0) ? n / 4 : 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) { // 4 samples: 16 bytes in, 32 out
        const int4 v = __ldg(reinterpret_cast<const int4*>(in) + i);
        const int w[4] = {v.x, v.y, v.z, v.w};
        float4 o[2];
        float* f = reinterpret_cast<float*>(o);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            f[2 * k] = __fmul_rn((float)(short)(w[k] & 0xffff), scale);
            f[2 * k + 1] = __fmul_rn((float)(short)(w[k] >> 16), scale);
        }
        reinterpret_cast<float4*>(out)[2 * i] = o[0];
        reinterpret_cast<float4*>(out)[2 * i + 1] = o[1];
    }
    for (size_t i = 4 * n4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const short2 v = in[i];
        out[i] = make_float2(__fmul_rn((float)v.x, scale), __fmul_rn((float)v.y, scale));
    }
}

cudaError_t launch_convert_i16(const short2* in, float2* out, size_t n, float scale, int sm_count, cudaStream_t st)
{
    size_t blocks = (n / 4 + 255) / 256;
    const size_t cap = (size_t)sm_count * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    convert_i16_kernel<<<(unsigned)blocks, 256, 0, st>>>(in, out, n, scale);
    ++g_launches;
    return cudaGetLastError();
}

// int8 I/Q (RTL-SDR class front ends, or a gateway that trades resolution for link bytes): 2 bytes per sample over PCIe.
// The conversion is exact (an 8-bit integer is a float; one rounding in the multiply), so the results equal the float
// path run on the converted samples.
__global__ void __launch_bounds__(256) convert_i8_kernel(const char2* __restrict__ in, float2* __restrict__ out, size_t n, float scale)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t n8 = ((reinterpret_cast<uintptr_t>(in) & 15u) == 0 && (reinterpret_cast<uintptr_t>(out) & 15u) == 0) ? n / 8 : 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) { // 8 samples: 16 bytes in, 64 out
        const int4 v = __ldg(reinterpret_cast<const int4*>(in) + i);
        const int w[4] = {v.x, v.y, v.z, v.w};
        float4 o[4];
        float* f = reinterpret_cast<float*>(o);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int b = 0; b < 4; ++b) f[4 * k + b] = __fmul_rn((float)(signed char)((w[k] >> (8 * b)) & 0xff), scale);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) reinterpret_cast<float4*>(out)[4 * i + k] = o[k];
    }
    for (size_t i = 8 * n8 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const char2 v = in[i];
        out[i] = make_float2(__fmul_rn((float)v.x, scale), __fmul_rn((float)v.y, scale));
    }
}

cudaError_t launch_convert_i8(const char2* in, float2* out, size_t n, float scale, int sm_count, cudaStream_t st)
{
    size_t blocks = (n / 8 + 255) / 256;
    const size_t cap = (size_t)sm_count * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    convert_i8_kernel<<<(unsigned)blocks, 256, 0, st>>>(in, out, n, scale);
    ++g_launches;
    return cudaGetLastError();
}

} // namespace lora_b200
