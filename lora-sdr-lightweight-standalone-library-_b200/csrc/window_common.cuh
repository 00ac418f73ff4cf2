// window_common.cuh -- pieces shared by every kernel built on the register-blocked FFT: the synchronisation
// of the threads that own one window, and the (value, index) peak search of LoRaDetector.hpp:46-58.
#pragma once

#include "fft_blocked.cuh"

namespace lora_b200 {
namespace {

constexpr float kPiF = 3.14159274101257324f; // float(M_PI)
constexpr int kCta = 256;

// ---- synchronisation among the threads that share one window -------------------------------------------------
// TPW threads (a power of two, aligned) own a window: a sub-warp group, a warp, several warps, or the whole CTA.
template <int TPW>
struct Group {
    static PM_DEV unsigned mask()
    {
        if (TPW >= 32) return 0xffffffffu;
        const unsigned lane = threadIdx.x & 31u;
        return ((1u << (TPW & 31)) - 1u) << (lane & ~(unsigned)(TPW - 1));
    }
    static PM_DEV void sync(int slot)
    {
        if (TPW <= 32) __syncwarp(mask());
        else if (TPW == kCta) __syncthreads();
        else asm volatile("bar.sync %0, %1;" ::"r"(slot + 1), "r"(TPW) : "memory");
    }
};

// ---- peak search ---------------------------------------------------------------------------------------------
// (value, index) ordering of LoRaDetector.hpp:46-58: larger |X|^2 wins, equal values keep the lower index, and
// a value that is not > 0 (zero, NaN) never wins.  `second` = the largest value that did not win (fast flavour).
struct Top {
    float v;
    int i;
    float second;
};

PM_DEV Top top_merge(Top a, Top b)
{
    Top r;
    const bool take_b = b.v > a.v || (b.v == a.v && b.i < a.i);
    r.v = take_b ? b.v : a.v;
    r.i = take_b ? b.i : a.i;
    r.second = fmaxf(fmaxf(a.second, b.second), take_b ? a.v : b.v);
    return r;
}

template <int SF, bool FAITH>
PM_DEV Top thread_top(const c64 (&v)[16], int t)
{
    using B = Blk<SF>;
    constexpr int P = B::NP - 1, Rp = B::R(P), Gp = B::G(P);
    Top best{0.0f, 0, 0.0f};
    // ascending bin order for this thread is e-major, g-minor: bin = t + T*(g + G*e); strict '>' keeps the first
#pragma unroll
    for (int e = 0; e < Rp; ++e) {
#pragma unroll
        for (int g = 0; g < Gp; ++g) {
            const int i = g * Rp + e;
            const float m2 = mag2<FAITH>(v[i]);
            if (!FAITH) best.second = fmaxf(best.second, fminf(best.v, m2));
            if (m2 > best.v) { best.v = m2; best.i = bin_of<SF>(t, i); }
        }
    }
    return best;
}

// Reduces over the TPW threads of the window; the result is valid in the window's thread 0 (all threads for TPW<=32).
template <int TPW>
PM_DEV Top group_top(Top best, int t, int slot, Top* red)
{
    constexpr int W = TPW < 32 ? TPW : 32;
    const unsigned mask = Group<TPW>::mask();
#pragma unroll
    for (int o = W / 2; o > 0; o >>= 1) {
        Top other;
        other.v = __shfl_xor_sync(mask, best.v, o);
        other.i = __shfl_xor_sync(mask, best.i, o);
        other.second = __shfl_xor_sync(mask, best.second, o);
        best = top_merge(best, other);
    }
    if (TPW > 32) {
        constexpr int WARPS = TPW / 32;
        if ((t & 31) == 0) red[slot * WARPS + (t >> 5)] = best;
        Group<TPW>::sync(slot);
        if (t == 0) {
            best = red[slot * WARPS];
#pragma unroll
            for (int w = 1; w < WARPS; ++w) best = top_merge(best, red[slot * WARPS + w]);
        }
    }
    return best;
}

// Peak of one window for the demod kernels.  For TPW >= 32 the warp-level part uses three REDUX operations on
// the value bits (|X|^2 >= 0, so unsigned order = float order) instead of a 5-round (value, index, second)
// shuffle butterfly, and the bin index is only searched by the thread(s) that hold the maximum.
template <int SF, bool FAITH, int TPW>
PM_DEV Top window_top(const c64 (&v)[16], int t, int slot, Top* red)
{
    if (TPW < 32) return group_top<TPW>(thread_top<SF, FAITH>(v, t), t, slot, red);
    using B = Blk<SF>;
    constexpr int P = B::NP - 1, Rp = B::R(P), Gp = B::G(P);
    float m2[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) m2[i] = mag2<FAITH>(v[i]);
    // largest (and, fast flavour, second largest counting duplicates) of the thread's 16 values; NaN is ignored
    float hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { hi[i] = fmaxf(m2[2 * i], m2[2 * i + 1]); lo[i] = fminf(m2[2 * i], m2[2 * i + 1]); }
#pragma unroll
    for (int n = 4; n >= 1; n >>= 1) {
#pragma unroll
        for (int i = 0; i < n; ++i) {
            if (!FAITH) lo[i] = fmaxf(fminf(hi[i], hi[i + n]), fmaxf(lo[i], lo[i + n]));
            hi[i] = fmaxf(hi[i], hi[i + n]);
        }
    }
    const float mine_v = fmaxf(hi[0], 0.0f); // all-NaN -> 0
    const unsigned full = 0xffffffffu;
    Top r;
    r.v = __uint_as_float(__reduce_max_sync(full, __float_as_uint(mine_v)));
    const bool mine = mine_v == r.v && r.v > 0.0f; // a value that is not > 0 never wins: index 0
    int cand = 0x7fffffff;
    if (mine) {
        // lowest bin holding the maximum: descending scan, the last hit is the lowest
#pragma unroll
        for (int e = Rp - 1; e >= 0; --e) {
#pragma unroll
            for (int g = Gp - 1; g >= 0; --g) {
                const int i = g * Rp + e;
                if (m2[i] == r.v) cand = bin_of<SF>(t, i);
            }
        }
    }
    r.i = __reduce_min_sync(full, cand);
    if (r.i == 0x7fffffff) r.i = 0;
    r.second = 0.0f;
    if (!FAITH) {
        const unsigned ties = __ballot_sync(full, mine);
        const float x = mine ? fmaxf(lo[0], 0.0f) : mine_v;
        r.second = __uint_as_float(__reduce_max_sync(full, __float_as_uint(x)));
        if (__popc(ties) > 1) r.second = r.v;
    }
    if (TPW > 32) {
        constexpr int WARPS = TPW / 32;
        if ((t & 31) == 0) red[slot * WARPS + (t >> 5)] = r;
        Group<TPW>::sync(slot);
        if (t == 0) {
            r = red[slot * WARPS];
#pragma unroll
            for (int w = 1; w < WARPS; ++w) r = top_merge(r, red[slot * WARPS + w]);
        }
    }
    return r;
}

} // namespace
} // namespace lora_b200
