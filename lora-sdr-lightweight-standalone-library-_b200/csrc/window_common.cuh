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
// `tb` = the thread index the bins belong to (bin_of); differs from t under the half-warp mapping of fft_blocked.
template <int SF, bool FAITH, int TPW>
PM_DEV Top window_top(const c64 (&v)[16], int t, int slot, Top* red, int tb = -1)
{
    if (tb < 0) tb = t;
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
                if (m2[i] == r.v) cand = bin_of<SF>(tb, i);
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

// Peak of one window for the fused-math flavour, TPW >= 32.  Every thread only finds the largest of its 16 values;
// REDUX gives the warp's largest and the largest among the OTHER lanes; for windows spread over several warps one
// round through shared memory (two floats per warp) gives every thread the window's maximum.  Only the thread that
// holds that maximum then looks for the bin index and for the runner-up inside its own 16 values -- a branch that is
// taken by one thread per window -- and that thread, not thread 0, reports the window.  Returns true in the one
// reporting thread.  On an exact tie between threads the elected reporter sets second == max, i.e. untrusted: the
// bit-faithful recheck then settles the lowest-index rule.
template <int SF, int TPW>
PM_DEV bool window_top_fast(const c64 (&v)[16], int t, int slot, float* redw /* 2 * kCta/32 */, Top& out, int tb = -1)
{
    if (tb < 0) tb = t;
    static_assert(TPW >= 32, "sub-warp windows use group_top");
    using B = Blk<SF>;
    constexpr int P = B::NP - 1, Rp = B::R(P), Gp = B::G(P), WARPS = TPW / 32;
    const unsigned full = 0xffffffffu;
    float m2[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) m2[i] = mag2<false>(v[i]);
    float h8[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) h8[i] = fmaxf(m2[2 * i], m2[2 * i + 1]);
    const float hi = fmaxf(fmaxf(fmaxf(fmaxf(h8[0], h8[1]), fmaxf(h8[2], h8[3])), fmaxf(fmaxf(h8[4], h8[5]), fmaxf(h8[6], h8[7]))), 0.0f);
    const float wmax = __uint_as_float(__reduce_max_sync(full, __float_as_uint(hi)));
    const bool top_lane = hi == wmax;
    const unsigned lanes = __ballot_sync(full, top_lane);
    float w2 = __uint_as_float(__reduce_max_sync(full, top_lane ? 0u : __float_as_uint(hi)));
    if (__popc(lanes) > 1) w2 = wmax;
    float M = wmax, other = w2;
    float wm[WARPS], ws[WARPS];
    if (WARPS > 1) {
        float* rm = redw + slot * 2 * WARPS; // [0, WARPS): warp maxima, [WARPS, 2*WARPS): warp runners-up
        if ((t & 31) == 0) { rm[t >> 5] = wmax; rm[WARPS + (t >> 5)] = w2; }
        Group<TPW>::sync(slot);
#pragma unroll
        for (int w = 0; w < WARPS; ++w) wm[w] = rm[w];
        M = wm[0];
#pragma unroll
        for (int w = 1; w < WARPS; ++w) M = fmaxf(M, wm[w]);
    }
    // Exactly ONE thread reports a window: on an exact tie between threads (equal-power bins, structured inputs) the
    // lowest lane of the lowest warp that holds the maximum is elected, so the recheck list receives at most one entry
    // per window and can never overflow its n_windows capacity.
    const bool cand = top_lane && hi == M && M > 0.0f; // a value that is not > 0 never wins
    const unsigned cands = __ballot_sync(full, cand);
    int first_holder = 0;
    if (WARPS > 1) {
#pragma unroll
        for (int w = WARPS - 1; w >= 0; --w) if (wm[w] == M) first_holder = w;
    }
    const bool winner = cand && (int)(t & 31) == __ffs(cands) - 1 && (WARPS == 1 || (t >> 5) == first_holder);
    if (winner) {
        if (WARPS > 1) {
            const float* rm = redw + slot * 2 * WARPS;
            int holders = 0;
            other = 0.0f;
#pragma unroll
            for (int w = 0; w < WARPS; ++w) {
                ws[w] = rm[WARPS + w];
                const bool holds = wm[w] == M;
                holders += holds ? 1 : 0;
                other = fmaxf(other, holds ? ws[w] : wm[w]);
            }
            if (holders > 1) other = M;
        }
        int idx = 0, dup = 0;
        float lo = 0.0f;
        // descending scan over this thread's bins: the last hit is the lowest bin holding the maximum
#pragma unroll
        for (int e = Rp - 1; e >= 0; --e) {
#pragma unroll
            for (int g = Gp - 1; g >= 0; --g) {
                const int i = g * Rp + e;
                const bool hit = m2[i] == M;
                if (hit) { idx = bin_of<SF>(tb, i); ++dup; }
                else lo = fmaxf(lo, m2[i]);
            }
        }
        out.v = M;
        out.i = idx;
        out.second = dup > 1 ? M : fmaxf(other, lo);
        return true;
    }
    if (t == 0 && !(M > 0.0f)) { // all-zero / all-NaN window: index 0, never trusted
        out.v = M;
        out.i = 0;
        out.second = M;
        return true;
    }
    return false;
}

} // namespace
} // namespace lora_b200
