// fft_blocked.cuh -- the reference's 2^SF-point FFT as a register-blocked schedule of the SAME butterfly graph.
//
// kissfft for N = 2^SF is a decimation-in-time recursion with radix list 4,4,...[,2] (kissfft.hh:80-97): stage s
// (0 = outermost) has remainder m_s = N/(r_0..r_s) and twiddle stride F_s = r_0..r_{s-1}; its butterflies
// (kf_bfly4 :164-185, kf_bfly2 :155-162) combine positions base + q*m_s with twiddles tw[k*F_s*q],
// k = position mod m_s.  A butterfly is a pure function of its operands, so any schedule of the same graph
// gives the same bits.  The schedule here:
//
//   * a window is owned by T = N/16 threads, 16 points each, held in registers as packed (re,im) pairs;
//   * the stages are grouped, innermost first, into passes of two stages (radix 16 = 4*4, or 8 = 2*4 for the
//     innermost pass of an odd SF, or a lone radix-4 outermost stage): SF7 8.16, SF8 16.16, SF9 8.16.4,
//     SF10 16.16.4, SF11 8.16.16, SF12 16.16.16;
//   * pass 0 needs, per thread, exactly the input samples t + T*j (j = 0..15): the mixed-radix digit reversal
//     of kissfft's leaf copy (:114-120) becomes a compile-time register permutation, and the global loads are
//     coalesced;
//   * between passes the points are exchanged through shared memory at pad(position); pad() is chosen per SF
//     so that every 64-bit exchange access is bank-conflict free (searched by scripts/fft_schedule_model.py),
//     and pad(base + e*m) = pad(base) + pad(e*m), so every access is one LDS/STS with an immediate offset;
//   * after the last pass thread t holds bins t + T*(g + G*e) in natural order, and the |X|^2 argmax runs on
//     registers -- the spectrum is never written anywhere.
//
// FAITH = true keeps every multiplication the reference performs (including the ones by tw[0] = 1+0j) as
// separately rounded IEEE operations: bit-identical to the CPU.  FAITH = false fuses multiply-adds and skips
// the multiplications by 1.
#pragma once

#include "packed_math.cuh"

// A/B switch: the fused flavour derives w^2k, w^3k from w^k for SF >= this value (13 = never)
#ifndef LORA_V_DERIVE_FROM
#define LORA_V_DERIVE_FROM 11
#endif

namespace lora_b200 {

template <int SF>
struct Blk {
    static constexpr int N = 1 << SF;
    static constexpr int L4 = SF / 2;        // radix-4 stages
    static constexpr int L = (SF + 1) / 2;   // stages in all
    static constexpr int T = N / 16;         // threads per window
    static constexpr int NP = (L + 1) / 2;   // passes
    __host__ __device__ static constexpr int radix(int s) { return s < L4 ? 4 : 2; }
    // Closed forms (shifts only).  Written as product loops with a division, these reached ptxas as run-time
    // divisions by 16 and 64 inside the window loop (I2F/MUFU.RCP/IMAD.HI sequences in the SASS); the loop forms are
    // kept below as m_ref/F_ref/toff_ref and compared by static_assert.
    __host__ __device__ static constexpr int lp(int s) { return s < L4 ? 2 * (s + 1) : 2 * L4 + (s - L4 + 1); } // log2(r_0..r_s)
    __host__ __device__ static constexpr int m(int s) { return N >> lp(s); }
    __host__ __device__ static constexpr int F(int s) { return s == 0 ? 1 : 1 << lp(s - 1); }
    static constexpr int m_ref(int s) { int p = 1; for (int u = 0; u <= s; ++u) p *= radix(u); return N / p; }
    static constexpr int F_ref(int s) { int p = 1; for (int u = 0; u < s; ++u) p *= radix(u); return p; }
    // pass p (0 = innermost, executed first) covers stages pa(p)..pb(p)
    __host__ __device__ static constexpr int pb(int p) { return L - 1 - 2 * p; }
    __host__ __device__ static constexpr int pa(int p) { return pb(p) >= 1 ? pb(p) - 1 : 0; }
    __host__ __device__ static constexpr int R(int p) { return pa(p) == pb(p) ? radix(pb(p)) : radix(pa(p)) * radix(pb(p)); }
    __host__ __device__ static constexpr int G(int p) { return 16 / R(p); }
    __host__ __device__ static constexpr int pad(int X)
    {
        return SF == 7 ? X + (X >> 3) + (X >> 4) : SF == 8 ? X + (X >> 4) : X + (X >> 4) + (X >> 8);
    }
    // shared-memory footprint of one window, in c64 units; chosen = 8 (mod 16) for SF7 so that the four
    // windows of a warp interleave over all banks
    static constexpr int WIN = SF == 7 ? 152 : SF == 8 ? 272 : ((pad(N - 1) + 1 + 1) & ~1);
    // Per-stage twiddle table: stage s, multiplier q (1..radix-1), k in [0, m_s) lives at toff(s) + (q-1)*m_s + k
    // and holds tw[k*F_s*q] -- the same values kissfft indexes, laid out so that the threads of a window (consecutive
    // k) read consecutive entries.
    // = sum over u < s of (radix(u) - 1) * m(u); every stage before the last is radix 4, so the sum telescopes
    __host__ __device__ static constexpr int toff(int s) { return N - (N >> (2 * s)); }
    static constexpr int toff_ref(int s) { int o = 0; for (int u = 0; u < s; ++u) o += (radix(u) - 1) * m_ref(u); return o; }
    static constexpr int TWP = N + 16; // entries allocated (the table needs N-1, or N for odd SF)
    // digit of element e (position order inside the pass) at stage s
    __host__ __device__ static constexpr int digit(int p, int s, int e) { return (e >> (lp(pb(p)) - lp(s))) & (radix(s) - 1); }
    // pass 0: register slot of the j-th loaded sample (input index t + T*j)
    __host__ __device__ static constexpr int slot0(int j)
    {
        const int g = j % G(0), jj = j / G(0);
        int e = 0, rem = jj;
        for (int s = pa(0); s <= pb(0); ++s) { e += (rem & (radix(s) - 1)) << (lp(pb(0)) - lp(s)); rem >>= (radix(s) == 4 ? 2 : 1); }
        return g * R(0) + e;
    }
};

template <int SF>
constexpr bool blk_closed_forms_ok()
{
    using B = Blk<SF>;
    for (int s = 0; s < B::L; ++s)
        if (B::m(s) != B::m_ref(s) || B::F(s) != B::F_ref(s) || B::toff(s) != B::toff_ref(s)) return false;
    for (int p = 0; p < B::NP; ++p) {
        int r = 1;
        for (int s = B::pa(p); s <= B::pb(p); ++s) r *= B::radix(s);
        if (r != B::R(p)) return false;
        for (int s = B::pa(p); s <= B::pb(p); ++s)
            for (int e = 0; e < 16; ++e)
                if (B::digit(p, s, e) != (e / (B::m_ref(s) / B::m_ref(B::pb(p)))) % B::radix(s)) return false;
    }
    return true;
}
static_assert(blk_closed_forms_ok<7>() && blk_closed_forms_ok<8>() && blk_closed_forms_ok<9>() && blk_closed_forms_ok<10>() &&
              blk_closed_forms_ok<11>() && blk_closed_forms_ok<12>(), "closed forms of the stage geometry");

// The 16 twiddles pass 0 can need, tw0[k*4+q] = tw[k * F(pa(0)) * q]; thread-independent, so they travel as
// kernel arguments (constant bank) instead of being loaded.
struct Tw0 {
    float2 w[16];
};

template <int SF, bool FAITH, bool GUARD>
PM_DEV void bfly4(c64& f0, c64& f1, c64& f2, c64& f3, c64 w1, c64 w2, c64 w3, bool trivial, float one)
{
    c64 s0 = f1, s1 = f2, s2 = f3;
    if (FAITH || !trivial) { s0 = cmul<FAITH>(f1, w1, one); s1 = cmul<FAITH>(f2, w2, one); s2 = cmul<FAITH>(f3, w3, one); }
    const c64 s5 = GUARD ? csub_g<FAITH>(f0, s1, one) : sub2(f0, s1);
    const c64 g0 = GUARD ? cadd_g<FAITH>(f0, s1, one) : add2(f0, s1);
    const c64 s3 = add2(s0, s2);
    const c64 s4 = mul_mj(sub2(s0, s2));
    f2 = sub2(g0, s3);
    f0 = add2(g0, s3);
    f1 = add2(s5, s4);
    f3 = sub2(s5, s4);
}

template <int SF, bool FAITH, bool GUARD>
PM_DEV void bfly2(c64& f0, c64& f1, c64 w1, bool trivial, float one)
{
    c64 x = f1;
    if (FAITH || !trivial) x = cmul<FAITH>(f1, w1, one);
    f1 = GUARD ? csub_g<FAITH>(f0, x, one) : sub2(f0, x);
    f0 = GUARD ? cadd_g<FAITH>(f0, x, one) : add2(f0, x);
}

// Twiddles of one pass held in registers across windows (fused flavour, small SF): a thread's twiddles depend on its
// index and the stage only, never on the window, so a slot that walks many windows can load them once.  Per group:
// entry [0] = the inner stage of the pass (k = klow), entries [1 + c] = the outer stage at k = klow + c * m(b); three
// multipliers (w^k, w^2k, w^3k) each.  15 complex values for a radix-16 pass, 3 per group for a lone radix-4 stage.
template <int SF, int P>
PM_DEV int group_base_klow(int t, int g, int& klow) // klow of group g of thread t in pass P (see group_base below)
{
    using B = Blk<SF>;
    constexpr int mb = B::m(B::pb(P));
    const int gam = t + B::T * g;
    klow = P == 0 ? 0 : (gam & (mb - 1));
    return klow;
}

struct PreTw {
    c64 w[4][5][3];
};

template <int SF, int P>
PM_DEV void preload_twiddles(PreTw& pre, int t, const c64* __restrict__ tw)
{
    using B = Blk<SF>;
    constexpr int a = B::pa(P), b = B::pb(P), Gp = B::G(P);
#pragma unroll
    for (int g = 0; g < Gp; ++g) {
        int klow;
        (void)group_base_klow<SF, P>(t, g, klow);
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            if (q < B::radix(b) - 1) pre.w[g][0][q] = __ldg(tw + klow + B::toff(b) + q * B::m(b));
        }
        if (a != b) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (c < B::radix(b)) {
#pragma unroll
                    for (int q = 0; q < 3; ++q)
                        if (q < B::radix(a) - 1) pre.w[g][1 + c][q] = __ldg(tw + klow + B::toff(a) + c * B::m(b) + q * B::m(a));
                }
            }
        }
    }
}

// All stages of pass P on the 16 register-resident points of one thread.  klow[g] = (position mod m_b) of
// group g (0 in pass 0).  `tw` is the per-stage table (Blk::toff).
template <int SF, int P, bool FAITH, bool PRE = false>
PM_DEV void pass_stages(c64 (&v)[16], const int (&klow)[4], const c64* __restrict__ tw, const Tw0& tw0, float one,
                        const PreTw* pre = nullptr)
{
    using B = Blk<SF>;
    constexpr int a = B::pa(P), b = B::pb(P), Rp = B::R(P), Gp = B::G(P), mb = B::m(b);
#pragma unroll
    for (int s = b; s >= a; --s) {
        const int w = 1 << (B::lp(b) - B::lp(s)); // = m(s) / mb, register stride of this stage's digit
        const int rs = B::radix(s);
#pragma unroll
        for (int g = 0; g < Gp; ++g) {
#pragma unroll
            for (int e0 = 0; e0 < Rp; ++e0) {
                if (B::digit(P, s, e0) != 0) continue;
                int cpart = 0; // contribution of the in-pass digits below stage s to k
#pragma unroll
                for (int u = s + 1; u <= b; ++u) cpart += B::digit(P, u, e0) * B::m(u);
                const int i0 = g * Rp + e0;
                c64 w1, w2 = 0, w3 = 0;
                bool trivial;
                if (P == 0) { // k = cpart, thread independent
                    trivial = cpart == 0;
                    const int k = cpart; // < radix(b) <= 4, and only the outer stage of the pass has k != 0
                    w1 = pk(tw0.w[k * 4 + 1].x, tw0.w[k * 4 + 1].y);
                    if (rs == 4) {
                        w2 = pk(tw0.w[k * 4 + 2].x, tw0.w[k * 4 + 2].y);
                        w3 = pk(tw0.w[k * 4 + 3].x, tw0.w[k * 4 + 3].y);
                    }
                } else if (PRE) {
                    trivial = false;
                    const int ki = (s == b) ? 0 : 1 + cpart / mb; // compile-time after unrolling
                    w1 = pre->w[g][ki][0];
                    if (rs == 4) { w2 = pre->w[g][ki][1]; w3 = pre->w[g][ki][2]; }
                } else {
                    trivial = false;
                    const c64* wp = tw + klow[g] + (B::toff(s) + cpart); // per-stage table, immediate offsets
                    w1 = __ldg(wp);
                    if (rs == 4) {
                        // Fused flavour, SF >= 11: w^2k and w^3k from w^k (4 packed instructions instead of two
                        // loads).  Measured +5 % / +3 % at SF11 / SF12, where load latency binds, and -1 ... -3 % at
                        // SF9 / SF10 (deriving only w^3k: +-1 % everywhere); the extra 1e-7 relative error is far inside the margin of the argmax filter.
                        if (FAITH || SF < LORA_V_DERIVE_FROM) { w2 = __ldg(wp + B::m(s)); w3 = __ldg(wp + 2 * B::m(s)); }
                        else {
                            w2 = cmul<false>(w1, w1, one);
                            w3 = cmul<false>(w2, w1, one);
                        }
                    }
                }
                constexpr bool dummy = false; (void)dummy;
                if (rs == 4) {
                    if (P == 0 && s == b) bfly4<SF, FAITH, true>(v[i0], v[i0 + w], v[i0 + 2 * w], v[i0 + 3 * w], w1, w2, w3, trivial, one);
                    else bfly4<SF, FAITH, false>(v[i0], v[i0 + w], v[i0 + 2 * w], v[i0 + 3 * w], w1, w2, w3, trivial, one);
                } else {
                    if (P == 0 && s == b) bfly2<SF, FAITH, true>(v[i0], v[i0 + w], w1, trivial, one);
                    else bfly2<SF, FAITH, false>(v[i0], v[i0 + w], w1, trivial, one);
                }
            }
        }
    }
}

// Position base of group g of thread t in pass P, and the klow it implies.
template <int SF, int P>
PM_DEV int group_base(int t, int g, int& klow)
{
    using B = Blk<SF>;
    constexpr int a = B::pa(P), mb = B::m(B::pb(P)), Rp = B::R(P);
    const int gam = t + B::T * g;
    if (P == 0) { // digit reversal of the input base index over the stages outside the pass (all radix 4)
        int X = 0, rem = gam;
#pragma unroll
        for (int s = 0; s < a; ++s) { X += (rem & 3) * B::m(s); rem >>= 2; }
        klow = 0;
        return X;
    }
    klow = gam & (mb - 1);
    return (gam / mb) * (Rp * mb) + klow;
}

template <int SF, int P>
PM_DEV void exchange_store(const c64 (&v)[16], c64* __restrict__ S, int t)
{
    using B = Blk<SF>;
    constexpr int Rp = B::R(P), Gp = B::G(P), mb = B::m(B::pb(P));
#pragma unroll
    for (int g = 0; g < Gp; ++g) {
        int kl;
        const int A = B::pad(group_base<SF, P>(t, g, kl));
#pragma unroll
        for (int e = 0; e < Rp; ++e) S[A + B::pad(e * mb)] = v[g * Rp + e];
    }
}

template <int SF, int P>
PM_DEV void exchange_load(c64 (&v)[16], int (&klow)[4], const c64* __restrict__ S, int t)
{
    using B = Blk<SF>;
    constexpr int Rp = B::R(P), Gp = B::G(P), mb = B::m(B::pb(P));
#pragma unroll
    for (int g = 0; g < Gp; ++g) {
        const int A = B::pad(group_base<SF, P>(t, g, klow[g]));
#pragma unroll
        for (int e = 0; e < Rp; ++e) v[g * Rp + e] = S[A + B::pad(e * mb)];
    }
}

// Whole transform.  `in[j]` = preprocessed input sample t + T*j; on return `v[i]` holds bin bin_of<SF>(t, i).
// S = this window's shared-memory region (Blk<SF>::WIN entries); sync() = barrier over the window's T threads;
// tw = the per-stage twiddle table.  The caller must sync() before S is written again (next window).
// Half-warp mapping of the last two passes (SF11, SF12: three passes, the last two radix 16).  The sixteen threads that
// combine at one frequency in the last pass -- those sharing klow of pass 1 -- are made the sixteen lanes of an aligned
// half-warp by running passes 1 and 2 on the logical index u = (t & 15) * m1 + (t >> 4), m1 = m(pb(1)).  Pass 2 of
// thread u then reads only what threads of its own half-warp wrote in pass 1, so the SECOND exchange needs
// __syncwarp() instead of a barrier over the whole window (8 or 4 warps).  Pass 0 keeps the natural index (its global
// loads must stay coalesced), so the first exchange is unchanged.  Pass-1 twiddles become two addresses per warp
// (broadcast); pass-2 twiddles would be strided, so they come from a copy of the table permuted by the same map
// (stored right behind the table, at +TWP: every block of T entries of stages 0 and 1 reordered so that thread t finds
// the entry of u at offset t).
template <int SF>
struct HalfWarp {
    using B = Blk<SF>;
    static constexpr bool OK = B::NP == 3 && B::R(1) == 16 && B::R(2) == 16 && B::G(1) == 1 && B::G(2) == 1;
    static constexpr int M1 = B::m(B::pb(1));
    __host__ __device__ static constexpr int u_of(int t) { return (t & 15) * M1 + (t >> 4); }
};

// PREMASK: bit p set = pass p takes its twiddles from pre[p] (preload_twiddles) instead of loading them per window.
// HW: the half-warp mapping above (only where HalfWarp<SF>::OK); the caller must use HalfWarp<SF>::u_of(t) for bin_of.
template <int SF, bool FAITH, int PREMASK = 0, bool HW = false, typename Sync>
PM_DEV void fft_blocked(const c64 (&in)[16], c64 (&v)[16], c64* __restrict__ S, const c64* __restrict__ tw,
                        const Tw0& tw0, int t, float one, Sync sync, const PreTw* pre = nullptr)
{
    using B = Blk<SF>;
    static_assert(!HW || HalfWarp<SF>::OK, "half-warp mapping needs two trailing radix-16 passes");
    const int u = HW ? HalfWarp<SF>::u_of(t) : t;
    int klow[4] = {0, 0, 0, 0};
#pragma unroll
    for (int j = 0; j < 16; ++j) v[B::slot0(j)] = in[j];
    pass_stages<SF, 0, FAITH>(v, klow, tw, tw0, one);
    exchange_store<SF, 0>(v, S, t);
    sync();
    exchange_load<SF, 1>(v, klow, S, u);
    pass_stages<SF, 1, FAITH, (PREMASK & 2) != 0>(v, klow, tw, tw0, one, pre ? pre + 1 : nullptr);
    if (B::NP == 3) {
        exchange_store<SF, 1>(v, S, u); // in place: a thread overwrites exactly the positions it read

        if (HW) __syncwarp(); // writers and readers of this exchange share a half-warp
        else sync();
        exchange_load<SF, (B::NP == 3 ? 2 : 1)>(v, klow, S, u);
        if (HW) klow[0] = t; // the permuted table is indexed by the physical thread
        pass_stages<SF, (B::NP == 3 ? 2 : 1), FAITH, (PREMASK & 4) != 0>(v, klow, HW ? tw + B::TWP : tw, tw0, one, pre ? pre + 2 : nullptr);
    }
}

// Natural-order bin held in register i of thread t after fft_blocked.
template <int SF>
__host__ __device__ constexpr int bin_of(int t, int i)
{
    using B = Blk<SF>;
    constexpr int P = B::NP - 1;
    return t + B::T * (i / B::R(P)) + (i % B::R(P)) * B::m(B::pb(P));
}

} // namespace lora_b200
