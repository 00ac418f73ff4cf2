// stream_kernels.cu -- continuous-stream frame synchroniser (LoRa-SDR/LoRaDemod.cpp:145-327) and full-frame
// generator (LoRa-SDR/LoRaMod.cpp:109-238).
//
// The synchroniser is a serial state machine per stream: how far the input advances depends on the bin detected
// in the window just processed.  Parallelism therefore comes from (a) the T = N/16 threads that share each window
// (dechirp, fine-tune rotation, bit-faithful FFT, peak search) and (b) many independent streams -- the gateway's
// channels, or the segments a host cuts a long capture into -- one per slot of a persistent grid.
// Every thread of a slot runs the state machine on identical inputs, so no decision needs broadcasting.
#include "stream_kernels.h"

#include "std_chain.cuh"

#include <cmath>

#include "demod_kernels.h" // g_launches
#include "bulk_copy.cuh"
#include "chirp_pipeline.cuh"
#include "exact_math.h"
#include "window_common.cuh"

// LORA_V_STREAM_PF = 1: the synchroniser asks for the window after next to be pulled into L2 -- every window is a first
// touch of its samples.  Same-box A/B on configs[3]: +2.5 / +3 / +1 / +1.7 % at SF7 / 8 / 9 / 10 (profiles/r2_ab_stream_pf.txt).
// Resident CTAs per SM of the synchroniser (A/B switch; 3 means 80 registers per thread).
#ifndef LORA_V_STREAM_CTAS
#define LORA_V_STREAM_CTAS 2
#endif
#ifndef LORA_V_STREAM_PF
#define LORA_V_STREAM_PF 1
#endif

namespace lora_b200 {

namespace {

// ---- the fine-tune index walk ---------------------------------------------------------------------------------
// LoRaDemod.cpp:157-165: per sample, an int index is converted to float, reduced by fine_err*128 (a float), truncated
// back to int and wrapped once.  The rounding of the float subtraction depends on the binade of the difference, so
// the walk is piecewise regular: a closed form (see fine_walk) with breaks at binade crossings and ties.
// It is evaluated exactly without serialising it: guess the closed form from a known-correct start, let every
// thread check "one exact step from my guessed index gives the next guessed index" for its own positions, and
// accept the prefix before the first failed check (induction over positions).  The step after that prefix is
// recomputed exactly and becomes the next start.  A round costs 16 exact steps per thread and one group minimum;
// the number of rounds is the number of breaks (usually one round per window).
template <int N>
PM_DEV int fine_step(int idx, float d)
{
    constexpr int W = N * STREAM_FINE_STEPS;
    int r = __float2int_rz(__fsub_rn(__int2float_rn(idx), d));
    if (r < 0) r += W;
    else if (r >= W) r -= W;
    // The reference indexes its table with whatever comes out (undefined behaviour once |d| >= W or d is NaN);
    // here such an index is folded to 0 so the table read stays inside the allocation.
    if ((unsigned)r >= (unsigned)W) r = 0;
    return r;
}

template <int T>
PM_DEV int group_min(int x, int t, int slot, int* scratch /* 2 * (T/32) ints per slot when T > 32 */, int round)
{
    if (T <= 32) {
        const unsigned mask = Group<T>::mask();
#pragma unroll
        for (int o = T / 2; o > 0; o >>= 1) x = min(x, __shfl_xor_sync(mask, x, o));
        return x;
    } else {
        constexpr int WARPS = T / 32;
        x = __reduce_min_sync(0xffffffffu, x);
        int* s = scratch + (slot * 2 + (round & 1)) * WARPS; // double-buffered: one barrier per round
        if ((t & 31) == 0) s[t >> 5] = x;
        Group<T>::sync(slot);
        int r = s[0];
#pragma unroll
        for (int w = 1; w < WARPS; ++w) r = min(r, s[w]);
        return r;
    }
}

// Same-box A/B of the out-of-line rounds (profiles/r2_ab_stream_cold.txt): SF7 -2.6 %, SF8 +4.6 %, SF9 +1.7 %, SF10 -1.8 %.
#ifndef LORA_V_STREAM_COLD_WALK
#define LORA_V_STREAM_COLD_WALK(SF) ((SF) == 8 || (SF) == 9)
#endif
template <int SF>
PM_DEV void fine_walk_rounds(int idx0, float d, unsigned Mod, int t, int slot, int* scratch, int (&idx)[16], int& carry);
// out of line: kept out of the synchroniser's main loop, which is larger than the instruction cache likes
template <int SF>
__device__ __noinline__ void fine_walk_checked(int idx0, float d, unsigned Mod, int t, int slot, int* scratch, int (&idx)[16], int& carry)
{
    fine_walk_rounds<SF>(idx0, d, Mod, t, slot, scratch, idx, carry);
}

template <int SF>
PM_DEV void fine_walk(int idx0, float fine_err, int t, int slot, int* scratch, int (&idx)[16], int& carry)
{
    using B = Blk<SF>;
    constexpr int N = B::N, T = B::T;
    constexpr unsigned W = N * STREAM_FINE_STEPS;
    const float d = __fmul_rn(fine_err, (float)STREAM_FINE_STEPS);
    if (d == 0.0f) { // (int)((float)idx - 0) == idx
#pragma unroll
        for (int j = 0; j < 16; ++j) idx[j] = idx0;
        carry = idx0;
        return;
    }
    // The walk is an arithmetic progression modulo W while the float difference stays non-negative (truncation =
    // floor).  For a positive fractional d the difference goes negative just before every wrap, truncation toward
    // zero then takes one unit less, and the wrap adds W: together that is the SAME stride modulo W + 1.  (The one
    // exception, an index equal to floor(d), and the rare d whose fraction is below the float grid of the
    // difference are caught by the check below like any other break.)
    const unsigned Mod = (d > 0.0f && d != truncf(d)) ? W + 1u : W;
    // Shortcut, taken by ~7 windows in 8: when |d| < W and its fraction is either zero or at least 1/16 away from
    // an integer, the float subtraction (whose result is below 2^20, so rounded to a grid of 1/16 or finer) can
    // never round across an integer, truncation gives exactly what real arithmetic would, and the closed form needs
    // no checking -- except that modulo W + 1 it produces W where the true walk, passing through floor(d), goes to
    // 0.  A guess equal to W at any position sends the window to the checked path below.
    {
        const float ad = fabsf(d), fl = floorf(ad), fr = ad - fl;
        if (ad < (float)W && (fr == 0.0f || (fr >= 0.0625f && fr <= 0.9375f))) {
            const unsigned whole = (unsigned)fl;
            // stride subtracted per step, modulo Mod: d > 0 -> floor(d) (+1 if fractional); d < 0 -> adding floor(|d|)
            const unsigned s = d > 0.0f ? whole + (fr != 0.0f ? 1u : 0u) : (whole == 0u ? 0u : Mod - whole);
            const unsigned step_t = ((unsigned)T * s) % Mod;
            unsigned h = ((unsigned)idx0 + Mod - ((unsigned)t * s) % Mod) % Mod;
            bool hit_w = false;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                idx[j] = (int)h;
                hit_w = hit_w || h == W;
                h = h >= step_t ? h - step_t : h + Mod - step_t;
            }
            const unsigned last = ((unsigned)idx0 + Mod - ((unsigned)N * s) % Mod) % Mod; // position N: N*s < 2^32
            hit_w = hit_w || last == W;
            if (group_min<T>(hit_w ? 0 : 1, t, slot, scratch, 0) != 0) {
                carry = (int)last;
                return;
            }
            Group<T>::sync(slot); // the scratch buffer of round 0 is written again below
        }
    }
    if (LORA_V_STREAM_COLD_WALK(SF)) { fine_walk_checked<SF>(idx0, d, Mod, t, slot, scratch, idx, carry); return; }
    fine_walk_rounds<SF>(idx0, d, Mod, t, slot, scratch, idx, carry);
}

// The checked rounds of the walk (about one window in eight gets here).
template <int SF>
PM_DEV void fine_walk_rounds(int idx0, float d, unsigned Mod, int t, int slot, int* scratch, int (&idx)[16], int& carry)
{
    using B = Blk<SF>;
    constexpr int N = B::N, T = B::T;
    int base_pos = 0, base_idx = idx0;
    for (int round = 0;; ++round) {
        const int diff = base_idx - fine_step<N>(base_idx, d);                // in (-W, W)
        const unsigned s = diff < 0 ? (unsigned)(diff + (int)Mod) : (unsigned)diff; // guessed stride in [0, Mod)
        const unsigned step_t = ((unsigned)T * s) % Mod;                          // T*s < 2^8 * 2^20: 32-bit arithmetic throughout
        const int j0 = base_pos <= t ? 0 : (base_pos - t + T - 1) / T;        // first of this thread's positions >= base_pos
        const unsigned k0 = (unsigned)(t + T * j0 - base_pos);                // < T
        unsigned h = ((unsigned)base_idx + Mod - (k0 * s) % Mod) % Mod;
        int fail = N;
        int guess[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            if (j >= j0) {
                const unsigned hn = h >= s ? h - s : h + Mod - s;
                guess[j] = (int)h;
                if ((unsigned)fine_step<N>((int)h, d) != hn) fail = min(fail, t + T * j);
                h = h >= step_t ? h - step_t : h + Mod - step_t;
            }
        }
        const int pf = group_min<T>(fail, t, slot, scratch, round); // first position whose successor is not the guess
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const int p = t + T * j;
            if (j >= j0 && p <= pf) idx[j] = guess[j];
        }
        const unsigned span = (unsigned)((pf >= N ? N : pf) - base_pos);      // steps from the base to the last accepted guess
        const unsigned h_last = ((unsigned)base_idx + Mod - (span * s) % Mod) % Mod; // span*s <= 2^12 * 2^19 + 2^12 < 2^32
        if (pf >= N) { // every check passed, including the one at N-1 that validates position N
            carry = (int)h_last;
            return;
        }
        base_idx = fine_step<N>((int)h_last, d); // exact successor of the last accepted position
        base_pos = pf + 1;
        if (base_pos >= N) { carry = base_idx; return; }
    }
}

// ---- one window through the detector ---------------------------------------------------------------------------
struct Det {
    int value;
    float power, power_avg, f_index;
};

template <int SF>
struct StreamCfg {
    static constexpr int T = Blk<SF>::T;
    static constexpr int SLOTS = kCta / T;
    static constexpr int SMEM_C64 = SLOTS * Blk<SF>::WIN;
};

struct SlotShared {
    Top* red;        // kCta/32
    double* redd;    // kCta/32
    int* redi;       // 2 * kCta/32
    float4* bc;      // per slot: (value bits, max bits, total lo, total hi)
};

template <int SF>
PM_DEV Det detect_window(const StreamDetectArgs& a, const c64* __restrict__ x, const c64* __restrict__ chirp, int& ft,
                         float fine_err, c64* S, const SlotShared& sh, int t, int slot)
{
    using B = Blk<SF>;
    constexpr int N = B::N, T = B::T;
    int idx[16];
    int carry;
    fine_walk<SF>(ft, fine_err, t, slot, sh.redi, idx, carry);
    ft = carry;
    c64 in[16], v[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) { // LoRaDemod.cpp:157-160: (samp * chirp) * fineTune, two rounded complex products
        const c64 s = __ldg(x + (size_t)(t + T * j) * a.osr);
        const c64 dc = cmul<true>(s, __ldg(chirp + t + T * j), a.one);
        in[j] = cmul<true>(dc, __ldg(a.fine + idx[j]), a.one);
    }
    fft_blocked<SF, true>(in, v, S, a.twp, a.tw0, t, a.one, [slot] { Group<T>::sync(slot); });
    Top top = window_top<SF, true, T>(v, t, slot, sh.red);
    // LoRaDetector.hpp:46-58 accumulates |X|^2 in a double, bin after bin; here the same doubles are added as a tree
    // (per thread, then across the slot).  Each term has a 24-bit mantissa, so the double sum is exact -- and the
    // order irrelevant -- unless the terms span more than 2^17 in magnitude; beyond that the two orders can differ in
    // the last bits of the double, far below the float the total is narrowed to at :60.
    double tot = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) tot = __dadd_rn(tot, (double)mag2<true>(v[i]));
    {
        constexpr int Wd = T < 32 ? T : 32;
        const unsigned mask = Group<T>::mask();
#pragma unroll
        for (int o = Wd / 2; o > 0; o >>= 1) tot = __dadd_rn(tot, __shfl_xor_sync(mask, tot, o));
    }
    Group<T>::sync(slot); // exchange loads done: the region now takes the natural-order spectrum
#pragma unroll
    for (int i = 0; i < 16; ++i) S[bin_of<SF>(t, i)] = v[i];
    if (T > 32) {
        constexpr int WARPS = T / 32;
        if ((t & 31) == 0) sh.redd[slot * WARPS + (t >> 5)] = tot;
        Group<T>::sync(slot);
        if (t == 0) {
            double sum = sh.redd[slot * WARPS];
#pragma unroll
            for (int w = 1; w < WARPS; ++w) sum = __dadd_rn(sum, sh.redd[slot * WARPS + w]);
            sh.bc[slot] = make_float4(__int_as_float(top.i), top.v, __int_as_float(__double2loint(sum)),
                                      __int_as_float(__double2hiint(sum)));
        }
        Group<T>::sync(slot);
        const float4 b = sh.bc[slot];
        top.i = __float_as_int(b.x);
        top.v = b.y;
        tot = __hiloint2double(__float_as_int(b.w), __float_as_int(b.z));
    } else {
        Group<T>::sync(slot);
    }
    Det r;
    r.value = top.i;
    // LoRaDetector.hpp:60-71
    const float noise = __fsqrt_rn((float)__dsub_rn(tot, (double)top.v));
    const float fundamental = __fsqrt_rn(top.v);
    r.power_avg = __fsub_rn(__fmul_rn(20.0f, log10f(noise)), a.power_scale);
    r.power = __fsub_rn(__fmul_rn(20.0f, log10f(fundamental)), a.power_scale);
    const float2 Lb = up(S[top.i > 0 ? top.i - 1 : N - 1]);
    const float2 Rb = up(S[top.i < N - 1 ? top.i + 1 : 0]);
    const float left = xm_hypotf(Lb.x, Lb.y), right = xm_hypotf(Rb.x, Rb.y);
    const double demon = __dsub_rn(__dsub_rn(__dmul_rn(2.0, (double)fundamental), (double)right), (double)left);
    r.f_index = (demon == 0.0) ? 0.0f : (float)__ddiv_rn(__dmul_rn(0.5, (double)__fsub_rn(right, left)), demon);
    Group<T>::sync(slot); // S is rewritten by the next window
    return r;
}

enum { FRAMESYNC, DOWNCHIRP0, DOWNCHIRP1, QUARTERCHIRP, DATASYMBOLS };

// The look-ahead of FRAMESYNC (:191-204) runs the detector a second time on the same stream.  It is rare (one window in
// a frame), so it calls an out-of-line copy: inlined twice, the detector made the kernel ~6000 instructions (96 KB), more
// than the instruction cache holds, and `no_instruction` was the second largest stall of the kernel.
#ifndef LORA_V_STREAM_COLD
#define LORA_V_STREAM_COLD 1
#endif
template <int SF>
__device__ __noinline__ Det detect_window_cold(const StreamDetectArgs& a, const c64* __restrict__ x, const c64* __restrict__ chirp,
                                               int& ft, float fine_err, c64* S, const SlotShared& sh, int t, int slot)
{
    return detect_window<SF>(a, x, chirp, ft, fine_err, S, sh, t, slot);
}

template <int SF>
__global__ void __launch_bounds__(kCta, LORA_V_STREAM_CTAS) stream_detect_kernel(const __grid_constant__ StreamDetectArgs a)
{
    using B = Blk<SF>;
    constexpr int N = B::N, T = B::T, SLOTS = StreamCfg<SF>::SLOTS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ Top red[kCta / 32];
    __shared__ double redd[kCta / 32];
    __shared__ int redi[2 * (kCta / 32)];
    __shared__ float4 bc[kCta / 8];
    const int slot = threadIdx.x / T, t = threadIdx.x % T;
    c64* S = reinterpret_cast<c64*>(smem_raw) + slot * B::WIN;
    const SlotShared sh{red, redd, redi, bc};
    const unsigned sync_hi = a.sync >> 4, sync_lo = a.sync & 0xfu;

    // whole slots walk the streams; slots never synchronise with each other
    for (int st = blockIdx.x * SLOTS + slot; st < a.n_streams; st += gridDim.x * SLOTS) {
        const c64* x = reinterpret_cast<const c64*>(a.iq) + (size_t)st * a.stream_stride;
        int16_t* out = a.symbols + (size_t)st * a.sym_cap;
        StreamPacket* pk_out = a.packets + (size_t)st * a.pkt_cap;
        // activate(), LoRaDemod.cpp:138-142; _prevValue/_freqError/_finefreqError start at zero (the reference leaves
        // them uninitialised, :68-75)
        int state = FRAMESYNC;
        const c64* chirp = a.up;
        int fine_idx = 0, freq_err = 0;
        float fine_err = 0.0f;
        short prev_value = 0;
        size_t pos = 0, cur_start = 0;
        unsigned sym_count = 0, n_pkt = 0, frame_limit = a.mtu;
        unsigned long long hdr_lo = 0, hdr_hi = 0;
        size_t sym_used = 0;

        const size_t avail = (a.count + a.osr - 1) / a.osr; // samples 0, osr, 2*osr, ... of the stream
        while (pos + 2 * (size_t)N <= avail) { // :148
            if (LORA_V_STREAM_PF && t == 0) { // every window is a first touch of its samples: pull the one after next into L2
                const size_t ahead = pos + 2 * (size_t)N;
                if (ahead + (size_t)N <= avail) l2_prefetch(x + ahead * a.osr, sizeof(c64) * (size_t)N * a.osr);
            }
            Det d = detect_window<SF>(a, x + pos * a.osr, chirp, fine_idx, fine_err, S, sh, t, slot);
            const unsigned value = (unsigned)d.value;
            const bool squelched = __fsub_rn(d.power, d.power_avg) < a.thresh; // :174
            size_t total = 0;
            switch (state) {
            case FRAMESYNC: { // :180-237
                const bool syncd = !squelched && (prev_value + 4) / 8 == 0;
                const bool match0 = (value + 4) / 8 == sync_hi;
                bool match1 = false;
                if (syncd && match0) {
                    int ft = fine_idx; // the look-ahead walks a copy (:191-203) but overwrites power/powerAvg/fIndex (:204)
                    d = LORA_V_STREAM_COLD ? detect_window_cold<SF>(a, x + (pos + N) * a.osr, chirp, ft, fine_err, S, sh, t, slot)
                                           : detect_window<SF>(a, x + (pos + N) * a.osr, chirp, ft, fine_err, S, sh, t, slot);
                    match1 = ((unsigned)d.value + 4) / 8 == sync_lo;
                }
                if (syncd && match0 && match1) {
                    total = 2 * (size_t)N;
                    state = DOWNCHIRP0;
                    chirp = a.down;
                } else if (!squelched) {
                    total = (size_t)N - value;
                    fine_err = __fadd_rn(fine_err, d.f_index);
                } else {
                    total = N;
                    fine_err = 0.0f;
                    fine_idx = 0;
                }
            } break;
            case DOWNCHIRP0: { // :242-250
                state = DOWNCHIRP1;
                total = N;
                int error = (int)value;
                if (value > (unsigned)N / 2) error -= N;
                freq_err = error;
            } break;
            case DOWNCHIRP1: { // :255-270
                state = QUARTERCHIRP;
                total = N;
                chirp = a.up;
                int error = (int)value;
                if (value > (unsigned)N / 2) error -= N;
                freq_err = (freq_err + error) / 2;
            } break;
            case QUARTERCHIRP: { // :275-284
                state = DATASYMBOLS;
                total = (size_t)(N / 4) + (size_t)(long long)(freq_err / 2); // size_t + int, modulo 2^64 like the reference
                fine_err = __fadd_rn(fine_err, (float)(freq_err / 2));
                sym_count = 0;
                hdr_lo = hdr_hi = 0;
                frame_limit = a.mtu;
                cur_start = pos + total;
            } break;
            default: { // DATASYMBOLS, :289-310
                total = N;
                if (t == 0 && sym_used + sym_count < a.sym_cap) out[sym_used + sym_count] = (int16_t)value;
                if (a.frame_header && sym_count < 8) { // keep the header block: four 16-bit symbols per word
                    const unsigned long long v16 = (unsigned long long)(value & 0xffffu) << (16 * (sym_count & 3));
                    if (sym_count < 4) hdr_lo |= v16; else hdr_hi |= v16;
                }
                ++sym_count;
                if (a.frame_header && sym_count == 8) { // not in the reference: frame length from the explicit header
                    uint16_t hs[8];
#pragma unroll
                    for (int k = 0; k < 4; ++k) { hs[k] = (uint16_t)(hdr_lo >> (16 * k)); hs[4 + k] = (uint16_t)(hdr_hi >> (16 * k)); }
                    const unsigned want = stdchain::std_frame_symbols(SF, a.frame_header == 2, hs);
                    frame_limit = want ? min(want, a.mtu) : 8u; // an invalid header ends the frame here
                }
                if (sym_count >= frame_limit || squelched) {
                    if (t == 0 && n_pkt < a.pkt_cap) {
                        StreamPacket p;
                        p.start = cur_start;
                        p.first_symbol = (uint32_t)sym_used;
                        p.n_symbols = sym_count;
                        p.freq_error = freq_err;
                        p.reserved = 0;
                        pk_out[n_pkt] = p;
                    }
                    ++n_pkt;
                    sym_used += sym_count;
                    if (sym_used > a.sym_cap) sym_used = a.sym_cap;
                    fine_err = 0.0f;
                    state = FRAMESYNC;
                }
            } break;
            }
            prev_value = (short)value; // :326
            pos += total;              // :321
        }
        if (t == 0) {
            StreamResult r;
            r.consumed = pos;
            r.n_packets = n_pkt;
            r.n_symbols = (uint32_t)sym_used;
            a.results[st] = r;
        }
    }
}

// ---- frame generator --------------------------------------------------------------------------------------------
// Same engine as mod_kernels.cu (chirp_pipeline.cuh): genChirp (LoRa-SDR/ChirpGenerator.hpp) is a serial float recurrence
// with the phase carried through the whole frame, so a lane owns a frame for the recurrence and the warp shares the
// cos/sin work.  This kernel describes the frame of LoRaMod.cpp:138-208: 10 up-chirps, 2 sync chirps, 2 down-chirps, a
// quarter down-chirp, the data symbols; then the zero padding of :217-228.
constexpr int FM_WARPS = 4;

template <int PPW>
__global__ void __launch_bounds__(32 * FM_WARPS) frame_modulate_kernel(FrameModArgs a)
{
    __shared__ float ph[FM_WARPS][2][PPW][CHIRP_CHUNK + 1];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int pk0 = (blockIdx.x * FM_WARPS + warp) * PPW;
    if (pk0 >= a.n_packets) return; // whole warp
    const int pk = pk0 + lane;
    const bool live = lane < PPW && pk < a.n_packets;
    const int N = 1 << a.sf;
    const int NN = N * a.osr;
    const double pi = 3.14159265358979323846;
    ChirpSweep sw;
    sw.f_min = (float)__ddiv_rn(-pi, (double)a.osr);
    sw.f_max = (float)__ddiv_rn(pi, (double)a.osr);
    sw.f_step = (float)__ddiv_rn(2 * pi, (double)(N * a.osr * a.osr));
    const uint16_t* sym = a.symbols + (size_t)(live ? pk : pk0) * a.symbol_stride;
    const int segments = 15 + (int)a.n_symbols; // 10 up, 2 sync, 2 down, 1/4 down, data (LoRaMod.cpp:138-208)
    const int warp_live = min(PPW, a.n_packets - pk0);
    auto segment = [&](int s) {
        int v = 0;
        if (s == 10) v = (int)(a.sync >> 4) * 8;          // :150
        else if (s == 11) v = (int)(a.sync & 0xfu) * 8;   // :162
        else if (s >= 15) v = sym[s - 15];                // :205
        // freq = (2*M_PI*sym)/NN in double, narrowed to the float argument of genChirp
        return ChirpSegment{s == 14 ? NN / 4 : NN, (float)__ddiv_rn(__dmul_rn(2 * pi, (double)v), (double)NN), s >= 12 && s < 15};
    };
    const size_t n_samples = (size_t)(14 + a.n_symbols) * NN + NN / 4;
    float2* out = a.iq + (size_t)pk0 * a.packet_stride;
    chirp_pipeline<PPW>(segment, segments, n_samples, sw, a.amplitude, out, a.packet_stride, warp_live, ph[warp]);
    // :224-226 zero padding
    const size_t pad = (size_t)(a.padding ? a.padding : 1u) * NN; // the counter is incremented before the test (:219-221)
    for (int q = 0; q < warp_live; ++q)
        for (size_t i = lane; i < pad; i += 32) out[(size_t)q * a.packet_stride + n_samples + i] = make_float2(0.0f, 0.0f);
}

// One frame per CTA (a caller generating frames one at a time): chirp_one_packet of chirp_pipeline.cuh.
__global__ void __launch_bounds__(32 * CHIRP_ONE_WARPS) frame_modulate_few_kernel(FrameModArgs a)
{
    __shared__ float ring[CHIRP_ONE_SLOTS][32];
    const int pk = blockIdx.x;
    const int N = 1 << a.sf;
    const int NN = N * a.osr;
    const double pi = 3.14159265358979323846;
    ChirpSweep sw;
    sw.f_min = (float)__ddiv_rn(-pi, (double)a.osr);
    sw.f_max = (float)__ddiv_rn(pi, (double)a.osr);
    sw.f_step = (float)__ddiv_rn(2 * pi, (double)(N * a.osr * a.osr));
    const uint16_t* sym = a.symbols + (size_t)pk * a.symbol_stride;
    const int segments = 15 + (int)a.n_symbols;
    auto segment = [&](int s) {
        int v = 0;
        if (s == 10) v = (int)(a.sync >> 4) * 8;
        else if (s == 11) v = (int)(a.sync & 0xfu) * 8;
        else if (s >= 15) v = sym[s - 15];
        return ChirpSegment{s == 14 ? NN / 4 : NN, (float)__ddiv_rn(__dmul_rn(2 * pi, (double)v), (double)NN), s >= 12 && s < 15};
    };
    const size_t n_samples = (size_t)(14 + a.n_symbols) * NN + NN / 4;
    float2* out = a.iq + (size_t)pk * a.packet_stride;
    const size_t pad = (size_t)(a.padding ? a.padding : 1u) * NN;
    for (size_t i = threadIdx.x; i < pad; i += blockDim.x) out[n_samples + i] = make_float2(0.0f, 0.0f); // :224-226
    chirp_one_packet(segment, segments, n_samples, sw, a.amplitude, out, ring);
}

template <int SF>
cudaError_t launch_stream_sf(const StreamDetectArgs& a, int sm_count, cudaStream_t st)
{
    constexpr int SLOTS = StreamCfg<SF>::SLOTS;
    const size_t smem = sizeof(c64) * StreamCfg<SF>::SMEM_C64;
    static bool configured_dev[kMaxDevices] = {};
    bool& configured = configured_dev[current_device_slot()];
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(stream_detect_kernel<SF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, stream_detect_kernel<SF>, kCta, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    int grid = (a.n_streams + SLOTS - 1) / SLOTS;
    if (grid > sm_count * per_sm) grid = sm_count * per_sm;
    if (grid < 1) grid = 1;
    stream_detect_kernel<SF><<<grid, kCta, smem, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

} // namespace

void build_stream_tables(int sf, std::vector<float2>& up, std::vector<float2>& down, std::vector<float2>& fine)
{
    const size_t n = (size_t)1 << sf;
    up.resize(n);
    down.resize(n);
    fine.resize(n * STREAM_FINE_STEPS);
    float phase = (float)(-M_PI); // LoRaDemod.cpp:99
    double accum = 0.0;
    for (size_t i = 0; i < n; ++i) {
        accum += phase;
        const double c = std::cos(accum), s = std::sin(accum); // std::polar(1.0, accum)
        up[i] = make_float2((float)c, (float)(-s));
        down[i] = make_float2((float)c, (float)s);
        phase = (float)((double)phase + (2 * M_PI) / (double)n);
    }
    accum = 0.0;
    phase = (float)(2.0 * M_PI / (double)(n * STREAM_FINE_STEPS)); // :111
    for (size_t i = 0; i < n * STREAM_FINE_STEPS; ++i) {
        accum += phase;
        fine[i] = make_float2((float)std::cos(accum), (float)std::sin(accum));
    }
}

size_t frame_samples(int sf, int osr, size_t n_symbols, unsigned padding)
{
    const size_t nn = ((size_t)1 << sf) * (size_t)osr;
    return 14 * nn + nn / 4 + n_symbols * nn + (size_t)(padding ? padding : 1u) * nn; // LoRaMod.cpp:219-221: never fewer than one
}

cudaError_t launch_stream_detect(int sf, const StreamDetectArgs& a, int sm_count, cudaStream_t st)
{
    switch (sf) {
    case 7: return launch_stream_sf<7>(a, sm_count, st);
    case 8: return launch_stream_sf<8>(a, sm_count, st);
    case 9: return launch_stream_sf<9>(a, sm_count, st);
    case 10: return launch_stream_sf<10>(a, sm_count, st);
    case 11: return launch_stream_sf<11>(a, sm_count, st);
    case 12: return launch_stream_sf<12>(a, sm_count, st);
    default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_frame_modulate(const FrameModArgs& a, int sm_count, cudaStream_t st)
{
    if (a.n_packets < 8) {
        frame_modulate_few_kernel<<<(unsigned)a.n_packets, 32 * CHIRP_ONE_WARPS, 0, st>>>(a);
        ++g_launches;
        return cudaGetLastError();
    }
    const int ppw = chirp_packets_per_warp(a.n_packets, sm_count);
    const int per_block = ppw * FM_WARPS;
    const unsigned grid = (unsigned)((a.n_packets + per_block - 1) / per_block);
    if (ppw == 32) frame_modulate_kernel<32><<<grid, 32 * FM_WARPS, 0, st>>>(a);
    else if (ppw == 16) frame_modulate_kernel<16><<<grid, 32 * FM_WARPS, 0, st>>>(a);
    else frame_modulate_kernel<8><<<grid, 32 * FM_WARPS, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

} // namespace lora_b200
