// chirp_pipeline.cuh -- the warp-level engine shared by the two chirp generators (mod_kernels.cu: lora_modulate,
// stream_kernels.cu: the Pothos frame generator).
//
// genChirp (include/lora_phy/ChirpGenerator.hpp:24-50, LoRa-SDR/ChirpGenerator.hpp) is a serial float recurrence over
// every sample of a packet: f += fStep (wrapped at fMax), phase +-= f, the phase carried from chirp to chirp with one
// double-precision wrap per chirp (:48).  Float addition is not associative, so only packet-level parallelism
// reproduces it.  A warp owns 32 packets and walks them in chunks of 32 samples:
//   * Recurrence: lane l advances packet l, two or three dependent float operations per sample -- a latency chain, not
//     work.  It runs ONE CHUNK AHEAD, interleaved step by step with the evaluation below (double-buffered in shared
//     memory), so its latency hides behind the expensive part instead of preceding it.  Chirp bookkeeping (start
//     frequency, the wrap, the end of the packet) happens once per chunk whenever the chunk lies inside one chirp --
//     always, for SF >= 7 -- and per sample otherwise.
//   * Evaluation: phases -> (ampl*cos, ampl*sin) with glibc's cosf/sinf bits (exact_math.h, double-precision
//     polynomials), lanes across the 32 consecutive samples of one packet so every store is a full 256-byte line.
//     Four packets are evaluated side by side with the branch-free xm_sincosf_paths: four independent dependency chains
//     per thread, because with one lane per packet a batch of P packets only ever has P/32 warps to hide latency with
//     (14 per SM at 65,536 packets), and ONE warp-uniform choice of range reduction per step (phases inside a chirp are
//     mostly on one side of glibc's |y| < 120 test), never a divergent branch.
// The caller describes the packet as a list of segments (chirps): segment(s) -> length, start frequency offset, sweep
// direction.  All packets of a launch share the lengths, so every branch on them is uniform.
#pragma once

#include <cstdlib>

#include <cuda_runtime.h>
#include <stdint.h>

#include "exact_math.h"

namespace lora_b200 {

constexpr int CHIRP_CHUNK = 32;

struct ChirpSegment {
    int len;    // samples
    float f0;   // start frequency offset (radians per sample), added to f_min
    bool down;  // phase -= f instead of += f
};

struct ChirpSweep {
    float f_min, f_max, f_step; // ChirpGenerator.hpp:27-29, each evaluated in double and stored as float
};

// ph: this warp's two chunk buffers, [2][PPW][CHIRP_CHUNK + 1].  out + q * packet_stride = packet q of the warp.
// (Calls of fewer than 8 packets -- the reference's per-packet API -- use chirp_one_packet below instead.)
// PPW = packets per warp (32, 16 or 8): the first PPW lanes run a recurrence each, all 32 lanes evaluate.  Fewer
// packets per warp means more warps for the same batch -- the evaluation is a long dependent chain, so the kernel needs
// a full complement of resident warps (chirp_packets_per_warp picks PPW from the batch size).
template <int PPW = 32, typename SegFn>
__device__ __forceinline__ void chirp_pipeline(const SegFn& segment, int n_segments, size_t n_samples, const ChirpSweep sw,
                                               float amplitude, float2* out, size_t packet_stride, int warp_live,
                                               float (*ph)[PPW][CHIRP_CHUNK + 1])
{
    const int lane = threadIdx.x & 31;
    const double pi = 3.14159265358979323846;
    const float span = __fsub_rn(sw.f_max, sw.f_min);
    float f = 0.0f, phase = 0.0f; // LoRaMod.cpp:13 / LoRa-SDR LoRaMod.cpp:131: the phase starts at zero
    int s = 0, j = 0, len = 0;
    bool down = false;
    auto load_segment = [&]() {
        if (s < n_segments) {
            const ChirpSegment g = segment(s);
            len = g.len;
            down = g.down;
            f = __fadd_rn(sw.f_min, g.f0);
        }
    };
    auto step = [&]() -> float { // ChirpGenerator.hpp:33-46
        f = __fadd_rn(f, sw.f_step);
        if (f > sw.f_max) f = __fsub_rn(f, span);
        phase = down ? __fsub_rn(phase, f) : __fadd_rn(phase, f);
        return phase;
    };
    auto finish_segment = [&]() { // :48 -- wrap once per chirp, in double, narrowed back into the float accumulator
        const double pd = (double)phase;
        phase = (float)__dsub_rn(pd, __dmul_rn(__dmul_rn(floor(__ddiv_rn(pd, 2 * pi)), 2.0), pi));
        j = 0;
        ++s;
        load_segment();
    };
    auto slow_step = [&]() -> float { // a chunk that straddles chirps, or lies past the packet's end
        if (s >= n_segments) return 0.0f;
        const float r = step();
        if (++j == len) finish_segment();
        return r;
    };
    load_segment();
    while (s < n_segments && len == 0) finish_segment(); // zero-length segments produce nothing
    bool fast = false;
    auto begin_chunk = [&]() { fast = s < n_segments && len - j >= CHIRP_CHUNK; };
    auto end_chunk = [&]() {
        if (fast) {
            j += CHIRP_CHUNK;
            if (j == len) finish_segment();
        }
    };

    begin_chunk();
    const bool owner = PPW == 32 || lane < PPW; // lanes beyond PPW own no packet and no row of the buffers
    for (int i = 0; i < CHIRP_CHUNK; ++i) {
        const float v = fast ? step() : slow_step();
        if (owner) ph[0][lane][i] = v;
    }
    end_chunk();
    const size_t n_chunks = (n_samples + CHIRP_CHUNK - 1) / CHIRP_CHUNK;
    for (size_t k = 0; k < n_chunks; ++k) {
        __syncwarp();
        const float(*cur)[CHIRP_CHUNK + 1] = ph[k & 1];
        float(*nxt)[CHIRP_CHUNK + 1] = ph[(k + 1) & 1];
        const size_t at = k * CHIRP_CHUNK + lane; // this lane's sample in every packet of the warp
        const bool in_range = at < n_samples;
        begin_chunk();
        constexpr int RS = 4 * 32 / PPW; // recurrence steps per iteration: the chunk's 32 spread over PPW / 4 iterations
        static_assert(PPW == 32 || PPW == 16 || PPW == 8, "packets per warp");
#pragma unroll 1
        for (int q = 0; q < PPW; q += 4) {
            float y[4], sn[4], cs[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) y[i] = cur[q + i][lane];
            const int r0 = (q >> 2) * RS;
            if (fast) { // the recurrence, one chunk ahead
#pragma unroll
                for (int i = 0; i < RS; ++i) { const float v = step(); if (owner) nxt[lane][r0 + i] = v; }
            } else {
#pragma unroll
                for (int i = 0; i < RS; ++i) { const float v = slow_step(); if (owner) nxt[lane][r0 + i] = v; }
            }
            int large = 0; // std::polar -> cosf/sinf
#pragma unroll
            for (int i = 0; i < 4; ++i) large += xm_is_large_arg(y[i]);
            if (__all_sync(0xffffffffu, large == 0)) {
#pragma unroll
                for (int i = 0; i < 4; ++i) xm_sincosf_paths(y[i], &sn[i], &cs[i], 1, 0);
            } else if (__all_sync(0xffffffffu, large == 4)) {
#pragma unroll
                for (int i = 0; i < 4; ++i) xm_sincosf_paths(y[i], &sn[i], &cs[i], 0, 1);
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) xm_sincosf_paths(y[i], &sn[i], &cs[i], 1, 1);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (in_range && q + i < warp_live)
                    out[(size_t)(q + i) * packet_stride + at] = make_float2(__fmul_rn(amplitude, cs[i]), __fmul_rn(amplitude, sn[i]));
        }
        end_chunk();
    }
}

// ---- one packet per CTA: the per-packet API of the reference (lora_modulate called for one packet) -----------------
// With a single packet the recurrence IS the critical path (three dependent float operations per sample), and in the
// warp-per-packets engine above it shares one instruction stream with a ~100-instruction sincos per chunk.  Here a CTA
// of four warps takes one packet: warp 0 only runs the recurrence -- all 32 lanes redundantly, lane l keeping sample l of
// the chunk, so no transpose is needed -- and hands each chunk of 32 phases to one of three evaluating warps through a
// ring of one slot per evaluating warp in shared memory.  Hand-over is the producer/consumer form of the named barriers: the producer
// `bar.arrive`s on a slot's "full" barrier after storing, the consumer `bar.sync`s on it before loading, and the other
// way round on the slot's "empty" barrier.  Per chunk the packet then costs max(recurrence, evaluation / 3).
// Warps per packet: the recurrence warp plus the evaluating ones.  Measured: 8 is no faster than 4 (2326 vs 2172 us for
// an SF12 packet), and with the evaluation compiled out the time does not change either: the recurrence warp is the
// limit -- ~15 cycles of the nominal clock per sample for a 4-instruction step, i.e. a single small kernel on an
// otherwise idle GPU runs at the idle clock.
#ifndef LORA_V_ONE_WARPS
#define LORA_V_ONE_WARPS 4
#endif
// one slot per evaluating warp: slot k % SLOTS is always evaluated by warp 1 + k % SLOTS (2 * SLOTS barrier ids <= 14)
constexpr int CHIRP_ONE_WARPS = LORA_V_ONE_WARPS, CHIRP_ONE_SLOTS = LORA_V_ONE_WARPS - 1;

__device__ __forceinline__ void chirp_bar_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void chirp_bar_arrive(int id) { asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory"); }

// ring: [CHIRP_ONE_SLOTS][32] floats of shared memory.  Called by all 32 * CHIRP_ONE_WARPS threads of the CTA.
template <typename SegFn>
__device__ __forceinline__ void chirp_one_packet(const SegFn& segment, int n_segments, size_t n_samples, const ChirpSweep sw,
                                                 float amplitude, float2* out, float (*ring)[32])
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t n_chunks = (n_samples + CHIRP_CHUNK - 1) / CHIRP_CHUNK;
    constexpr int FULL = 1, EMPTY = 1 + CHIRP_ONE_SLOTS; // barrier ids 1..SLOTS and SLOTS+1..2*SLOTS (0 is __syncthreads)
    static_assert(2 * CHIRP_ONE_SLOTS <= 14 && CHIRP_ONE_SLOTS >= 1, "named barriers 1..15");
    if (warp == 0) {
        // the recurrence of chirp_pipeline, every lane running the same packet
        const double pi = 3.14159265358979323846;
        const float span = __fsub_rn(sw.f_max, sw.f_min);
        float f = 0.0f, phase = 0.0f;
        int s = 0, j = 0, len = 0;
        bool down = false;
        unsigned flip = 0u; // sign bit applied to f before it enters the phase sum
        // The descriptor of a chirp (a symbol read from global memory and a division) is fetched one chirp ahead: on
        // this warp's critical path its latency would cost as much as a chunk at SF7.
        ChirpSegment ahead = n_segments > 0 ? segment(0) : ChirpSegment{0, 0.0f, false};
        auto load_segment = [&]() {
            if (s < n_segments) {
                const ChirpSegment g = ahead;
                len = g.len;
                down = g.down;
                flip = down ? 0x80000000u : 0u;
                f = __fadd_rn(sw.f_min, g.f0);
                if (s + 1 < n_segments) ahead = segment(s + 1);
            }
        };
        // phase -= f is phase += (-f) bit for bit, so the direction is a sign flip of f (off the phase chain), not a
        // select between two sums on it
        auto signed_f = [&]() { return __uint_as_float(__float_as_uint(f) ^ flip); };
        auto step = [&]() -> float { // ChirpGenerator.hpp:33-46
            f = __fadd_rn(f, sw.f_step);
            if (f > sw.f_max) f = __fsub_rn(f, span);
            phase = __fadd_rn(phase, signed_f());
            return phase;
        };
        auto finish_segment = [&]() { // :48
            const double pd = (double)phase;
            phase = (float)__dsub_rn(pd, __dmul_rn(__dmul_rn(floor(__ddiv_rn(pd, 2 * pi)), 2.0), pi));
            j = 0;
            ++s;
            load_segment();
        };
        auto slow_step = [&]() -> float {
            if (s >= n_segments) return 0.0f;
            const float r = step();
            if (++j == len) finish_segment();
            return r;
        };
        load_segment();
        while (s < n_segments && len == 0) finish_segment();
        for (size_t k = 0; k < n_chunks; ++k) {
            const int slot = (int)(k % CHIRP_ONE_SLOTS);
            float mine = 0.0f;
            if (s < n_segments && len - j >= CHIRP_CHUNK) { // the chunk lies inside one chirp
                // The frequency wraps once per chirp.  When it provably cannot within this chunk -- the float sum after
                // 32 steps stays below f + 40 * f_step -- the compare-and-subtract leaves the dependency chain.
                if (__fadd_rn(f, __fmul_rn(40.0f, sw.f_step)) < sw.f_max) {
#pragma unroll
                    for (int i = 0; i < CHIRP_CHUNK; ++i) {
                        f = __fadd_rn(f, sw.f_step);
                        phase = __fadd_rn(phase, signed_f());
                        if (i == lane) mine = phase;
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < CHIRP_CHUNK; ++i) { const float v = step(); if (i == lane) mine = v; }
                }
                j += CHIRP_CHUNK;
                if (j == len) finish_segment();
            } else {
#pragma unroll 1
                for (int i = 0; i < CHIRP_CHUNK; ++i) { const float v = slow_step(); if (i == lane) mine = v; }
            }
            if (k >= (size_t)CHIRP_ONE_SLOTS) chirp_bar_sync(EMPTY + slot); // its previous occupant has been read
            ring[slot][lane] = mine;
            chirp_bar_arrive(FULL + slot);
        }
    } else {
        for (size_t k = (size_t)(warp - 1); k < n_chunks; k += CHIRP_ONE_WARPS - 1) {
            const int slot = (int)(k % CHIRP_ONE_SLOTS);
            chirp_bar_sync(FULL + slot);
            const float y = ring[slot][lane];
            chirp_bar_arrive(EMPTY + slot);
            float sn, cs; // std::polar -> cosf/sinf, one warp-uniform choice of range reduction
            const int large = xm_is_large_arg(y);
            if (__all_sync(0xffffffffu, large == 0)) xm_sincosf_paths(y, &sn, &cs, 1, 0);
            else if (__all_sync(0xffffffffu, large == 1)) xm_sincosf_paths(y, &sn, &cs, 0, 1);
            else xm_sincosf_paths(y, &sn, &cs, 1, 1);
            const size_t at = k * CHIRP_CHUNK + lane;
            if (at < n_samples) out[at] = make_float2(__fmul_rn(amplitude, cs), __fmul_rn(amplitude, sn));
        }
    }
}

// Packets per warp for a batch: 32 when that still gives every SM its ~24 resident warps, else 16, else 8.
inline int chirp_packets_per_warp(long long n_packets, int sm_count)
{
    static const int forced = getenv("LORA_B200_PPW") ? atoi(getenv("LORA_B200_PPW")) : 0; // experiments: 8, 16 or 32
    if (forced == 8 || forced == 16 || forced == 32) return forced;
    const long long want = (long long)sm_count * 24;
    if (n_packets / 32 >= want) return 32;
    if (n_packets / 16 >= want) return 16;
    return 8;
}

} // namespace lora_b200
