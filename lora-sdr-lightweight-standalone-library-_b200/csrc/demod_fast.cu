// demod_fast.cu -- the receive hot path for SF7..SF12 on the register-blocked FFT (fft_blocked.cuh).
//
//   prepass_kernel   per packet: max-abs normalisation scan (LoRaDemod.cpp:59-77) and the offset estimate over
//                    the first windows (phy.cpp:78-145 / LoRaDemod.cpp:79-135), bit-faithful.
//   demod_kernel     per window: dechirp * CFO-rotate * window -> FFT -> |X|^2 argmax (phy.cpp:200-234,
//                    LoRaDemod.cpp:141-175), in two arithmetic flavours of ONE butterfly graph:
//                      FAITH  every operation rounded as the reference rounds it, cosf/sinf with glibc's bits;
//                             symbol index AND peak value are bit-identical to the CPU.
//                      fast   FMA-fused butterflies, and dechirp+rotation done in the phase domain (one MUFU
//                             sincos of  down_phase[i] + start + rate*i  per sample).  |X|^2 differs from the
//                             faithful value by ~1e-6 relative, so the argmax is only trusted when the top two
//                             bins are separated by more than `margin`; every other window is appended to a
//                             list and recomputed by the FAITH flavour in a second launch (a floating-point
//                             filter: the cheap evaluation decides whenever its error bound allows, the exact
//                             one otherwise).  Symbols are therefore bit-exact in both precisions; peak values
//                             are exact under PRECISION_EXACT and within 1e-4 relative under PRECISION_FAST.
#include "demod_fast.h"

#include <cstdlib>
#include <type_traits>

#include "bulk_copy.cuh"
#include "demod_kernels.h"
#include "exact_math.h"
#include "window_common.cuh"

namespace lora_b200 {

namespace {


// ---- per-sample preparation -----------------------------------------------------------------------------------
// Faithful: ((iq * down) * scale) * rot, then * window, each step rounded as phy.cpp:216-225 /
// LoRaDemod.cpp:73-75,153-160 / tests/performance_test.cpp:109-114 round it.  `down` points at the entry of
// the DOUBLED table (2N entries, so no index wrap) that multiplies in-window sample 0.
struct PrepFaith {
    const c64* down;     // nullptr = no dechirp
    const float* window; // nullptr = none
    float scale;
    bool rotate;
    float rate, start, one;
    PM_DEV c64 operator()(c64 v, int i) const
    {
        if (down) v = cmul<true>(v, __ldg(down + i), one);
        v = cscale(v, scale);
        if (rotate) {
            const float ph = __fadd_rn(start, __fmul_rn(rate, (float)i));
            float sn, cs;
            xm_sincosf(ph, &sn, &cs); // glibc cosf/sinf bits, exact_math.h
            v = cmul<true>(v, pk(cs, sn), one);
        }
        if (window) v = cscale(v, __ldg(window + i));
        return v;
    }
};

// Fast flavour: dechirp and CFO rotation are one unit-modulus factor per sample,
//     rot[i] = exp(j*(down_phase[ofs + i] + rate*i)),
// and the window's start phase is dropped: it multiplies the whole window by one unit complex number, which
// cannot change any |X[k]|^2.  rot[] therefore depends on the packet (rate) and on the table offset only, not on
// the window, so each thread computes its 16 factors once per run of windows (one MUFU sin/cos pair each) and
// keeps them in registers; per window a sample costs one fused complex multiply.  (Keeping them in shared memory
// instead shrinks L1 below the twiddle table and was measured slower.)  The normalisation scale is applied to the peak value instead (|X|^2 is homogeneous of degree 2).
template <int T>
PM_DEV void compute_rotors(c64 (&rot)[16], const float* down_phase, float rate, int t)
{
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        float ph = rate * (float)(t + T * j);
        if (down_phase) ph += __ldg(down_phase + T * j);
        float sn, cs;
        __sincosf(ph, &sn, &cs);
        rot[j] = pk(cs, sn);
    }
}

// What every thread of a window needs to know about its packet and window (phy.cpp:197-214, LoRaDemod.cpp:137-151).
struct PktParams {
    int t_off;
    float rate, scale;
};

// COHERENT: the estimate was written by another CTA of the SAME launch (receive_merged_kernel), so it must not come
// through the non-coherent path.
template <int SF, bool COHERENT = false>
PM_DEV PktParams packet_params(const BlockedDemodArgs& a, int p)
{
    constexpr int N = 1 << SF;
    const float4 e = COHERENT ? __ldcg(a.est + p) : __ldg(a.est + p);
    PktParams k;
    k.t_off = (int)roundf(e.y);                                                  // phy.cpp:197
    // :198  -2*pi*cfo / N; N is a power of two, so the division is the exact scaling by 1/N
    k.rate = __fmul_rn(__fmul_rn(-2.0f * kPiF, e.x), 1.0f / (float)N);
    k.scale = a.use_scale ? e.z : 1.0f;
    return k;
}

struct WinParams {
    size_t base;
    float start;
    int down_ofs;
};

template <int SF, bool FAITH, bool OSR1>
PM_DEV WinParams window_params(const BlockedDemodArgs& a, const PktParams& k, int s)
{
    constexpr int N = 1 << SF;
    WinParams q;
    const size_t step = OSR1 ? (size_t)N : (size_t)N * a.osr;
    size_t base = (size_t)s * step;                                              // phy.cpp:205-212
    if (k.t_off > 0) {
        if (base + (size_t)k.t_off + step <= a.count) base += (size_t)k.t_off;
    } else if (k.t_off < 0) {
        const size_t off = (size_t)(-(long long)k.t_off);
        if (off <= base) base -= off;
    }
    q.base = base;
    const float toff_f = (OSR1 || a.osr == 1) ? (float)k.t_off : __fdiv_rn((float)k.t_off, (float)a.osr);
    // :213-214.  The fused flavour only uses `start` to widen its trust margin, so it skips the 64-bit conversion
    // (float(s)*N equals float(s*N) whenever s < 2^24, and any value of that size serves the bound).
    q.start = FAITH ? __fmul_rn(k.rate, __fadd_rn((float)((size_t)s * N), toff_f))
                    : k.rate * ((float)s * (float)N + toff_f);
    q.down_ofs = a.dechirp_absolute ? (int)(base & (size_t)(N - 1)) : 0;
    return q;
}

// Sample loads.  LORA_V_NOALLOC = 1: ld.global.nc.L1::no_allocate -- the samples are read once, so they need not
// push the twiddle and phase tables (re-read by every window) out of L1.  Measured +1.0 / +2.9 / +0.7 % at SF9 / 11 / 12
// (profiles/r2_ab_kernel_variants.txt); no L2 prefetch at all costs 2-3 %, a two-window prefetch distance changes nothing.
#ifndef LORA_V_NOALLOC
#define LORA_V_NOALLOC 1
#endif
PM_DEV c64 load_sample(const c64* p)
{
#if LORA_V_NOALLOC
    c64 v;
    asm("ld.global.nc.L1::no_allocate.b64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
#else
    return __ldg(p);
#endif
}
// How many windows ahead the L2 prefetch of a run reaches (0 = none).
#ifndef LORA_V_PFDIST
#define LORA_V_PFDIST 1
#endif

// ---- the hot kernel ---------------------------------------------------------------------------------------------
// A CTA of 256 threads holds 256/T window slots; each slot walks its own sequence of window chunks (a chunk =
// `chunk` consecutive windows, normally of one packet, so the per-packet values are loaded once), and slots never
// wait for each other (sub-warp / named-barrier synchronisation only).  LIST = window numbers come from the
// recheck list instead of 0..n_windows-1.  OSR1 = the common configuration, oversampling ratio 1 and no Hann
// window (every BASELINE config): sample addresses become one pointer per window plus immediate offsets and
// the window multiply disappears; the other instantiation handles any osr and an optional window.
template <int SF>
struct DemodCfg {
    static constexpr int T = Blk<SF>::T;
    static constexpr int SLOTS = kCta / T;
    static constexpr int SMEM_C64 = SLOTS * Blk<SF>::WIN;
};

// A/B switches (scripts/ab_build.sh -D...): LORA_V_PREFETCH = the fused flavour keeps the NEXT window's samples of a
// run in registers (loads issued a whole window ahead; 2 CTAs/SM at <= 128 registers instead of 3 at 80).
#ifndef LORA_V_PREFETCH
#define LORA_V_PREFETCH 0
#endif
// Resident CTAs per SM of the fused flavour.  Measured on one box (profiles/r2_ab_kernel_variants.txt): two CTAs with up
// to 128 registers beat three with 80 at SF7-SF10 (+4.6 / +3.1 / +4.3 / +2.6 %: no spills, the stage geometry stays in
// registers) and lose at SF11 / SF12 (-1 / -6 %: one window per CTA there, so a third CTA is a third independent phase).
#ifndef LORA_V_CTAS
#define LORA_V_CTAS(SF) (LORA_V_PREFETCH ? 2 : ((SF) <= 10 ? 2 : 3))
#endif
// Passes whose twiddles stay in registers across the windows a slot walks (bit p = pass p; fft_blocked.cuh PreTw).
// Measured (profiles/r2_ab_kernel_variants.txt): SF8 +3.7 %, SF9 +1.7 %, SF10 +1.3 %, SF7 +0.8 % -- fewer L1 wavefronts,
// the unit that binds these sizes.  SF7/8: the 15 twiddles of the radix-16 pass; SF9/10: the 12 of the radix-4 pass.
// The second exchange of SF11 / SF12 inside a half-warp (fft_blocked.cuh HalfWarp): one barrier fewer per window.
// Measured on two boxes: SF11 +0.8 / +1.1 %, SF12 +1.3 / +1.3 % (profiles/r2_ab_kernel_variants.txt).
#ifndef LORA_V_HALFWARP
#define LORA_V_HALFWARP 1
#endif
#ifndef LORA_V_PRE
#define LORA_V_PRE(SF) ((SF) <= 8 ? 2 : ((SF) <= 10 ? 4 : 0))
#endif
// LORA_V_TMA = 1: the fused flavour's samples are staged by ONE bulk asynchronous copy per window (cp.async.bulk,
// completion on an mbarrier) into a per-slot buffer, issued by the slot's first thread right after the first exchange
// barrier of the window before (every thread has consumed the buffer by then, so no extra barrier is needed); the
// threads then read their sixteen samples with LDS instead of sixteen LDGs.  A window that does not start on a 16-byte
// boundary is copied with one sample of slack on either side.  
#ifndef LORA_V_TMA
#define LORA_V_TMA 0
#endif
struct Stage {
    c64* buf;                 // N + 2 samples, 16-byte aligned
    unsigned long long* bar;  // its mbarrier
    unsigned phase;           // parity the next wait expects
};

// One work item of the demod kernels: a run of consecutive windows of one packet (or, LIST, one window named by the
// recheck list), walked by the T threads of one slot.
template <int SF, bool FAITH, bool LIST, bool OSR1, int PREMASK, bool HW, bool COHERENT = false, bool TMA = false>
PM_DEV void demod_run(const BlockedDemodArgs& a, unsigned c, c64* S, Top* red, float* redw, const PreTw* pre, int slot, int t,
                      Stage* stage = nullptr)
{
    using B = Blk<SF>;
    constexpr int N = B::N, T = B::T;
    static_assert(!TMA || (!FAITH && !LIST && OSR1), "staged samples: fused flavour, osr 1 only");
    // a window can be staged when its 16-byte aligned superset stays inside the packet's samples
    auto stage_mis = [&](const c64* xw) { return (int)((reinterpret_cast<uintptr_t>(xw) >> 3) & 1u); };
    auto stage_ok = [&](const WinParams& w, const c64* xw) { return w.base + (size_t)N + (size_t)stage_mis(xw) <= a.count; };
    auto stage_issue = [&](const c64* xw) {
        const int mis = stage_mis(xw);
        const unsigned bytes = (unsigned)((N + 2 * mis) * sizeof(c64));
        mbar_expect_tx(stage->bar, bytes);
        bulk_load(stage->buf, xw - mis, bytes, stage->bar);
    };
    bool staged_in_flight = false; // the copy of the window about to be processed has been issued
    int p, s, s_end;
    if (LIST) { // one window, by number
        const unsigned wn = a.flag_list[c];
        p = a.div_total.d == 1 ? (int)wn : (int)(__umulhi(wn, a.div_total.magic) >> a.div_total.shift);
        s = (int)(wn - (unsigned)p * (unsigned)a.total);
        s_end = s + 1;
    } else {    // a run of windows inside one packet
        p = a.div_cpp.d == 1 ? (int)c : (int)(__umulhi(c, a.div_cpp.magic) >> a.div_cpp.shift);
        s = (int)(c - (unsigned)p * a.chunks_per_packet) * (int)a.chunk_len;
        s_end = min(s + (int)a.chunk_len, a.total);
    }
    const PktParams pk_ = packet_params<SF, COHERENT>(a, p);
    const float peak_scale = pk_.scale * pk_.scale;
    int rot_ofs = -1;
    c64 rot[16];
    c64 nxt[LORA_V_PREFETCH ? 16 : 1];
    bool have_nxt = false;
    for (; s < s_end; ++s) {
        const WinParams q = window_params<SF, FAITH, OSR1>(a, pk_, s);
        const unsigned wn = (unsigned)p * (unsigned)a.total + (unsigned)s;
        const c64* x = reinterpret_cast<const c64*>(a.iq) + (size_t)p * a.packet_stride + q.base;
        if (!TMA && !LIST && LORA_V_PFDIST > 0 && t == 0) { // samples after this window: a later window of the run, or of the packet
            const size_t span = OSR1 ? (size_t)N : (size_t)N * a.osr;
            const size_t ahead = span * LORA_V_PFDIST;
            const size_t left = a.count - min(a.count, q.base + ahead);
            l2_prefetch(x + ahead, sizeof(c64) * min(span, left));
        }
        c64 in[16], v[16];
        constexpr bool PF = LORA_V_PREFETCH && !FAITH && !LIST && OSR1;
        if (TMA && stage_ok(q, x)) {
            if (!staged_in_flight && t == 0) stage_issue(x); // first window of the run
            mbar_wait(stage->bar, stage->phase);
            stage->phase ^= 1u;
            const c64* sb = stage->buf + stage_mis(x) + t;
#pragma unroll
            for (int j = 0; j < 16; ++j) in[j] = sb[T * j];
        } else if (PF && have_nxt) {
#pragma unroll
            for (int j = 0; j < 16; ++j) in[j] = nxt[j];
        } else if (OSR1) {
            const c64* xt = x + t;
#pragma unroll
            for (int j = 0; j < 16; ++j) in[j] = load_sample(xt + T * j);
        } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) in[j] = __ldg(x + (size_t)(t + T * j) * a.osr);
        }
        if (PF) { // the next window of this run: its loads travel while this window is transformed
            have_nxt = s + 1 < s_end;
            if (have_nxt) {
                const WinParams qn = window_params<SF, FAITH, OSR1>(a, pk_, s + 1);
                const c64* xn = reinterpret_cast<const c64*>(a.iq) + (size_t)p * a.packet_stride + qn.base + t;
#pragma unroll
                for (int j = 0; j < 16; ++j) nxt[j] = __ldg(xn + T * j);
            }
        }
        float margin = 0.0f;
        if (FAITH) {
            const PrepFaith prep{a.dechirp ? a.down + q.down_ofs : nullptr, a.window, pk_.scale, true, pk_.rate, q.start, a.one};
#pragma unroll
            for (int j = 0; j < 16; ++j) in[j] = prep(in[j], t + T * j);
        } else {
            if (q.down_ofs != rot_ofs) { // first window of the run, or a window whose shift was not applied
                compute_rotors<T>(rot, a.dechirp ? a.down_phase + q.down_ofs + t : nullptr, pk_.rate, t);
                rot_ofs = q.down_ofs;
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                in[j] = cmul<false>(in[j], rot[j], 1.0f);
                // unconditional multiply where a window is possible: a select between packed values costs moves
                if (!OSR1) in[j] = cscale(in[j], a.window ? __ldg(a.window + t + T * j) : 1.0f);
            }
            // the reference rounds start + rate*i to float; the fast path does not reproduce that rounding, so
            // the trusted margin widens with the magnitude of the phase
            const float ph_max = fmaxf(fabsf(q.start), fabsf(q.start + pk_.rate * (float)N));
            margin = a.margin + 8.0f * ph_max * 1.1920929e-7f;
        }
        if (TMA) {
            // the next window of the run: its copy is issued behind the first exchange barrier of this one
            const c64* xn = x;
            bool next_ok = false;
            if (s + 1 < s_end) {
                const WinParams qn = window_params<SF, FAITH, OSR1>(a, pk_, s + 1);
                xn = reinterpret_cast<const c64*>(a.iq) + (size_t)p * a.packet_stride + qn.base;
                next_ok = stage_ok(qn, xn);
            }
            int syncs = 0;
            fft_blocked<SF, FAITH, PREMASK, HW>(in, v, S, a.twp, a.tw0, t, a.one,
                [&] { Group<T>::sync(slot); if (syncs++ == 0 && next_ok && t == 0) stage_issue(xn); }, PREMASK ? pre : nullptr);
            staged_in_flight = next_ok;
        } else
        fft_blocked<SF, FAITH, PREMASK, HW>(in, v, S, a.twp, a.tw0, t, a.one, [slot] { Group<T>::sync(slot); }, PREMASK ? pre : nullptr);
        const int tb = HW ? HalfWarp<SF>::u_of(t) : t; // whose bins this thread holds
        Top best{0.0f, 0, 0.0f};
        bool owner; // the thread that reports this window
        // measured: the winner-reports search is faster for windows of 2 or 4 warps (SF10 +3.6 %, SF11 +5 %) and
        // slower for one warp (every warp takes the winner branch) and for eight (SF12 -6 %, longer barrier tail)
        if (!FAITH && (T == 64 || T == 128)) {
            owner = window_top_fast<SF, (T >= 32 ? T : 32)>(v, t, slot, redw, best, tb);
        } else {
            best = window_top<SF, FAITH, T>(v, t, slot, red, tb);
            owner = t == 0;
        }
        if (owner) {
            if (!FAITH) {
                const bool trusted = best.second < best.v * (1.0f - margin); // false for 0, NaN, Inf and near-ties
                best.v *= peak_scale;
                if (!trusted) {
                    const unsigned k = atomicAdd(a.flag_count, 1u);
                    if (k < a.flag_cap) a.flag_list[k] = wn;
                }
            }
            if (!a.have_sync) a.symbols[(size_t)p * a.symbol_stride + s] = (uint16_t)best.i;
            else if (s >= 2) a.symbols[(size_t)p * a.symbol_stride + (s - 2)] = (uint16_t)best.i;
            else a.sync_raw[2 * (size_t)p + s] = (uint16_t)best.i;
            if (a.peaks) a.peaks[wn] = best.v;
        }
        // No barrier here.  The peak search is itself a rendezvous of the window's threads that comes after every
        // thread's last exchange load (warp shuffles / REDUX for T <= 32, the barrier inside window_top for
        // T > 32), so the exchange region is already free; and red[] is not written again before the next
        // window's first barrier, which thread 0 only reaches after it has read red[] here.
    }
}

template <int SF, bool FAITH, bool LIST, bool OSR1>
__global__ void __launch_bounds__(kCta, FAITH ? 2 : LORA_V_CTAS(SF)) demod_blocked_kernel(const __grid_constant__ BlockedDemodArgs a)
{
    using B = Blk<SF>;
    constexpr int T = B::T, SLOTS = DemodCfg<SF>::SLOTS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    c64* smem = reinterpret_cast<c64*>(smem_raw);
    __shared__ Top red[T > 32 ? kCta / 32 : 1];
    __shared__ float redw[T > 32 ? 2 * (kCta / 32) : 1];
    const int slot = threadIdx.x / T;
    const int t = threadIdx.x % T;
    c64* S = smem + slot * B::WIN;
    constexpr int PREMASK = (!FAITH && !LIST) ? (LORA_V_PRE(SF)) : 0;
    constexpr bool HW = LORA_V_HALFWARP && !FAITH && !LIST && HalfWarp<SF>::OK;
    PreTw pre[PREMASK ? 3 : 1];
    if (PREMASK & 2) preload_twiddles<SF, 1>(pre[1], t, a.twp);
    if (PREMASK & 4) preload_twiddles<SF, (B::NP == 3 ? 2 : 1)>(pre[2], t, a.twp);
    const unsigned n_items = LIST ? min(*a.flag_count, a.flag_cap) : a.n_packets * a.chunks_per_packet;
    if (LIST && blockIdx.x == 0 && threadIdx.x == 0 && n_items) atomicAdd(a.stat_rechecked, (unsigned long long)n_items);

    constexpr bool TMA = LORA_V_TMA && !FAITH && !LIST && OSR1;
    static_assert(!TMA || DemodCfg<SF>::SMEM_C64 % 2 == 0, "the staging buffers start 16-byte aligned");
    Stage stage{nullptr, nullptr, 0u};
    if (TMA) {
        c64* stage0 = smem + DemodCfg<SF>::SMEM_C64;                                  // after the exchange regions
        stage.buf = stage0 + slot * (B::N + 2);
        stage.bar = reinterpret_cast<unsigned long long*>(stage0 + SLOTS * (B::N + 2)) + slot;
        if (t == 0) mbar_init(stage.bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        __syncthreads();
    }
    for (unsigned c = blockIdx.x * SLOTS + slot; c < n_items; c += gridDim.x * SLOTS)
        demod_run<SF, FAITH, LIST, OSR1, PREMASK, HW, false, TMA>(a, c, S, red, redw, pre, slot, t, &stage);
}

// Windows of the fast scan form whose loads are issued together, as a function of the 16-byte loads a thread issues
// per window (PER: 2 at SF7, 4 at SF8, 8 from SF9 up).  Measured (profiles/r2_ab_merged_schedule.txt): one window at a
// time costs 22 % at SF7 and 5 % at SF8, two at a time cost 3 % at SF9 and above, so the rule keeps eight loads in
// flight per thread everywhere.
#ifndef LORA_V_SCAN_WUNROLL
#define LORA_V_SCAN_WUNROLL(PER) ((PER) >= 8 ? 1 : ((PER) >= 4 ? 2 : 4))
#endif
// ---- prepass: normalisation scan + offset estimate, one packet per group of TPK threads --------------------------
template <int SF>
struct PreCfg {
    static constexpr int T = Blk<SF>::T;
    static constexpr int TPK = T > 32 ? T : 32;  // threads per packet
    static constexpr int SLOTS = kCta / TPK;
    static constexpr int SMEM_C64 = SLOTS * Blk<SF>::WIN;
};

// The prepass of ONE packet by the TPK threads of one slot: normalisation scan, then the offset estimate; thread u == 0
// writes est[p].  Returns with the slot's threads past their last barrier.
template <int SF>
PM_DEV void prepass_packet(const PrepassArgs& a, int p, c64* S, Top* red, float* redf, int* redp, int slot, int u)
{
    using B = Blk<SF>;
    constexpr int N = B::N, T = B::T, TPK = PreCfg<SF>::TPK;
    const c64* x = reinterpret_cast<const c64*>(a.iq) + (size_t)p * a.packet_stride;
    // scale_from_est: the light scan kernel (scan_kernels.cu) already wrote scale and max_amp for this packet
    float4 e = a.scale_from_est ? a.est[p] : make_float4(0.0f, 0.0f, 1.0f, 0.0f);

    // LoRaDemod.cpp:59-77: max over all samples of max(|re|,|im|), comparisons written so a NaN never wins.
    // In fused-dechirp mode the values examined are iq * down[n mod N] (performance_test.cpp:109-114).
    if (a.do_maxabs) {
        float mx = 0.0f;
        const bool aligned = (reinterpret_cast<uintptr_t>(x) & 15u) == 0;
        // Exact form: the reference's two comparisons per sample, so a NaN real part drops the sample and a NaN
        // imaginary part drops only itself.
        auto take = [&](c64 v, size_t n) {
            if (a.dechirp) v = cmul<true>(v, __ldg(a.down + (n & (size_t)(N - 1))), a.one);
            const float2 f = up(v);
            const float r = fabsf(f.x), im = fabsf(f.y);
            const float m = (r < im) ? im : r;
            if (m > mx) mx = m;
        };
        auto scan_exact = [&](size_t first) {
            if (aligned && (first & 1) == 0) {
                // 16-byte aligned: two samples per load, eight loads in flight per thread (a max is order-free)
                const ulonglong2* x2 = reinterpret_cast<const ulonglong2*>(x);
                const size_t n2 = a.count >> 1;
#pragma unroll 8
                for (size_t i = (first >> 1) + u; i < n2; i += TPK) {
                    const ulonglong2 v = __ldg(x2 + i);
                    take(v.x, 2 * i);
                    take(v.y, 2 * i + 1);
                }
                if ((a.count & 1) && u == 0) take(__ldg(x + a.count - 1), a.count - 1);
            } else {
#pragma unroll 8
                for (size_t n = first + u; n < a.count; n += TPK) take(__ldg(x + n), n);
            }
        };
        // Fast form, whole windows of an aligned packet: for finite data the two comparisons are a plain maximum
        // (one 3-input FMNMX with |.| modifiers per sample), and the thread's down-chirp entries repeat every window
        // (its sample positions are congruent mod N), so they live in registers.  Whether the data IS finite is
        // tracked by one packed FMA per sample (0*v + acc turns NaN as soon as any examined component is NaN or
        // Inf); in that case the packet is rescanned in the exact form, so the result never differs.
        constexpr int PER = N / (2 * TPK); // 16-byte pairs per thread per window
        constexpr int kScanWUnroll = LORA_V_SCAN_WUNROLL(PER); // (#pragma unroll takes a constant expression, not a macro)
        bool poisoned = false;
        size_t done = 0;
        if (aligned && PER >= 1) {
            const ulonglong2* x2 = reinterpret_cast<const ulonglong2*>(x);
            const size_t n_win = a.count / N;
            c64 nan_acc = pk(0.0f, 0.0f);
            const c64 zero2 = pk(0.0f, 0.0f);
            auto windows = [&](auto dech) {
                ulonglong2 dn[PER > 0 ? PER : 1];
                if (dech.value) {
#pragma unroll
                    for (int k = 0; k < PER; ++k) dn[k] = __ldg(reinterpret_cast<const ulonglong2*>(a.down) + u + k * TPK);
                }
                auto one_sample = [&](c64 v, c64 d) {
                    if (dech.value) v = cmul<true>(v, d, a.one);
                    nan_acc = fma2(v, zero2, nan_acc);
                    const float2 f = up(v);
                    mx = fmaxf(mx, fmaxf(fabsf(f.x), fabsf(f.y)));
                };
#pragma unroll kScanWUnroll
                for (size_t w = 0; w < n_win; ++w) {
                    const ulonglong2* xw = x2 + w * (size_t)(N / 2) + u;
                    ulonglong2 v[PER > 0 ? PER : 1];
#pragma unroll
                    for (int k = 0; k < PER; ++k) v[k] = __ldg(xw + k * TPK);
#pragma unroll
                    for (int k = 0; k < PER; ++k) {
                        one_sample(v[k].x, dn[k].x);
                        one_sample(v[k].y, dn[k].y);
                    }
                }
            };
            if (a.dechirp) windows(std::true_type{});
            else windows(std::false_type{});
            done = n_win * (size_t)N;
            const float2 na = up(nan_acc);
            poisoned = na.x != na.x || na.y != na.y;
        }
        scan_exact(done); // the samples after the last whole window (or everything, for an unaligned packet)
        auto reduce = [&] {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { const float y = __shfl_xor_sync(0xffffffffu, mx, o); if (y > mx) mx = y; }
            bool any = __any_sync(0xffffffffu, poisoned);
            if (TPK > 32) {
                if ((u & 31) == 0) { redf[slot * (TPK / 32) + (u >> 5)] = mx; redp[slot * (TPK / 32) + (u >> 5)] = any ? 1 : 0; }
                Group<TPK>::sync(slot);
                mx = redf[slot * (TPK / 32)];
                any = redp[slot * (TPK / 32)] != 0;
#pragma unroll
                for (int w = 1; w < TPK / 32; ++w) {
                    if (redf[slot * (TPK / 32) + w] > mx) mx = redf[slot * (TPK / 32) + w];
                    any = any || redp[slot * (TPK / 32) + w] != 0;
                }
                Group<TPK>::sync(slot); // redf/redp may be written again by the rescan
            }
            return any;
        };
        if (reduce()) { // NaN or Inf somewhere in the packet: the reference's comparison order matters
            mx = 0.0f;
            poisoned = false;
            scan_exact(0);
            reduce();
        }
        e.w = mx;
        e.z = (mx > 1.0f) ? __fdiv_rn(1.0f, mx) : 1.0f;
    }

    // phy.cpp:78-145 and its twin LoRaDemod.cpp:79-135
    float sum_index = 0.0f, phase_diff = 0.0f, prev_phase = 0.0f;
    bool have_prev = false;
    unsigned sum_t = 0;
    const size_t step = (size_t)N * a.osr;
    const PrepFaith prep{a.dechirp ? a.down : nullptr, a.window, a.use_scale ? e.z : 1.0f, false, 0.0f, 0.0f, a.one};
    for (int s = 0; s < a.symbols; ++s) {
        float best_p = -1e30f, best_f = 0.0f;
        int best_idx = 0;
        unsigned best_t = 0;
        float2 best_bin = make_float2(0.0f, 0.0f);
        for (unsigned ph = 0; ph < (unsigned)a.osr; ++ph) {
            Top pk_{0.0f, 0, 0.0f};
            if (u < T) {
                c64 in[16], v[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) in[j] = prep(__ldg(x + s * step + ph + (size_t)(u + T * j) * a.osr), u + T * j);
                fft_blocked<SF, true>(in, v, S, a.twp, a.tw0, u, a.one, [slot] { Group<T>::sync(slot); });
                pk_ = thread_top<SF, true>(v, u);
                Group<T>::sync(slot); // all exchange loads done: the region now receives the natural-order spectrum
#pragma unroll
                for (int i = 0; i < 16; ++i) S[bin_of<SF>(u, i)] = v[i];
            }
            Group<TPK>::sync(slot);
            pk_ = group_top<TPK>(pk_, u, slot, red);
            if (u == 0) { // LoRaDetector.hpp:60-71
                const float fundamental = __fsqrt_rn(pk_.v);
                const float pw = __fsub_rn(__fmul_rn(20.0f, log10f(fundamental)), a.power_scale);
                const float2 Lb = up(S[pk_.i > 0 ? pk_.i - 1 : N - 1]);
                const float2 Rb = up(S[pk_.i < N - 1 ? pk_.i + 1 : 0]);
                const float left = xm_hypotf(Lb.x, Lb.y), right = xm_hypotf(Rb.x, Rb.y);
                const double demon = __dsub_rn(__dsub_rn(__dmul_rn(2.0, (double)fundamental), (double)right), (double)left);
                const float fi = (demon == 0.0) ? 0.0f : (float)__ddiv_rn(__dmul_rn(0.5, (double)__fsub_rn(right, left)), demon);
                if (pw > best_p || (a.legacy_tie && pw == best_p && pk_.i < best_idx)) {
                    best_p = pw; best_idx = pk_.i; best_f = fi; best_t = ph; best_bin = up(S[pk_.i]);
                }
            }
            Group<TPK>::sync(slot);
        }
        if (u == 0) {
            sum_t += best_t;
            sum_index = __fadd_rn(sum_index, __fadd_rn((float)best_idx, best_f));
            const float phase = xm_atan2f(best_bin.y, best_bin.x); // std::arg
            if (have_prev) {
                float d = __fsub_rn(phase, prev_phase);
                const float two_pi = __fmul_rn(2.0f, kPiF);
                while (d > kPiF) d = __fsub_rn(d, two_pi);
                while (d < -kPiF) d = __fadd_rn(d, two_pi);
                phase_diff = __fadd_rn(phase_diff, d);
            }
            prev_phase = phase;
            have_prev = true;
        }
    }
    if (u == 0) {
        if (a.symbols > 0) {
            const float Sn = (float)a.symbols;
            const float avg_index = __fdiv_rn(sum_index, Sn);
            const float cfo_coarse = __fdiv_rn(avg_index, (float)N);
            float cfo_fine = 0.0f;
            if (a.symbols > 1)
                cfo_fine = __fdiv_rn(__fdiv_rn(phase_diff, (float)(a.symbols - 1)), __fmul_rn(__fmul_rn(2.0f, kPiF), (float)N));
            e.x = __fadd_rn(cfo_coarse, cfo_fine);
            const float frac = __fsub_rn(avg_index, floorf(__fadd_rn(avg_index, 0.5f)));
            const float avg_t = __fdiv_rn((float)sum_t, Sn);
            e.y = __fsub_rn(avg_t, __fmul_rn(__fmul_rn(frac, (float)N), (float)a.osr));
        }
        a.est[p] = e;
    }
}

template <int SF>
__global__ void __maxnreg__(96) prepass_blocked_kernel(const __grid_constant__ PrepassArgs a)
{
    using B = Blk<SF>;
    constexpr int TPK = PreCfg<SF>::TPK, SLOTS = PreCfg<SF>::SLOTS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    c64* smem = reinterpret_cast<c64*>(smem_raw);
    __shared__ Top red[kCta / 32];
    __shared__ float redf[kCta / 32];
    __shared__ int redp[kCta / 32];
    const int slot = threadIdx.x / TPK;
    const int u = threadIdx.x % TPK;
    c64* S = smem + slot * B::WIN;
    const int p = blockIdx.x * SLOTS + slot;
    if (p >= a.n_packets) return; // whole slots leave together; slots never synchronise with each other
    prepass_packet<SF>(a, p, S, red, redf, redp, slot, u);
}

// ---- prepass and demod in ONE persistent launch -----------------------------------------------------------------
// The prepass is HBM-bound (it only compares), the fused demod is issue-bound (it barely loads): run back to back they
// leave the other resource idle, 19 + 33 ms of a 53 ms bench step.  Here both are work items of one queue that the
// resident CTAs pull from, so at any moment some CTAs of every SM scan while the others transform.  The dependency --
// a packet's windows need its estimate -- is carried by one word per packet: the prepass item publishes est[p] with a
// release store, the demod item acquires it (and reads est[] through the coherent path).  The queue is ordered so that
// the prepass of a packet group is fetched `lead` groups before the group's demod items, which makes the wait a
// formality; it can never deadlock, because a prepass item waits for nothing and every fetched item belongs to a
// running CTA.  Same arithmetic, same results: the two bodies are the functions the separate kernels call.
//
// Queue layout, in groups of kMergedGroup packets:  [prepass of groups 0..lead-1]  then for g = 0, 1, ...:
// [prepass of group g+lead (pre_items entries)] [demod of group g (dem_items entries)].
constexpr int kMergedGroup = 32; // packets per group: a multiple of the prepass slots per CTA of every SF

#ifndef LORA_V_MERGED_CTAS
#define LORA_V_MERGED_CTAS(SF) 2
#endif

PM_DEV unsigned ld_acquire(const unsigned* p)
{
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
PM_DEV void st_release(unsigned* p, unsigned v) { asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

template <int SF, bool OSR1>
__global__ void __launch_bounds__(kCta, LORA_V_MERGED_CTAS(SF)) receive_merged_kernel(const __grid_constant__ MergedArgs m)
{
    using B = Blk<SF>;
    constexpr int T = B::T, SLOTS = DemodCfg<SF>::SLOTS, TPK = PreCfg<SF>::TPK, PSLOTS = PreCfg<SF>::SLOTS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    c64* smem = reinterpret_cast<c64*>(smem_raw);
    __shared__ Top red[kCta / 32];
    __shared__ float redw[2 * (kCta / 32)];
    __shared__ float redf[kCta / 32];
    __shared__ int redp[kCta / 32];
    __shared__ unsigned item_s;
    constexpr int PREMASK = LORA_V_PRE(SF);
    constexpr bool HW = LORA_V_HALFWARP && HalfWarp<SF>::OK;
    const unsigned head = m.lead * m.pre_items, per_group = m.pre_items + m.dem_items;
    for (;;) {
        if (threadIdx.x == 0) item_s = atomicAdd(m.queue, 1u);
        __syncthreads();
        const unsigned item = item_s;
        __syncthreads(); // item_s is rewritten only after every thread has read it
        if (item >= m.n_items) break;
        bool is_pre;
        unsigned idx, group = 0;
        if (item < head) { is_pre = true; idx = item; }
        else {
            const unsigned j = item - head, r = j % per_group;
            group = j / per_group;
            is_pre = r < m.pre_items;
            idx = is_pre ? (group + m.lead) * m.pre_items + r : r - m.pre_items;
        }
        if (is_pre) {
            const int slot = threadIdx.x / TPK, u = threadIdx.x % TPK;
            const long long p = (long long)idx * PSLOTS + slot;
            if (p < m.pre.n_packets) {
                prepass_packet<SF>(m.pre, (int)p, smem + slot * B::WIN, red, redf, redp, slot, u);
                if (u == 0) st_release(m.ready + p, 1u); // est[p] was stored by this thread just above
            }
        } else {
            const int slot = threadIdx.x / T, t = threadIdx.x % T;
            const unsigned run = idx * SLOTS + slot, runs_per_group = kMergedGroup * m.dem.chunks_per_packet;
            const unsigned c = group * runs_per_group + run;
            const unsigned p = m.dem.div_cpp.d == 1 ? c : (__umulhi(c, m.dem.div_cpp.magic) >> m.dem.div_cpp.shift);
            if (run < runs_per_group && p < m.dem.n_packets) {
                while (ld_acquire(m.ready + p) == 0u) __nanosleep(200);
                PreTw pre[PREMASK ? 3 : 1];
                if (PREMASK & 2) preload_twiddles<SF, 1>(pre[1], t, m.dem.twp);
                if (PREMASK & 4) preload_twiddles<SF, (B::NP == 3 ? 2 : 1)>(pre[2], t, m.dem.twp);
                demod_run<SF, false, false, OSR1, PREMASK, HW, true>(m.dem, c, smem + slot * B::WIN, red, redw, pre, slot, t);
            }
        }
    }
}

} // namespace

// ---- launchers ---------------------------------------------------------------------------------------------------
FastDiv make_fastdiv(unsigned d)
{
    FastDiv f{d, 0u, 0u};
    if (d <= 1) return f;
    unsigned lg = 0;
    while ((1ull << lg) < d) ++lg;                    // ceil(log2 d)
    const unsigned k = 31 + lg;
    f.magic = (unsigned)(((1ull << k) + d - 1) / d);   // ceil(2^k / d) < 2^32
    f.shift = k - 32;
    return f;
}

bool blocked_supported(int sf) { return sf >= 7 && sf <= 12; }

template <int SF>
static void fill_tw0_sf(const float2* h_tw, Tw0* out)
{
    using B = Blk<SF>;
    const int Fa = B::F(B::pa(0));
    for (int k = 0; k < 4; ++k)
        for (int q = 0; q < 4; ++q) out->w[k * 4 + q] = h_tw[(k * Fa * q) & (B::N - 1)];
}

template <int SF>
static void stage_twiddles_sf(const float2* h_tw, std::vector<float2>& out)
{
    using B = Blk<SF>;
    out.assign(2 * B::TWP, make_float2(1.0f, 0.0f));
    for (int s = 0; s < B::L; ++s)
        for (int q = 1; q < B::radix(s); ++q)
            for (int k = 0; k < B::m(s); ++k) out[B::toff(s) + (q - 1) * B::m(s) + k] = h_tw[k * B::F(s) * q];
    // second copy at +TWP: for the half-warp mapping the entries of the last pass (stages 0 and 1, [0, toff(2))) are
    // reordered inside every block of T so that physical thread t finds the entry of logical thread u_of(t) at +t
    for (int i = 0; i < B::TWP; ++i) out[B::TWP + i] = out[i];
    if (HalfWarp<SF>::OK)
        for (int base = 0; base + B::T <= B::toff(2); base += B::T)
            for (int t = 0; t < B::T; ++t) out[B::TWP + base + t] = out[base + HalfWarp<SF>::u_of(t)];
}

std::vector<float2> build_stage_twiddles(int sf, const float2* h_tw)
{
    std::vector<float2> out;
    switch (sf) {
    case 7: stage_twiddles_sf<7>(h_tw, out); break;
    case 8: stage_twiddles_sf<8>(h_tw, out); break;
    case 9: stage_twiddles_sf<9>(h_tw, out); break;
    case 10: stage_twiddles_sf<10>(h_tw, out); break;
    case 11: stage_twiddles_sf<11>(h_tw, out); break;
    case 12: stage_twiddles_sf<12>(h_tw, out); break;
    default: break;
    }
    return out;
}

void fill_tw0(int sf, const float2* h_tw, Tw0* out)
{
    switch (sf) {
    case 7: fill_tw0_sf<7>(h_tw, out); break;
    case 8: fill_tw0_sf<8>(h_tw, out); break;
    case 9: fill_tw0_sf<9>(h_tw, out); break;
    case 10: fill_tw0_sf<10>(h_tw, out); break;
    case 11: fill_tw0_sf<11>(h_tw, out); break;
    case 12: fill_tw0_sf<12>(h_tw, out); break;
    default: break;
    }
}

template <int SF>
static cudaError_t launch_prepass_sf(const PrepassArgs& a, cudaStream_t st)
{
    constexpr int SLOTS = PreCfg<SF>::SLOTS;
    constexpr size_t smem = sizeof(c64) * PreCfg<SF>::SMEM_C64;
    static bool attr_done_dev[kMaxDevices] = {};
    bool& attr_done = attr_done_dev[current_device_slot()];
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(prepass_blocked_kernel<SF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    const int grid = (a.n_packets + SLOTS - 1) / SLOTS;
    prepass_blocked_kernel<SF><<<grid, kCta, smem, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

template <int SF, bool FAITH, bool LIST, bool OSR1>
static cudaError_t launch_demod_one(BlockedDemodArgs a, int sm_count, cudaStream_t st)
{
    constexpr int SLOTS = DemodCfg<SF>::SLOTS;
    // exchange regions, plus (fast flavour) 16 rotation factors per thread
    constexpr bool TMA = LORA_V_TMA && !FAITH && !LIST && OSR1;
    constexpr size_t smem = sizeof(c64) * DemodCfg<SF>::SMEM_C64 +
                            (TMA ? (sizeof(c64) * (Blk<SF>::N + 2) + sizeof(unsigned long long)) * SLOTS : 0);
    static int per_sm_dev[kMaxDevices] = {};
    int& per_sm = per_sm_dev[current_device_slot()];
    if (per_sm == 0) {
        cudaError_t e = cudaFuncSetAttribute(demod_blocked_kernel<SF, FAITH, LIST, OSR1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, demod_blocked_kernel<SF, FAITH, LIST, OSR1>, kCta, smem);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) per_sm = 1;
    }
    // whole waves of resident CTAs; slots stride over the work items.  max_ctas_per_sm leaves room on every SM for
    // the prepass of the next sub-batch, which runs concurrently on another stream (capi.cu)
    const int use = (a.max_ctas_per_sm > 0 && a.max_ctas_per_sm < per_sm) ? a.max_ctas_per_sm : per_sm;
    long long grid = (long long)sm_count * use;
    const long long slots = grid * SLOTS;
    long long items;
    if (LIST) {
        items = (long long)sm_count * SLOTS; // the list is short; its length is only known on the device
    } else {
        // cut every packet into runs of consecutive windows: as long as possible (the per-packet work is done once
        // per run), but at least ~16 runs per slot so the slots finish together
        long long cpp = (16 * slots + a.n_packets - 1) / a.n_packets;
        if (cpp > a.total) cpp = a.total;
        if (cpp < 1) cpp = 1;
        a.chunk_len = (unsigned)((a.total + cpp - 1) / cpp);
        a.chunks_per_packet = (unsigned)((a.total + a.chunk_len - 1) / a.chunk_len);
        a.div_cpp = make_fastdiv(a.chunks_per_packet);
        items = (long long)a.n_packets * a.chunks_per_packet;
    }
    const long long need = (items + SLOTS - 1) / SLOTS;
    if (grid > need) grid = need;
    if (grid < 1) grid = 1;
    demod_blocked_kernel<SF, FAITH, LIST, OSR1><<<(unsigned)grid, kCta, smem, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

template <int SF, bool OSR1>
static cudaError_t launch_demod_osr(const BlockedDemodArgs& a, int precision, int sm_count, cudaStream_t st)
{
    if (precision == PRECISION_EXACT) return launch_demod_one<SF, true, false, OSR1>(a, sm_count, st);
    cudaError_t e = cudaMemsetAsync(a.flag_count, 0, sizeof(unsigned), st);
    if (e != cudaSuccess) return e;
    e = launch_demod_one<SF, false, false, OSR1>(a, sm_count, st);
    if (e != cudaSuccess) return e;
    // the recheck list is short (about 1e-3 of the windows on noise); one CTA per SM is plenty, and the grid
    // size must not depend on a device-side count
    return launch_demod_one<SF, true, true, OSR1>(a, sm_count, st);
}

template <int SF>
static cudaError_t launch_demod_sf(const BlockedDemodArgs& a, int precision, int sm_count, cudaStream_t st)
{
    return (a.osr == 1 && a.window == nullptr) ? launch_demod_osr<SF, true>(a, precision, sm_count, st)
                      : launch_demod_osr<SF, false>(a, precision, sm_count, st);
}

#define LORA_BLOCKED_SWITCH(fn, ...)          \
    switch (sf) {                             \
    case 7: return fn<7>(__VA_ARGS__);        \
    case 8: return fn<8>(__VA_ARGS__);        \
    case 9: return fn<9>(__VA_ARGS__);        \
    case 10: return fn<10>(__VA_ARGS__);      \
    case 11: return fn<11>(__VA_ARGS__);      \
    case 12: return fn<12>(__VA_ARGS__);      \
    default: return cudaErrorInvalidValue;    \
    }

template <int SF, bool OSR1>
static cudaError_t launch_merged_osr(MergedArgs m, int sm_count, cudaStream_t st)
{
    constexpr int SLOTS = DemodCfg<SF>::SLOTS, PSLOTS = PreCfg<SF>::SLOTS;
    constexpr size_t smem = sizeof(c64) * DemodCfg<SF>::SMEM_C64; // >= the prepass's: it has no more slots than the demod
    static_assert(kMergedGroup % PSLOTS == 0, "a group is a whole number of prepass items");
    static int per_sm_dev[kMaxDevices] = {};
    int& per_sm = per_sm_dev[current_device_slot()];
    if (per_sm == 0) {
        cudaError_t e = cudaFuncSetAttribute(receive_merged_kernel<SF, OSR1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, receive_merged_kernel<SF, OSR1>, kCta, smem);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) per_sm = 1;
    }
    BlockedDemodArgs& a = m.dem;
    const long long grid = (long long)sm_count * per_sm; // persistent: exactly the CTAs that are resident together
    const long long slots = grid * SLOTS;
    long long cpp = (16 * slots + a.n_packets - 1) / a.n_packets; // as launch_demod_one: >= ~16 runs per slot
    if (cpp > a.total) cpp = a.total;
    if (cpp < 1) cpp = 1;
    a.chunk_len = (unsigned)((a.total + cpp - 1) / cpp);
    static const int env_chunk = getenv("LORA_B200_MERGED_CHUNK") ? atoi(getenv("LORA_B200_MERGED_CHUNK")) : 0;
    static const double env_lead = getenv("LORA_B200_MERGED_LEAD") ? atof(getenv("LORA_B200_MERGED_LEAD")) : 2.0;
    if (env_chunk > 0 && (unsigned)env_chunk < a.chunk_len) a.chunk_len = (unsigned)env_chunk;
    a.chunks_per_packet = (unsigned)((a.total + a.chunk_len - 1) / a.chunk_len);
    a.div_cpp = make_fastdiv(a.chunks_per_packet);
    m.pre_items = kMergedGroup / PSLOTS;
    m.dem_items = (kMergedGroup * a.chunks_per_packet + SLOTS - 1) / SLOTS;
    const unsigned n_groups = (a.n_packets + kMergedGroup - 1) / kMergedGroup;
    // the prepass of a group is queued ~two waves of CTAs ahead of the group's demod items
    unsigned lead = (unsigned)((long long)(env_lead * grid + m.pre_items + m.dem_items - 1) / (m.pre_items + m.dem_items));
    if (lead < 1) lead = 1;
    if (lead > n_groups) lead = n_groups;
    m.lead = lead;
    m.n_items = lead * m.pre_items + n_groups * (m.pre_items + m.dem_items);
    cudaError_t e = cudaMemsetAsync(m.queue, 0, sizeof(unsigned) * ((size_t)a.n_packets + 1), st); // queue head + ready[]
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(a.flag_count, 0, sizeof(unsigned), st);
    if (e != cudaSuccess) return e;
    receive_merged_kernel<SF, OSR1><<<(unsigned)grid, kCta, smem, st>>>(m);
    ++g_launches;
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    return launch_demod_one<SF, true, true, OSR1>(a, sm_count, st); // the recheck list, as after the separate kernels
}

template <int SF>
static cudaError_t launch_merged_sf(const MergedArgs& m, int sm_count, cudaStream_t st)
{
    return (m.dem.osr == 1 && m.dem.window == nullptr) ? launch_merged_osr<SF, true>(m, sm_count, st)
                                                       : launch_merged_osr<SF, false>(m, sm_count, st);
}

cudaError_t launch_prepass_blocked(int sf, const PrepassArgs& a, cudaStream_t st) { LORA_BLOCKED_SWITCH(launch_prepass_sf, a, st) }

cudaError_t launch_demod_blocked(int sf, const BlockedDemodArgs& a, int precision, int sm_count, cudaStream_t st)
{
    LORA_BLOCKED_SWITCH(launch_demod_sf, a, precision, sm_count, st)
}

cudaError_t launch_receive_merged(int sf, const MergedArgs& m, int sm_count, cudaStream_t st)
{
    LORA_BLOCKED_SWITCH(launch_merged_sf, m, sm_count, st)
}

} // namespace lora_b200
