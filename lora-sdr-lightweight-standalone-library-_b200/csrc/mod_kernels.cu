// mod_kernels.cu -- batched chirp modulator with the reference's float32 phase-accumulator recurrence,
// and a counter-based AWGN generator for synthetic inputs.
#include "mod_kernels.h"

#include "demod_kernels.h" // g_launches
#include "exact_math.h"

namespace lora_b200 {

// genChirp (include/lora_phy/ChirpGenerator.hpp:24-50) is a serial float recurrence over every sample of a
// packet: f += fStep (wrapped at fMax), phase += f, and the phase is carried from symbol to symbol
// (LoRaMod.cpp:13,25-41) with one double-precision wrap per symbol (:48).  Float addition is not
// associative, so only packet-level parallelism reproduces it.  Layout: a warp owns 32 packets and walks them in
// chunks of 32 samples.
//   * Recurrence: lane l advances packet l, two dependent float adds per sample -- a latency chain, not work.  It
//     runs ONE CHUNK AHEAD, interleaved step by step with the evaluation below (double-buffered in shared memory),
//     so its latency is hidden behind the expensive part instead of preceding it.
//   * Evaluation: phases -> (ampl*cos, ampl*sin) with glibc's cosf/sinf bits (exact_math.h, double-precision
//     polynomials), lanes across the 32 consecutive samples of one packet so every store is a full 256-byte line.
//     Four packets are evaluated side by side with the branch-free xm_sincosf_bf: four independent dependency
//     chains per thread, because with one lane per packet a batch of P packets only ever has P/32 warps to hide
//     latency with (14 per SM at 65,536 packets).
// Algorithmic bytes: 8*N*osr written per symbol (+2 read); bound: issue slots of the fp64-exact sincos (~70
// instructions per sample), about 0.6 of the HBM write roofline.
constexpr int MOD_CHUNK = 32;
constexpr int MOD_WARPS = 4;

struct ChirpState { // one packet's genChirp / lora_modulate state, held by one lane
    float f, phase;
    int s, j;       // symbol index (sync words first) and sample index inside it
};

__global__ void __launch_bounds__(32 * MOD_WARPS) modulate_kernel(ModulateArgs a)
{
    __shared__ float ph[MOD_WARPS][2][32][MOD_CHUNK + 1];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int pk0 = (blockIdx.x * MOD_WARPS + warp) * 32;
    if (pk0 >= a.n_packets) return; // whole warp
    const int pk = pk0 + lane;
    const bool live = pk < a.n_packets;
    const int N = 1 << a.sf;
    const int step = N * a.osr;
    const double pi = 3.14159265358979323846;
    // :27-29 -- evaluated in double, stored as float
    const float f_min = (float)__ddiv_rn(__dmul_rn(-pi, (double)a.bw_scale), (double)a.osr);
    const float f_max = (float)__ddiv_rn(__dmul_rn(pi, (double)a.bw_scale), (double)a.osr);
    const float f_step = (float)__ddiv_rn(__dmul_rn(2 * pi, (double)a.bw_scale), (double)(N * a.osr * a.osr));
    const float span = __fsub_rn(f_max, f_min);
    const float pi_f = 3.14159274101257324f;
    const float denom = __fmul_rn((float)N, (float)a.osr);
    const uint16_t* sym = a.symbols + (size_t)(live ? pk : pk0) * a.symbol_stride;
    const int total = (int)a.n_symbols + 2;
    const int warp_live = min(32, a.n_packets - pk0);
    const size_t n_samples = (size_t)total * step;
    const size_t n_chunks = (n_samples + MOD_CHUNK - 1) / MOD_CHUNK;

    ChirpState st{0.0f, 0.0f, 0, 0}; // LoRaMod.cpp:13: phase = 0
    // One step of the recurrence: the phase of the next sample of this lane's packet (any value past the packet's end).
    auto next_phase = [&]() -> float {
        if (st.s >= total) return 0.0f;
        if (st.j == 0) {
            const unsigned v = st.s == 0 ? a.sw0 : st.s == 1 ? a.sw1 : (unsigned)sym[st.s - 2];
            // LoRaMod.cpp:24,29,36: ((2*pi_f * v) * bw_scale) / (N * osr)
            const float f0 = __fdiv_rn(__fmul_rn(__fmul_rn(2.0f * pi_f, (float)v), a.bw_scale), denom);
            st.f = __fadd_rn(f_min, f0);
        }
        st.f = __fadd_rn(st.f, f_step);                 // ChirpGenerator.hpp:41-45 (up-chirp branch)
        if (st.f > f_max) st.f = __fsub_rn(st.f, span);
        st.phase = __fadd_rn(st.phase, st.f);
        const float out = st.phase;
        if (++st.j == step) { // :48 -- wrap once per chirp, in double, narrowed back into the float accumulator
            const double pd = (double)st.phase;
            st.phase = (float)__dsub_rn(pd, __dmul_rn(__dmul_rn(floor(__ddiv_rn(pd, 2 * pi)), 2.0), pi));
            st.j = 0;
            ++st.s;
        }
        return out;
    };

    for (int j = 0; j < MOD_CHUNK; ++j) ph[warp][0][lane][j] = next_phase();
    for (size_t k = 0; k < n_chunks; ++k) {
        __syncwarp();
        const float(*cur)[MOD_CHUNK + 1] = ph[warp][k & 1];
        float(*nxt)[MOD_CHUNK + 1] = ph[warp][(k + 1) & 1];
        const size_t at = k * MOD_CHUNK + lane;           // this lane's sample in every packet of the warp
        const bool in_range = at < n_samples;
#pragma unroll 1
        for (int q = 0; q < 32; q += 4) {
            float y[4], sn[4], cs[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) y[i] = cur[q + i][lane];
#pragma unroll
            for (int i = 0; i < 4; ++i) nxt[lane][q + i] = next_phase(); // the recurrence, one chunk ahead
            // std::polar -> cosf/sinf.  One warp-uniform choice of range reduction for all 128 arguments of this step
            // (phases inside a chirp are mostly on one side of glibc's |y| < 120 test); never a divergent branch.
            int large = 0;
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
                    a.iq[(size_t)(pk0 + q + i) * a.packet_stride + at] =
                        make_float2(__fmul_rn(a.amplitude, cs[i]), __fmul_rn(a.amplitude, sn[i]));
        }
    }
}

// ---- AWGN --------------------------------------------------------------------------------------------------
// Philox-4x32-10 keyed by the seed, counter = (sample index, packet index): the noise added to a sample
// depends on nothing else, so any launch geometry or sharding across GPUs gives the same stream.
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1)
{
    const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
    const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
    c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}

__global__ void __launch_bounds__(256) awgn_kernel(AwgnArgs a)
{
    const size_t total = (size_t)a.n_packets * a.count;
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (size_t)gridDim.x * blockDim.x) {
        const size_t p = g / a.count, n = g - p * a.count;
        uint32_t c[4] = {(uint32_t)n, (uint32_t)(n >> 32), (uint32_t)p, 0x4c6f5261u};
        uint32_t k0 = (uint32_t)a.seed, k1 = (uint32_t)(a.seed >> 32);
#pragma unroll
        for (int r = 0; r < 10; ++r) { philox_round(c, k0, k1); k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
        // Box-Muller on two 32-bit uniforms in (0, 1]
        const float u1 = ((float)c[0] + 1.0f) * 2.3283064365386963e-10f;
        const float u2 = (float)c[1] * 2.3283064365386963e-10f;
        const float rad = sqrtf(-2.0f * logf(u1));
        float sn, cs;
        sincospif(2.0f * u2, &sn, &cs);
        // tests/awgn_sweep_gtest.cpp:76-77: sigma = 10^(-snr/20), per component sigma/sqrt(2)
        const float sigma = exp10f(-a.snr_db[p] * 0.05f) * 0.70710678118654752f;
        float2* q = a.iq + p * a.packet_stride + n;
        float2 v = *q;
        v.x += sigma * rad * cs;
        v.y += sigma * rad * sn;
        *q = v;
    }
}

cudaError_t launch_modulate(const ModulateArgs& a, cudaStream_t st)
{
    const int per_block = 32 * MOD_WARPS;
    modulate_kernel<<<(a.n_packets + per_block - 1) / per_block, 32 * MOD_WARPS, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

cudaError_t launch_awgn(const AwgnArgs& a, int sm_count, cudaStream_t st)
{
    const size_t total = (size_t)a.n_packets * a.count;
    size_t blocks = (total + 255) / 256;
    const size_t cap = (size_t)sm_count * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    awgn_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

} // namespace lora_b200
