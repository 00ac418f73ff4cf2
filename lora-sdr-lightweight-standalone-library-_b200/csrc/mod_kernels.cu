// mod_kernels.cu -- batched chirp modulator with the reference's float32 phase-accumulator recurrence,
// and a counter-based AWGN generator for synthetic inputs.
#include "mod_kernels.h"

#include "demod_kernels.h" // g_launches
#include "chirp_pipeline.cuh"
#include "exact_math.h"

namespace lora_b200 {

// lora_modulate (src/phy/LoRaMod.cpp:8-43): two sync chirps and one up-chirp per symbol, every one a genChirp call
// that shares one float phase.  The recurrence / evaluation engine is chirp_pipeline.cuh; this kernel only describes the
// packet: segment s = sync word 0, sync word 1, then the data symbols, each N*osr samples.
// Algorithmic bytes: 8*N*osr written per symbol (+2 read); bound: issue slots of the fp64-exact sincos (~90
// instructions per sample), i.e. about a third of the HBM write roofline.
constexpr int MOD_WARPS = 4;

// Minimum resident CTAs per SM asked of ptxas (A/B switch: 8 means 64 registers per thread).
#ifndef LORA_V_MOD_CTAS
#define LORA_V_MOD_CTAS 1
#endif
template <int PPW>
__global__ void __launch_bounds__(32 * MOD_WARPS, LORA_V_MOD_CTAS) modulate_kernel(ModulateArgs a)
{
    __shared__ float ph[MOD_WARPS][2][PPW][CHIRP_CHUNK + 1];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int pk0 = (blockIdx.x * MOD_WARPS + warp) * PPW;
    if (pk0 >= a.n_packets) return; // whole warp
    const int pk = pk0 + lane;
    const bool live = lane < PPW && pk < a.n_packets;
    const int N = 1 << a.sf;
    const int step = N * a.osr;
    const double pi = 3.14159265358979323846;
    ChirpSweep sw; // ChirpGenerator.hpp:27-29 -- evaluated in double, stored as float
    sw.f_min = (float)__ddiv_rn(__dmul_rn(-pi, (double)a.bw_scale), (double)a.osr);
    sw.f_max = (float)__ddiv_rn(__dmul_rn(pi, (double)a.bw_scale), (double)a.osr);
    sw.f_step = (float)__ddiv_rn(__dmul_rn(2 * pi, (double)a.bw_scale), (double)(N * a.osr * a.osr));
    const float pi_f = 3.14159274101257324f;
    const float denom = __fmul_rn((float)N, (float)a.osr);
    const uint16_t* sym = a.symbols + (size_t)(live ? pk : pk0) * a.symbol_stride;
    const int total = (int)a.n_symbols + 2;
    auto segment = [&](int s) {
        const unsigned v = s == 0 ? a.sw0 : s == 1 ? a.sw1 : (unsigned)sym[s - 2];
        // LoRaMod.cpp:24,29,36: ((2*pi_f * v) * bw_scale) / (N * osr)
        return ChirpSegment{step, __fdiv_rn(__fmul_rn(__fmul_rn(2.0f * pi_f, (float)v), a.bw_scale), denom), false};
    };
    chirp_pipeline<PPW>(segment, total, (size_t)total * step, sw, a.amplitude, a.iq + (size_t)pk0 * a.packet_stride,
                        a.packet_stride, min(PPW, a.n_packets - pk0), ph[warp]);
}

// lora_modulate of ONE packet per CTA (the reference's per-packet API): chirp_one_packet of chirp_pipeline.cuh.
__global__ void __launch_bounds__(32 * CHIRP_ONE_WARPS) modulate_few_kernel(ModulateArgs a)
{
    __shared__ float ring[CHIRP_ONE_SLOTS][32];
    const int pk = blockIdx.x;
    const int N = 1 << a.sf;
    const int step = N * a.osr;
    const double pi = 3.14159265358979323846;
    ChirpSweep sw;
    sw.f_min = (float)__ddiv_rn(__dmul_rn(-pi, (double)a.bw_scale), (double)a.osr);
    sw.f_max = (float)__ddiv_rn(__dmul_rn(pi, (double)a.bw_scale), (double)a.osr);
    sw.f_step = (float)__ddiv_rn(__dmul_rn(2 * pi, (double)a.bw_scale), (double)(N * a.osr * a.osr));
    const float pi_f = 3.14159274101257324f;
    const float denom = __fmul_rn((float)N, (float)a.osr);
    const uint16_t* sym = a.symbols + (size_t)pk * a.symbol_stride;
    const int total = (int)a.n_symbols + 2;
    auto segment = [&](int s) {
        const unsigned v = s == 0 ? a.sw0 : s == 1 ? a.sw1 : (unsigned)sym[s - 2];
        return ChirpSegment{step, __fdiv_rn(__fmul_rn(__fmul_rn(2.0f * pi_f, (float)v), a.bw_scale), denom), false};
    };
    chirp_one_packet(segment, total, (size_t)total * step, sw, a.amplitude, a.iq + (size_t)pk * a.packet_stride, ring);
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

cudaError_t launch_modulate(const ModulateArgs& a, int sm_count, cudaStream_t st)
{
    if (a.n_packets < 8) { // a packet or a few: one CTA each, recurrence and evaluation on different warps
        modulate_few_kernel<<<(unsigned)a.n_packets, 32 * CHIRP_ONE_WARPS, 0, st>>>(a);
        ++g_launches;
        return cudaGetLastError();
    }
    const int ppw = chirp_packets_per_warp(a.n_packets, sm_count);
    const int per_block = ppw * MOD_WARPS;
    const unsigned grid = (unsigned)((a.n_packets + per_block - 1) / per_block);
    if (ppw == 32) modulate_kernel<32><<<grid, 32 * MOD_WARPS, 0, st>>>(a);
    else if (ppw == 16) modulate_kernel<16><<<grid, 32 * MOD_WARPS, 0, st>>>(a);
    else modulate_kernel<8><<<grid, 32 * MOD_WARPS, 0, st>>>(a);
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
