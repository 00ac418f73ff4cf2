// demod_kernels.h -- argument blocks and host-callable launchers of the receive kernels.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include <atomic>

#include "lora_phy_b200.h"

namespace lora_b200 {

// Per-packet scratch shared by the three demod kernels: x = cfo, y = time_offset, z = scale, w = max_amp.
struct MaxAbsArgs {
    const float2* iq;
    size_t packet_stride, count;
    const float2* down; // non-null: examine iq * down[n mod N] (fused caller-dechirp)
    int N;
    float4* est;
};

struct EstimateArgs {
    const float2* iq;
    size_t packet_stride;
    int symbols;       // min(total, 2)
    int osr;
    int dechirp;       // multiply by the down-chirp first (fused caller-dechirp only)
    int use_scale;     // apply est[p].z
    int legacy_tie;    // LoRaDemod.cpp:101 lowest-index rule on equal power
    const float2* tw;
    const float2* down;
    const float* window;
    float power_scale; // float(20*log10(N)), LoRaDetector.hpp:29
    float4* est;
};

struct DemodArgs {
    const float2* iq;
    size_t packet_stride, count;
    long long n_windows; // n_packets * total
    int total;           // windows per packet, sync included
    int osr;
    int dechirp, use_scale, have_sync;
    int dechirp_absolute; // fused caller-dechirp: table index = absolute sample position mod N
    const float2* tw;
    const float2* down;
    const float* window;
    const float4* est;
    uint16_t* symbols;
    size_t symbol_stride;
    uint16_t* sync_raw;  // 2 per packet
    float* peaks;        // optional, n_windows
};

struct FinalizeArgs {
    const float4* est;
    const uint16_t* sync_raw;
    lora_b200_pkt_metrics* metrics;
    int n_packets, sf, have_sync, legacy, n_symbols;
};

cudaError_t launch_maxabs(const MaxAbsArgs& a, int n_packets, cudaStream_t st);
cudaError_t launch_estimate(int sf, const EstimateArgs& a, int n_packets, cudaStream_t st);
cudaError_t launch_demod(int sf, const DemodArgs& a, int sm_count, cudaStream_t st);
cudaError_t launch_finalize(const FinalizeArgs& a, cudaStream_t st);
// compensate_offsets (phy.cpp:147-176): rotate sample n by rate*n, then shift by `offset` samples with zero fill.
struct CompensateArgs {
    const float2* in;
    float2* out;
    size_t count;
    float rate;      // -2*pi*cfo / (N*osr), formed on the host in the reference's float order
    long long offset; // round(time_offset); 0 when it would not be applied (|offset| >= count)
};
cudaError_t launch_compensate(const CompensateArgs& a, cudaStream_t st);
cudaError_t launch_fill_est(float4* est, int n_packets, cudaStream_t st);

// kernels launched by this library (bench.py's gpu_launches); incremented from whichever host thread launches
extern std::atomic<unsigned long long> g_launches;

// Function attributes and occupancy are properties of (kernel, device): launchers cache them per device ordinal.
constexpr int kMaxDevices = 64;
inline int current_device_slot()
{
    int d = 0;
    cudaGetDevice(&d);
    return (d >= 0 && d < kMaxDevices) ? d : 0;
}

// SDR sample ingest: interleaved int16 (I, Q) -> float32 * scale.  With scale = 2^-15 the conversion is exact.
cudaError_t launch_convert_i16(const short2* in, float2* out, size_t n, float scale, int sm_count, cudaStream_t st);
cudaError_t launch_convert_i8(const char2* in, float2* out, size_t n, float scale, int sm_count, cudaStream_t st);

} // namespace lora_b200
