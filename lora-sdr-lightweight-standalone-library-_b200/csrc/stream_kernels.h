// stream_kernels.h -- the continuous-stream receive front end and the full-frame generator of the upstream
// Pothos blocks (LoRa-SDR/LoRaDemod.cpp:145-327, LoRa-SDR/LoRaMod.cpp:109-238), one stream / one frame per slot.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include <vector>

#include "fft_blocked.cuh"

namespace lora_b200 {

constexpr int STREAM_FINE_STEPS = 128; // LoRaDemod.cpp:70

struct StreamPacket { // == lora_b200_stream_packet
    uint64_t start;       // sample index of the first data window
    uint32_t first_symbol, n_symbols;
    int32_t freq_error;   // averaged down-chirp bin error (LoRaDemod.cpp:242-270)
    uint32_t reserved;
};

struct StreamResult { // == lora_b200_stream_result
    uint64_t consumed;    // samples the state machine consumed
    uint32_t n_packets;   // packets found (may exceed the capacity; only the first pkt_cap are stored)
    uint32_t n_symbols;   // symbols stored
};

struct StreamDetectArgs {
    const float2* iq;
    size_t stream_stride, count;
    size_t osr;       // every osr-th sample is used (plain decimation, as phy.cpp:216 indexes iq[base + i*osr])
    int n_streams;
    const c64* twp;
    Tw0 tw0;
    float one;
    const c64* up;    // N: conj of the table entry (LoRaDemod.cpp:98-108)
    const c64* down;  // N
    const c64* fine;  // N * 128 (LoRaDemod.cpp:110-117)
    float power_scale, thresh;
    unsigned mtu, sync;
    int frame_header; // 0: symbols until mtu or squelch (LoRaDemod.cpp:286-310); 1 / 2: after eight data symbols read the
                      // deployed chain's explicit header (2 = LDRO) and stop at the frame length it announces
    int16_t* symbols;
    size_t sym_cap;
    StreamPacket* packets;
    unsigned pkt_cap;
    StreamResult* results;
};

struct FrameModArgs {
    const uint16_t* symbols;
    size_t n_symbols, symbol_stride;
    float2* iq;
    size_t packet_stride;
    int n_packets, sf, osr;
    float amplitude;
    unsigned sync, padding;
};

// Host-built tables of LoRaDemod.cpp:98-117 (double phase accumulation of a float frequency, polar in double).
void build_stream_tables(int sf, std::vector<float2>& up, std::vector<float2>& down, std::vector<float2>& fine);
size_t frame_samples(int sf, int osr, size_t n_symbols, unsigned padding);

cudaError_t launch_stream_detect(int sf, const StreamDetectArgs& a, int sm_count, cudaStream_t st);
cudaError_t launch_frame_modulate(const FrameModArgs& a, int sm_count, cudaStream_t st);

} // namespace lora_b200
