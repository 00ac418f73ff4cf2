// mod_kernels.h -- chirp modulator and synthetic-channel kernels.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace lora_b200 {

struct ModulateArgs { // lora_modulate, src/phy/LoRaMod.cpp:8-43
    const uint16_t* symbols;
    size_t n_symbols, symbol_stride;
    float2* iq;
    size_t packet_stride;
    int n_packets, sf, osr;
    float bw_scale, amplitude; // amplitude already clamped to [-1, 1] (:18)
    unsigned sw0, sw1;         // sync chirp symbols (:20-22)
};

struct AwgnArgs {
    float2* iq;
    size_t count, packet_stride;
    const float* snr_db;
    unsigned long long seed;
    int n_packets;
};

cudaError_t launch_modulate(const ModulateArgs& a, int sm_count, cudaStream_t st);
cudaError_t launch_awgn(const AwgnArgs& a, int sm_count, cudaStream_t st);

} // namespace lora_b200
