// std_kernels.h -- the LoRa coding chain as it is deployed on the air: header block at
// reduced rate, diagonal interleaver, Hamming 4/5..4/8, whitening LFSR x^8+x^6+x^5+x^4+1, CRC-16/CCITT with the last two
// payload bytes folded in.  The reference's own chain (code_kernels.cu) is the SX127x-style variant of the Pothos blocks;
// this one decodes the standards-style frames of the reference's golden capture, which the reference itself cannot.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "lora_phy_b200.h"

namespace lora_b200 {

struct StdDecodeArgs {
    lora_b200_std_cfg cfg;
    const uint16_t* symbols;
    size_t n_symbols, symbol_stride;
    uint8_t* payload;              // LORA_B200_PAYLOAD_STRIDE bytes per packet
    lora_b200_std_result* results;
    int n_packets;
    // optional: packets as the stream synchroniser reported them (see ChainDecodeArgs in code_kernels.h)
    const void* stream_packets;
    const void* stream_results;
    unsigned pkt_cap;
};

struct StdEncodeArgs {
    lora_b200_std_cfg cfg;
    const uint8_t* payload;
    size_t payload_len, payload_stride;
    uint16_t* symbols;
    size_t n_symbols, symbol_stride;
    int n_packets;
};

// Symbols the encoder emits for `len` payload bytes; 0 if cfg is invalid.
size_t std_symbol_count(const lora_b200_std_cfg& cfg, size_t len);

cudaError_t launch_std_decode(const StdDecodeArgs& a, cudaStream_t st);
cudaError_t launch_std_encode(const StdEncodeArgs& a, cudaStream_t st);

} // namespace lora_b200
