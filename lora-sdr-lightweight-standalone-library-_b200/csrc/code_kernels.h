// code_kernels.h -- integer back end: Gray, diagonal (de)interleave, Hamming / parity, whitening, CRC.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "lora_phy_b200.h"

namespace lora_b200 {

constexpr int WHITEN_LEN = 1024; // bytes of each whitening sequence kept on the device (max index used ~ 620)

struct SimpleDecodeArgs { // lora_decode + the CRC flag of decode(), phy.cpp:241-256
    const uint16_t* symbols;
    size_t n_symbols, symbol_stride;
    uint8_t* bytes;
    size_t byte_stride;
    uint8_t* crc_ok; // optional
    int n_packets;
};

struct SimpleEncodeArgs { // lora_encode, src/phy/LoRaEncoder.cpp:6-18
    const uint8_t* bytes;
    size_t n_bytes, byte_stride;
    uint16_t* symbols;
    size_t symbol_stride;
    int n_packets;
};

struct ChainDecodeArgs { // LoRaDecoder::work, LoRa-SDR/LoRaDecoder.cpp:196-397
    lora_b200_chain_cfg cfg;
    const uint16_t* symbols;
    size_t n_symbols, symbol_stride;
    uint8_t* payload; // LORA_B200_PAYLOAD_STRIDE bytes per packet
    lora_b200_chain_result* results;
    const uint8_t* whiten; // 2 x WHITEN_LEN
    int n_packets;
    // optional: packets as the stream synchroniser reported them (stream_kernels.h).  Packet slot p = stream * pkt_cap
    // + k takes its symbols from symbols[stream * symbol_stride + first_symbol ...] and its count from the record;
    // a slot beyond the stream's packet count yields n_bytes = -1.
    const void* stream_packets; // StreamPacket[n_streams * pkt_cap]
    const void* stream_results; // StreamResult[n_streams]
    unsigned pkt_cap;
};

struct ChainEncodeArgs { // LoRaEncoder::work, LoRa-SDR/LoRaEncoder.cpp:161-233
    lora_b200_chain_cfg cfg;
    const uint8_t* payload;
    size_t payload_len, payload_stride;
    uint16_t* symbols;
    size_t symbol_stride;
    const uint8_t* whiten;
    int n_packets;
};

// Symbols the chain encoder emits for a payload of `len` bytes (LoRaEncoder.cpp:174-175); 0 if cfg is invalid.
size_t chain_symbol_count(const lora_b200_chain_cfg& cfg, size_t len);

cudaError_t launch_simple_decode(const SimpleDecodeArgs& a, cudaStream_t st);
cudaError_t launch_simple_encode(const SimpleEncodeArgs& a, cudaStream_t st);
cudaError_t launch_chain_decode(const ChainDecodeArgs& a, cudaStream_t st);
cudaError_t launch_chain_encode(const ChainEncodeArgs& a, cudaStream_t st);

} // namespace lora_b200
