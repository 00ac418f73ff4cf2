// std_chain.cuh -- device primitives of the LoRa coding chain as deployed on the air, shared by the chain kernels
// (std_kernels.cu) and by the frame synchroniser when it reads the frame length from the header (stream_kernels.cu).
// Construction and what pins it: oracle/std_chain.c.
#pragma once

#include <stdint.h>

namespace lora_b200 {
namespace stdchain {

__device__ __forceinline__ unsigned sbit(unsigned v, unsigned i) { return (v >> i) & 1u; }

// nibble -> codeword, MSB first b0 b1 b2 b3 p0 p1 p2 p3 (b0 = nibble LSB); 4/5 carries the overall parity only
__device__ inline unsigned std_hamming_encode(unsigned nib, unsigned cr)
{
    const unsigned b0 = sbit(nib, 0), b1 = sbit(nib, 1), b2 = sbit(nib, 2), b3 = sbit(nib, 3);
    if (cr == 1) return b0 << 4 | b1 << 3 | b2 << 2 | b3 << 1 | (b0 ^ b1 ^ b2 ^ b3);
    const unsigned p0 = b0 ^ b1 ^ b2, p1 = b1 ^ b2 ^ b3, p2 = b0 ^ b1 ^ b3, p3 = b0 ^ b2 ^ b3;
    return (b0 << 7 | b1 << 6 | b2 << 5 | b3 << 4 | p0 << 3 | p1 << 2 | p2 << 1 | p3) >> (4 - cr);
}

// codeword -> nibble.  4/7 and 4/8 correct one bit (syndrome -> position), 4/5 and 4/6 only detect.
__device__ inline unsigned std_hamming_decode(unsigned cw, unsigned cr, bool& fixed, bool& bad)
{
    const unsigned len = 4 + cr;
    unsigned c0 = sbit(cw, len - 1), c1 = sbit(cw, len - 2), c2 = sbit(cw, len - 3), c3 = sbit(cw, len - 4);
    if (cr == 1) {
        if (c0 ^ c1 ^ c2 ^ c3 ^ sbit(cw, 0)) bad = true;
    } else {
        const unsigned s0 = c0 ^ c1 ^ c2 ^ sbit(cw, len - 5), s1 = c1 ^ c2 ^ c3 ^ sbit(cw, len - 6);
        if (cr == 2) {
            if (s0 | s1) bad = true;
        } else {
            const unsigned s2 = c0 ^ c1 ^ c3 ^ sbit(cw, len - 7);
            const unsigned s3 = cr == 4 ? (c0 ^ c2 ^ c3 ^ sbit(cw, 0)) : 0u;
            const unsigned syn = s0 | s1 << 1 | s2 << 2 | s3 << 3;
            if (syn) {
                // data-bit syndromes: c0 -> s0,s2(,s3); c1 -> s0,s1,s2; c2 -> s0,s1(,s3); c3 -> s1,s2(,s3)
                const unsigned e0 = cr == 3 ? 5u : 13u, e1 = 7u, e2 = cr == 3 ? 3u : 11u, e3 = cr == 3 ? 6u : 14u;
                if (syn == e0) { c0 ^= 1u; fixed = true; }
                else if (syn == e1) { c1 ^= 1u; fixed = true; }
                else if (syn == e2) { c2 ^= 1u; fixed = true; }
                else if (syn == e3) { c3 ^= 1u; fixed = true; }
                else if (cr == 3 || (syn & (syn - 1)) == 0) fixed = true; // a parity bit
                else bad = true;                                           // two errors (4/8 only)
            }
        }
    }
    return c0 | c1 << 1 | c2 << 2 | c3 << 3;
}

__device__ inline unsigned std_header_checksum(unsigned n0, unsigned n1, unsigned n2)
{
    const unsigned c4 = sbit(n0, 3) ^ sbit(n0, 2) ^ sbit(n0, 1) ^ sbit(n0, 0);
    const unsigned c3 = sbit(n0, 3) ^ sbit(n1, 3) ^ sbit(n1, 2) ^ sbit(n1, 1) ^ sbit(n2, 0);
    const unsigned c2 = sbit(n0, 2) ^ sbit(n1, 3) ^ sbit(n1, 0) ^ sbit(n2, 3) ^ sbit(n2, 1);
    const unsigned c1 = sbit(n0, 1) ^ sbit(n1, 2) ^ sbit(n1, 0) ^ sbit(n2, 2) ^ sbit(n2, 1) ^ sbit(n2, 0);
    const unsigned c0 = sbit(n0, 0) ^ sbit(n1, 1) ^ sbit(n2, 3) ^ sbit(n2, 2) ^ sbit(n2, 1) ^ sbit(n2, 0);
    return c4 << 4 | c3 << 3 | c2 << 2 | c1 << 1 | c0;
}

__device__ __forceinline__ unsigned whiten_next(unsigned v)
{
    return (v << 1 | ((v >> 7 ^ v >> 5 ^ v >> 4 ^ v >> 3) & 1u)) & 0xFFu;
}

__device__ __forceinline__ unsigned crc_byte(unsigned crc, unsigned byte)
{
    for (int b = 7; b >= 0; --b) {
        const unsigned top = (crc >> 15 & 1u) ^ (byte >> b & 1u);
        crc = (crc << 1) & 0xFFFFu;
        if (top) crc ^= 0x1021u;
    }
    return crc;
}

__device__ inline unsigned std_crc(const uint8_t* p, unsigned len)
{
    unsigned crc = 0;
    for (unsigned i = 0; i + 2 < len; ++i) crc = crc_byte(crc, p[i]);
    if (len >= 1) crc ^= p[len - 1];
    if (len >= 2) crc ^= (unsigned)p[len - 2] << 8;
    return crc;
}

// block of cw_len symbols -> n_cw codewords: bit i of codeword (i - j - 1) mod n_cw = bit j of symbol i (MSB first)
__device__ inline void std_deinterleave(const uint16_t* sym, unsigned n_cw, unsigned cw_len, unsigned sf, bool reduced, unsigned* cw)
{
    for (unsigned k = 0; k < n_cw; ++k) cw[k] = 0;
    const unsigned mask = (1u << sf) - 1u;
    for (unsigned i = 0; i < cw_len; ++i) {
        unsigned v = ((unsigned)sym[i] + mask) & mask; // the transmitter adds 1
        if (reduced) v >>= 2;
        v ^= v >> 1;
        for (unsigned j = 0; j < n_cw; ++j) {
            const unsigned k = (i + 2 * n_cw - j - 1) % n_cw;
            cw[k] |= sbit(v, n_cw - 1 - j) << (cw_len - 1 - i);
        }
    }
}

__device__ inline void std_interleave(const unsigned* cw, unsigned n_cw, unsigned cw_len, unsigned sf, bool reduced, uint16_t* out)
{
    for (unsigned i = 0; i < cw_len; ++i) {
        unsigned v = 0;
        for (unsigned j = 0; j < n_cw; ++j) v = v << 1 | sbit(cw[(i + 2 * n_cw - j - 1) % n_cw], cw_len - 1 - i);
        for (unsigned s = 1; s < sf; s <<= 1) v ^= v >> s; // Gray to binary
        if (reduced) v <<= 2;                              // reduced-rate symbols sit on every fourth bin
        out[i] = (uint16_t)((v + 1u) & ((1u << sf) - 1u));
    }
}

// Symbols in the frame whose explicit header sits in the eight reduced-rate symbols `sym` (bin indices; only their value
// modulo 2^sf matters); 0 when the header checksum fails, the coding rate is not 1..4 or the length is zero.
__device__ inline unsigned std_frame_symbols(unsigned sf, bool ldro, const uint16_t* sym)
{
    unsigned cw[12];
    std_deinterleave(sym, sf - 2, 8, sf, true, cw);
    bool fixed = false, bad = false;
    unsigned nib[5];
    for (unsigned k = 0; k < 5; ++k) nib[k] = std_hamming_decode(cw[k], 4, fixed, bad);
    const unsigned len = nib[0] << 4 | nib[1], cr = nib[2] >> 1, has_crc = nib[2] & 1u;
    const unsigned chk = (nib[3] & 1u) << 4 | nib[4];
    if (chk != std_header_checksum(nib[0], nib[1], nib[2]) || cr < 1 || cr > 4 || len == 0) return 0;
    const unsigned first = sf - 2, app = ldro ? sf - 2 : sf, nn = 5 + 2 * len + (has_crc ? 4u : 0u);
    const unsigned rest = nn > first ? nn - first : 0u;
    return 8 + (rest + app - 1) / app * (4 + cr);
}

} // namespace stdchain
} // namespace lora_b200
