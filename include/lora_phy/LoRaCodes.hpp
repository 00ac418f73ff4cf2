// lora_phy/LoRaCodes.hpp -- caller-side code helpers of the `lora_phy` API, B200 build.
//
// Reference callers (runners/lora_phy_vector_dump.cpp:105,153-160, tests/whitening_test.cpp, tests/hamming_cross_check.cpp)
// use these header-only helpers directly on a handful of host bytes; they are not the receive path (that is
// csrc/code_kernels.cu on the device) but part of the source-level contract (reference
// include/lora_phy/LoRaCodes.hpp:16-432): same names, signatures, results and in/out flag behaviour.  The bodies are
// this build's own: every parity is the parity of a masked word, whitening comes from one keystream generator, the
// CRC is table-free bit arithmetic.  tests/test_cpp_api.py checks every function against the reference's over its whole
// input range.
#pragma once

#include <cstddef>
#include <cstdint>

#define HEADER_RDD 4                           // the header block is always sent at 4/8 (reference :16)
#define N_HEADER_SYMBOLS (HEADER_RDD + 4)      // :17
#define N_HEADER_CODEWORDS 5                   // :18

namespace lora_codes_detail {

// parity (XOR of all bits) of v
static inline unsigned par(unsigned v)
{
    v ^= v >> 16;
    v ^= v >> 8;
    v ^= v >> 4;
    return (0x6996u >> (v & 0xfu)) & 1u;
}

// The SX127x (8,4) code: parity bit p_k (codeword bit 4+k) covers the data bits in DATA_MASK[k].
static const unsigned DATA_MASK[4] = {0x7u, 0xEu, 0xBu, 0xDu};

static inline unsigned parity_bits(unsigned nibble, unsigned count)
{
    unsigned p = 0;
    for (unsigned k = 0; k < count; ++k) p |= par(nibble & DATA_MASK[k]) << k;
    return p;
}

// syndrome bit k = parity over the data bits of p_k and p_k itself
static inline unsigned syndrome(unsigned word, unsigned count)
{
    return parity_bits(word & 0xfu, count) ^ ((word >> 4) & ((1u << count) - 1u));
}

// One step of the 8-bit keystream register behind the SX127x data whitening (feedback taps 0, 2, 3, 4 of the byte that
// leaves): 8 such registers run side by side in a 64-bit word, stepped one byte at a time.
static inline uint64_t lane_step(uint64_t r)
{
    return (r >> 8) | ((r ^ (r >> 16) ^ (r >> 24) ^ (r >> 32)) << 56);
}

// Keystream byte j (before masking to the codeword width) of the two interleaved registers; `single_parity` selects the
// seed pair the modem uses at coding rate 4/5.
struct Keystream {
    uint64_t r[2];
    unsigned n;
    explicit Keystream(bool single_parity) : n(0)
    {
        r[0] = single_parity ? 0x05121100F8ECFEEFull : 0x6572D100E85C2EFFull;
        r[1] = single_parity ? 0xF8ECFEEFEFEFEFEFull : 0xE85C2EFFFFFFFFFFull;
    }
    uint8_t next()
    {
        uint64_t& x = r[n++ & 1u];
        const uint8_t out = static_cast<uint8_t>(x);
        x = lane_step(x);
        return out;
    }
};

} // namespace lora_codes_detail

// ---- sizes and checksums ------------------------------------------------------------------------------------------------
static inline unsigned roundUp(unsigned num, unsigned factor) // :24
{
    const unsigned rem = num % factor;
    return rem ? num + (factor - rem) : num;
}

static inline uint8_t checksum8(const uint8_t* p, const size_t len) // :32  rotate right by one, then add
{
    unsigned acc = 0;
    for (size_t i = 0; i < len; ++i) acc = (((acc >> 1) | (acc << 7)) + p[i]) & 0xffu;
    return static_cast<uint8_t>(acc);
}

// 5 check bits over the 12 header bits w = h[0] | (h[1] & 0xf) << 8 (:43-67): each is the parity of a fixed subset.
static inline uint8_t headerChecksum(const uint8_t* h)
{
    using lora_codes_detail::par;
    const unsigned w = h[0] | ((h[1] & 0xfu) << 8);
    return static_cast<uint8_t>((par(w & 0x0F0u) << 4) | (par(w & 0x18Eu) << 3) | (par(w & 0xA49u) << 2) |
                                (par(w & 0x725u) << 1) | par(w & 0xF12u));
}

static inline uint16_t crc16sx(uint16_t crc, const uint16_t poly) // :69  eight MSB-first shifts
{
    unsigned c = crc;
    for (int i = 0; i < 8; ++i) c = ((c << 1) ^ ((c & 0x8000u) ? poly : 0u)) & 0xffffu;
    return static_cast<uint16_t>(c);
}

static inline uint8_t xsum8(uint8_t t) { return static_cast<uint8_t>(lora_codes_detail::par(t)); } // :81

// CCITT update that lags one byte behind the data, masked at the end by two bytes of an 8-bit LFSR (:92-105).
static inline uint16_t sx1272DataChecksum(const uint8_t* data, int length)
{
    unsigned res = 0, v = 0xff;
    for (int i = 0; i < length; ++i) {
        const unsigned shifted = crc16sx(static_cast<uint16_t>(res), 0x1021);
        v = ((v << 1) | xsum8(static_cast<uint8_t>(v & 0xB8u))) & 0xffu;
        res = shifted ^ data[i];
    }
    res ^= v;
    v = ((v << 1) | xsum8(static_cast<uint8_t>(v & 0xB8u))) & 0xffu;
    res ^= v << 8;
    return static_cast<uint16_t>(res);
}

// ---- whitening ------------------------------------------------------------------------------------------------------------
// SX1232-style packet whitening: 9-bit LFSR x^9 + x^5 + 1 from 0x1FF, its low byte XORed onto each data byte (:111-137).
static inline void SX1232RadioComputeWhitening(uint8_t* buffer, uint16_t bufferSize)
{
    unsigned state = 0x1FF;
    for (uint16_t j = 0; j < bufferSize; ++j) {
        buffer[j] ^= static_cast<uint8_t>(state);
        for (int k = 0; k < 8; ++k) state = (state >> 1) | (((state ^ (state >> 5)) & 1u) << 8);
    }
}

// SX1272 codeword whitening from the interleaved LFSRs (:176-189): skip bitOfs keystream bytes, then XOR one byte,
// masked to the 4+RDD codeword bits, onto each codeword.
static inline void Sx1272ComputeWhiteningLfsr(uint8_t* buffer, uint16_t bufferSize, const int bitOfs, const size_t RDD)
{
    lora_codes_detail::Keystream ks(RDD == 1);
    const uint8_t mask = static_cast<uint8_t>(0xffu >> (4 - RDD));
    for (int i = 0; i < bitOfs; ++i) ks.next();
    for (uint16_t j = 0; j < bufferSize; ++j) buffer[j] ^= ks.next() & mask;
}

// The table form the transmitter uses (:147-167).  Its 510-entry sequence is the keystream above, period 510: position
// (j + bitOfs) mod 510 for any offset the table form accepts (bitOfs >= -150).
static inline void Sx1272ComputeWhitening(uint8_t* buffer, uint16_t bufferSize, const int bitOfs, const int RDD)
{
    const int period = 510;
    int skip = (bitOfs % period + period) % period;
    lora_codes_detail::Keystream ks(RDD == 1);
    const uint8_t mask = static_cast<uint8_t>(0xffu >> (4 - RDD));
    while (skip-- > 0) ks.next();
    for (uint16_t j = 0; j < bufferSize; ++j) buffer[j] ^= ks.next() & mask;
}

// ---- Gray ------------------------------------------------------------------------------------------------------------------
static inline unsigned short binaryToGray16(unsigned short num) // :201
{
    return static_cast<unsigned short>(num ^ (num >> 1));
}

static inline unsigned short grayToBinary16(unsigned short num) // :212  prefix XOR from the top bit down
{
    unsigned v = num;
    for (unsigned s = 1; s < 16; s <<= 1) v ^= v >> s;
    return static_cast<unsigned short>(v);
}

// ---- Hamming(8,4), Hamming(7,4), parity (5,4) and (6,4), SX127x bit layout --------------------------------------------------
static inline unsigned char encodeHamming84sx(const unsigned char x) // :229
{
    return static_cast<unsigned char>((x & 0xfu) | (lora_codes_detail::parity_bits(x & 0xfu, 4) << 4));
}

// Single-error correction, double-error detection (:250-281): a syndrome with one bit set is a damaged parity bit (data
// untouched); 0xD, 0x7, 0xB, 0xE point at data bits 0..3; anything else sets `bad`.  `error` is set on any non-zero
// syndrome; neither flag is ever cleared.
static inline unsigned char decodeHamming84sx(const unsigned char b, bool& error, bool& bad)
{
    const unsigned s = lora_codes_detail::syndrome(b, 4);
    if (s) error = true;
    unsigned flip = 0;
    for (unsigned bit = 0; bit < 4; ++bit)
        if (s == (lora_codes_detail::parity_bits(1u << bit, 4))) flip = 1u << bit;
    if (!flip && (s & (s - 1u))) bad = true; // more than one syndrome bit and no data-bit signature
    return static_cast<unsigned char>((b ^ flip) & 0xfu);
}

static inline unsigned char encodeHamming74sx(const unsigned char x) // :287
{
    return static_cast<unsigned char>((x & 0xfu) | (lora_codes_detail::parity_bits(x & 0xfu, 3) << 4));
}

static inline unsigned char decodeHamming74sx(const unsigned char b, bool& error) // :306
{
    const unsigned s = lora_codes_detail::syndrome(b & 0x7fu, 3);
    if (s) error = true;
    unsigned flip = 0;
    for (unsigned bit = 0; bit < 4; ++bit)
        if (s == (lora_codes_detail::parity_bits(1u << bit, 3))) flip = 1u << bit;
    return static_cast<unsigned char>((b ^ flip) & 0xfu);
}

static inline unsigned char checkParity54(const unsigned char b, bool& error) // :340
{
    if (lora_codes_detail::par(b & 0x1fu)) error = true;
    return b & 0xf;
}

static inline unsigned char encodeParity54(const unsigned char b) // :347
{
    return static_cast<unsigned char>((b & 0xfu) | (lora_codes_detail::par(b & 0xfu) << 4));
}

static inline unsigned char checkParity64(const unsigned char b, bool& error) // :357  p0 over d0 d1 d2, p1 over d1 d2 d3
{
    using lora_codes_detail::par;
    if (par(b & 0x17u) | par(b & 0x2Eu)) error = true;
    return b & 0xf;
}

static inline unsigned char encodeParity64(const unsigned char b) // :367
{
    using lora_codes_detail::par;
    return static_cast<unsigned char>((b & 0xfu) | (par(b & 0x7u) << 4) | (par(b & 0xEu) << 5));
}

// ---- diagonal interleaver ---------------------------------------------------------------------------------------------------
// Block of PPM codewords (4+RDD bits each) <-> block of 4+RDD symbols (PPM bits each): symbol `row` bit `col` carries
// bit `row` of codeword (col + row) mod PPM (:376-412).  The interleaver assigns its output, the deinterleaver ORs into
// codewords the caller has zeroed.
static inline void diagonalInterleaveSx(const uint8_t* codewords, const size_t numCodewords, uint16_t* symbols,
                                        const size_t PPM, const size_t RDD)
{
    const size_t rows = 4 + RDD;
    for (size_t b = 0; b < numCodewords / PPM; ++b) {
        const uint8_t* cw = codewords + b * PPM;
        uint16_t* sym = symbols + b * rows;
        for (size_t row = 0; row < rows; ++row) {
            unsigned v = 0;
            for (size_t col = 0; col < PPM; ++col) v |= ((cw[(col + row) % PPM] >> row) & 1u) << col;
            sym[row] = static_cast<uint16_t>(v);
        }
    }
}

static inline void diagonalDeterleaveSx(const uint16_t* symbols, const size_t numSymbols, uint8_t* codewords,
                                        const size_t PPM, const size_t RDD)
{
    const size_t rows = 4 + RDD;
    for (size_t b = 0; b < numSymbols / rows; ++b) {
        const uint16_t* sym = symbols + b * rows;
        uint8_t* cw = codewords + b * PPM;
        for (size_t row = 0; row < rows; ++row)
            for (size_t col = 0; col < PPM; ++col)
                cw[(col + row) % PPM] |= static_cast<uint8_t>(((sym[row] >> col) & 1u) << row);
    }
}

// The reference's second deinterleaver (:415-432) walks PPM rows per block instead of 4+RDD, so it equals the one above
// only when PPM == 4+RDD; no caller uses it.  Kept with the reference's behaviour, including
// the rows it reads beyond the block when PPM > 4+RDD.
static inline void diagonalDeterleaveSx2(const uint16_t* symbols, const size_t numSymbols, uint8_t* codewords,
                                         const size_t PPM, const size_t RDD)
{
    const size_t rows = 4 + RDD;
    for (size_t b = 0; b < numSymbols / rows; ++b)
        for (size_t row = 0; row < PPM; ++row) {
            const unsigned v = symbols[b * rows + row];
            for (size_t col = 0; col < PPM; ++col)
                codewords[b * PPM + (row + col) % PPM] |= static_cast<uint8_t>(((v >> col) & 1u) << row);
        }
}
