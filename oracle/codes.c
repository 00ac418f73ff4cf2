/* codes.c -- integer code primitives of the LoRa PHY, restated in plain C.
 * TEST INFRASTRUCTURE ONLY (see lora_oracle.h).  Follows include/lora_phy/LoRaCodes.hpp. */
#include "lora_oracle.h"

static unsigned bit(unsigned v, unsigned i) { return (v >> i) & 1u; }

unsigned orc_round_up(unsigned num, unsigned factor) /* LoRaCodes.hpp:24-27 */
{
    return ((num + factor - 1) / factor) * factor;
}

uint8_t orc_checksum8(const uint8_t* p, size_t len) /* LoRaCodes.hpp:32-41 */
{
    uint8_t acc = 0;
    for (size_t i = 0; i < len; ++i) {
        acc = (uint8_t)((acc >> 1) + ((acc & 1u) << 7));
        acc = (uint8_t)(acc + p[i]);
    }
    return acc;
}

uint8_t orc_header_checksum(const uint8_t* h) /* LoRaCodes.hpp:43-67 */
{
    /* 12 header bits: a = high nibble of h[0], b = low nibble of h[0], c = low nibble of h[1] */
    const unsigned a0 = bit(h[0], 4), a1 = bit(h[0], 5), a2 = bit(h[0], 6), a3 = bit(h[0], 7);
    const unsigned b0 = bit(h[0], 0), b1 = bit(h[0], 1), b2 = bit(h[0], 2), b3 = bit(h[0], 3);
    const unsigned c0 = bit(h[1], 0), c1 = bit(h[1], 1), c2 = bit(h[1], 2), c3 = bit(h[1], 3);
    unsigned r = (a0 ^ a1 ^ a2 ^ a3) << 4;
    r |= (a3 ^ b1 ^ b2 ^ b3 ^ c0) << 3;
    r |= (a2 ^ b0 ^ b3 ^ c1 ^ c3) << 2;
    r |= (a1 ^ b0 ^ b2 ^ c0 ^ c1 ^ c2) << 1;
    r |= a0 ^ b1 ^ c0 ^ c1 ^ c2 ^ c3;
    return (uint8_t)r;
}

uint16_t orc_crc16sx(uint16_t crc, uint16_t poly) /* LoRaCodes.hpp:69-79 */
{
    for (int i = 0; i < 8; ++i)
        crc = (crc & 0x8000u) ? (uint16_t)((crc << 1) ^ poly) : (uint16_t)(crc << 1);
    return crc;
}

uint8_t orc_xsum8(uint8_t t) /* LoRaCodes.hpp:81-86 */
{
    t ^= t >> 4;
    t ^= t >> 2;
    t ^= t >> 1;
    return t & 1u;
}

uint16_t orc_data_checksum(const uint8_t* data, int length) /* LoRaCodes.hpp:92-105 */
{
    uint16_t res = 0, crc = 0;
    uint8_t v = 0xff;
    for (int i = 0; i < length; ++i) {
        crc = orc_crc16sx(res, 0x1021);
        v = (uint8_t)(orc_xsum8(v & 0xB8) | (v << 1));
        res = (uint16_t)(crc ^ data[i]);
    }
    res ^= v;
    v = (uint8_t)(orc_xsum8(v & 0xB8) | (v << 1));
    res ^= (uint16_t)(v << 8);
    (void)crc;
    return res;
}

void orc_whiten_sx1232(uint8_t* buf, uint16_t n) /* LoRaCodes.hpp:111-137 */
{
    uint8_t msb = 0x01, lsb = 0xFF;
    for (uint16_t j = 0; j < n; ++j) {
        buf[j] ^= lsb;
        for (int i = 0; i < 8; ++i) {
            const uint8_t prev = msb;
            msb = (uint8_t)((lsb & 1u) ^ ((lsb >> 5) & 1u));
            lsb = (uint8_t)(((lsb >> 1) & 0xFF) | ((prev << 7) & 0x80));
        }
    }
}

void orc_whiten_table(uint8_t* buf, uint16_t n, int bit_ofs, int rdd) /* LoRaCodes.hpp:147-167 */
{
    static const int ofs_std[8] = {6, 4, 2, 0, -112, -114, -302, -34};
    static const int ofs_p54[5] = {6, 4, 2, 0, -360};
    static const uint64_t seq[8] = {0x0102291EA751AAFFull, 0xD24B050A8D643A17ull,
                                    0x5B279B671120B8F4ull, 0x032B37B9F6FB55A2ull,
                                    0x994E0F87E95E2D16ull, 0x7CBCFC7631984C26ull,
                                    0x281C8E4F0DAEF7F9ull, 0x1741886EB7733B15ull};
    const int period = 510;
    const int* ofs = (rdd == 1) ? ofs_p54 : ofs_std;
    for (int j = 0; j < n; ++j) {
        uint8_t x = 0;
        for (int i = 0; i < 4 + rdd; ++i) {
            const int t = (ofs[i] + j + bit_ofs + period) % period;
            if (seq[t >> 6] & ((uint64_t)1 << (t & 0x3F))) x |= (uint8_t)(1u << i);
        }
        buf[j] ^= x;
    }
}

static uint64_t lfsr_step(uint64_t r) /* poly 0x1D, eight interleaved bit lanes */
{
    return (r >> 8) | (((r >> 32) ^ (r >> 24) ^ (r >> 16) ^ r) << 56);
}

void orc_whiten_lfsr(uint8_t* buf, uint16_t n, int bit_ofs, size_t rdd) /* LoRaCodes.hpp:176-189 */
{
    uint64_t r[2];
    if (rdd == 1) { r[0] = 0x05121100F8ECFEEFull; r[1] = 0xF8ECFEEFEFEFEFEFull; }
    else          { r[0] = 0x6572D100E85C2EFFull; r[1] = 0xE85C2EFFFFFFFFFFull; }
    const uint8_t mask = (uint8_t)(0xff >> (4 - rdd));
    int i = 0;
    for (; i < bit_ofs; ++i) r[i & 1] = lfsr_step(r[i & 1]);
    for (int j = 0; j < n; ++j, ++i) {
        buf[j] ^= (uint8_t)(r[i & 1] & mask);
        r[i & 1] = lfsr_step(r[i & 1]);
    }
}

uint16_t orc_bin_to_gray(uint16_t v) { return (uint16_t)(v ^ (v >> 1)); } /* :201-207 */

uint16_t orc_gray_to_bin(uint16_t v) /* LoRaCodes.hpp:212-222 */
{
    v = (uint16_t)(v ^ (v >> 8));
    v = (uint16_t)(v ^ (v >> 4));
    v = (uint16_t)(v ^ (v >> 2));
    v = (uint16_t)(v ^ (v >> 1));
    return v;
}

uint8_t orc_enc_h84(uint8_t x) /* LoRaCodes.hpp:229-243 */
{
    const unsigned d0 = bit(x, 0), d1 = bit(x, 1), d2 = bit(x, 2), d3 = bit(x, 3);
    unsigned b = x & 0xfu;
    b |= (d0 ^ d1 ^ d2) << 4;
    b |= (d1 ^ d2 ^ d3) << 5;
    b |= (d0 ^ d1 ^ d3) << 6;
    b |= (d0 ^ d2 ^ d3) << 7;
    return (uint8_t)b;
}

uint8_t orc_dec_h84(uint8_t b, int* error, int* bad) /* LoRaCodes.hpp:251-281 */
{
    const unsigned p0 = bit(b, 0) ^ bit(b, 1) ^ bit(b, 2) ^ bit(b, 4);
    const unsigned p1 = bit(b, 1) ^ bit(b, 2) ^ bit(b, 3) ^ bit(b, 5);
    const unsigned p2 = bit(b, 0) ^ bit(b, 1) ^ bit(b, 3) ^ bit(b, 6);
    const unsigned p3 = bit(b, 0) ^ bit(b, 2) ^ bit(b, 3) ^ bit(b, 7);
    const unsigned syn = p0 | (p1 << 1) | (p2 << 2) | (p3 << 3);
    if (syn != 0) *error = 1;
    switch (syn) {
    case 0xD: return (uint8_t)((b ^ 1) & 0xf);
    case 0x7: return (uint8_t)((b ^ 2) & 0xf);
    case 0xB: return (uint8_t)((b ^ 4) & 0xf);
    case 0xE: return (uint8_t)((b ^ 8) & 0xf);
    case 0x0: case 0x1: case 0x2: case 0x4: case 0x8: return b & 0xf;
    default: *bad = 1; return b & 0xf;
    }
}

uint8_t orc_enc_h74(uint8_t x) /* LoRaCodes.hpp:287-299 */
{
    const unsigned d0 = bit(x, 0), d1 = bit(x, 1), d2 = bit(x, 2), d3 = bit(x, 3);
    unsigned b = x & 0xfu;
    b |= (d0 ^ d1 ^ d2) << 4;
    b |= (d1 ^ d2 ^ d3) << 5;
    b |= (d0 ^ d1 ^ d3) << 6;
    return (uint8_t)b;
}

uint8_t orc_dec_h74(uint8_t b, int* error) /* LoRaCodes.hpp:306-334 */
{
    const unsigned p0 = bit(b, 0) ^ bit(b, 1) ^ bit(b, 2) ^ bit(b, 4);
    const unsigned p1 = bit(b, 1) ^ bit(b, 2) ^ bit(b, 3) ^ bit(b, 5);
    const unsigned p2 = bit(b, 0) ^ bit(b, 1) ^ bit(b, 3) ^ bit(b, 6);
    const unsigned syn = p0 | (p1 << 1) | (p2 << 2);
    if (syn != 0) *error = 1;
    switch (syn) {
    case 0x5: return (uint8_t)((b ^ 1) & 0xf);
    case 0x7: return (uint8_t)((b ^ 2) & 0xf);
    case 0x3: return (uint8_t)((b ^ 4) & 0xf);
    case 0x6: return (uint8_t)((b ^ 8) & 0xf);
    default: return b & 0xf;
    }
}

uint8_t orc_chk_p54(uint8_t b, int* error) /* LoRaCodes.hpp:340-345 */
{
    int x = b ^ (b >> 2);
    x = x ^ (x >> 1) ^ (b >> 4);
    if (x & 1) *error = 1;
    return b & 0xf;
}

uint8_t orc_enc_p54(uint8_t b) /* LoRaCodes.hpp:347-351 */
{
    int x = b ^ (b >> 2);
    x = x ^ (x >> 1);
    return (uint8_t)((b & 0xf) | ((x << 4) & 0x10));
}

uint8_t orc_chk_p64(uint8_t b, int* error) /* LoRaCodes.hpp:357-365 */
{
    int x = b ^ (b >> 1) ^ (b >> 2);
    int y = x ^ b ^ (b >> 3);
    x ^= b >> 4;
    y ^= b >> 5;
    if ((x | y) & 1) *error = 1;
    return b & 0xf;
}

uint8_t orc_enc_p64(uint8_t b) /* LoRaCodes.hpp:367-371 */
{
    const int x = b ^ (b >> 1) ^ (b >> 2);
    const int y = x ^ b ^ (b >> 3);
    return (uint8_t)(((x & 1) << 4) | ((y & 1) << 5) | (b & 0xf));
}

void orc_interleave(const uint8_t* cw, size_t ncw, uint16_t* sym, size_t ppm, size_t rdd)
{ /* LoRaCodes.hpp:376-393 */
    const size_t rows = 4 + rdd;
    for (size_t blk = 0; blk < ncw / ppm; ++blk) {
        const uint8_t* c = cw + blk * ppm;
        uint16_t* s = sym + blk * rows;
        for (size_t row = 0; row < rows; ++row) {
            uint16_t v = 0;
            for (size_t col = 0; col < ppm; ++col)
                v |= (uint16_t)(((c[(col + row) % ppm] >> row) & 1u) << col);
            s[row] = v;
        }
    }
}

void orc_deinterleave(const uint16_t* sym, size_t nsym, uint8_t* cw, size_t ppm, size_t rdd)
{ /* LoRaCodes.hpp:396-412; cw[] must be zero on entry (it is OR-ed into) */
    const size_t rows = 4 + rdd;
    for (size_t blk = 0; blk < nsym / rows; ++blk) {
        uint8_t* c = cw + blk * ppm;
        const uint16_t* s = sym + blk * rows;
        for (size_t row = 0; row < rows; ++row) {
            uint16_t v = s[row];
            for (size_t col = 0; col < ppm; ++col, v >>= 1)
                c[(col + row) % ppm] |= (uint8_t)((v & 1u) << row);
        }
    }
}
