/* std_chain.c -- CPU checker for the LoRa coding chain as it is deployed on the air.
 *
 * TEST INFRASTRUCTURE ONLY (see lora_oracle.h).
 *
 * Parity status: UNPINNED against reference code -- the reference has no such chain.  Its only standards-style
 * input is the golden capture vectors_binary/bw_125k_sf_7_cr_1_ldro_false_crc_true_implheader_false.unknown, which its
 * own receive path cannot decode (rx_runner cuts it into aligned windows and prints 1043 bytes of
 * noise).  What pins this file instead: all 16 frames of that capture decode to "hello world: 0" ... "hello world: 15"
 * with a valid explicit header (length 14/15, CR 4/5, CRC present, header checksum 3/5) and a valid payload CRC
 * (tests/test_oracle_std.py) -- 16 independent 16-bit checks plus 16 five-bit ones on data this repository did not
 * produce.  That covers SF7, CR 4/8 (header block) and CR 4/5, explicit header, no LDRO.  The other coding rates,
 * LDRO and implicit header follow the same published construction (Robyns et al., "A Multi-Channel Software
 * Decoder for the LoRa Modulation Scheme", 2018; Tapparel et al., "An Open-Source LoRa Physical Layer Prototype on
 * GNU Radio", 2020) and are covered by encoder/decoder round trips and error-correction properties only.
 *
 * The chain, receive direction:
 *   bin index s            -> v = (s - 1) mod 2^sf           (the transmitter adds 1)
 *   reduced rate           -> v >>= 2                        (first 8 symbols always; every block with LDRO)
 *   Gray                   -> g = v ^ (v >> 1)
 *   diagonal deinterleave  -> block of (4+cr) symbols x sf_app bits -> sf_app codewords of (4+cr) bits:
 *                             bit i (MSB first) of codeword (i - j - 1) mod sf_app = bit j (MSB first) of symbol i
 *   Hamming                -> codeword MSB first = b0 b1 b2 b3 p0 p1 p2 p3 (b0 = nibble LSB),
 *                             p0 = b0^b1^b2, p1 = b1^b2^b3, p2 = b0^b1^b3, p3 = b0^b2^b3; 4/5 carries b0^b1^b2^b3 only;
 *                             4/7 and 4/8 correct one bit, 4/5 and 4/6 detect
 *   header (explicit)      -> nibbles len_hi, len_lo, (cr << 1 | has_crc), c4, c3..c0 (five-bit checksum below)
 *   payload                -> bytes from nibble pairs, low nibble first, XOR whitening LFSR x^8+x^6+x^5+x^4+1 from 0xFF
 *   CRC                    -> two un-whitened bytes, low first: CRC-16/CCITT (0x1021, init 0) over the first len-2
 *                             payload bytes, XOR payload[len-1] XOR payload[len-2] << 8
 */
#include <string.h>

#include "lora_oracle.h"

static unsigned std_bit(unsigned v, unsigned i) { return (v >> i) & 1u; }

static unsigned std_hamming_encode(unsigned nib, unsigned cr)
{
    const unsigned b0 = std_bit(nib, 0), b1 = std_bit(nib, 1), b2 = std_bit(nib, 2), b3 = std_bit(nib, 3);
    if (cr == 1) return b0 << 4 | b1 << 3 | b2 << 2 | b3 << 1 | (b0 ^ b1 ^ b2 ^ b3);
    const unsigned p0 = b0 ^ b1 ^ b2, p1 = b1 ^ b2 ^ b3, p2 = b0 ^ b1 ^ b3, p3 = b0 ^ b2 ^ b3;
    return (b0 << 7 | b1 << 6 | b2 << 5 | b3 << 4 | p0 << 3 | p1 << 2 | p2 << 1 | p3) >> (4 - cr);
}

/* -> nibble; *fixed set when one bit was corrected, *bad when an error was detected that cannot be corrected */
static unsigned std_hamming_decode(unsigned cw, unsigned cr, int* fixed, int* bad)
{
    const unsigned len = 4 + cr;
    unsigned c[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (unsigned i = 0; i < len; ++i) c[i] = std_bit(cw, len - 1 - i);
    if (cr == 1) {
        if (c[0] ^ c[1] ^ c[2] ^ c[3] ^ c[4]) *bad = 1;
    } else {
        const unsigned s0 = c[0] ^ c[1] ^ c[2] ^ c[4], s1 = c[1] ^ c[2] ^ c[3] ^ c[5];
        if (cr == 2) {
            if (s0 | s1) *bad = 1;
        } else {
            const unsigned s2 = c[0] ^ c[1] ^ c[3] ^ c[6];
            const unsigned s3 = cr == 4 ? (c[0] ^ c[2] ^ c[3] ^ c[7]) : 0u;
            const unsigned syn = s0 | s1 << 1 | s2 << 2 | s3 << 3;
            if (syn) {
                int which = -1; /* data bit to flip, -2 = a parity bit */
                if (cr == 3) {
                    switch (syn) { case 5: which = 0; break; case 7: which = 1; break; case 3: which = 2; break;
                                   case 6: which = 3; break; default: which = -2; break; }
                } else {
                    switch (syn) { case 13: which = 0; break; case 7: which = 1; break; case 11: which = 2; break;
                                   case 14: which = 3; break; case 1: case 2: case 4: case 8: which = -2; break;
                                   default: which = -1; break; }
                }
                if (which >= 0) { c[which] ^= 1u; *fixed = 1; }
                else if (which == -2) *fixed = 1;
                else *bad = 1;
            }
        }
    }
    return c[0] | c[1] << 1 | c[2] << 2 | c[3] << 3;
}

uint8_t orc_std_header_checksum(unsigned n0, unsigned n1, unsigned n2)
{
    const unsigned c4 = std_bit(n0, 3) ^ std_bit(n0, 2) ^ std_bit(n0, 1) ^ std_bit(n0, 0);
    const unsigned c3 = std_bit(n0, 3) ^ std_bit(n1, 3) ^ std_bit(n1, 2) ^ std_bit(n1, 1) ^ std_bit(n2, 0);
    const unsigned c2 = std_bit(n0, 2) ^ std_bit(n1, 3) ^ std_bit(n1, 0) ^ std_bit(n2, 3) ^ std_bit(n2, 1);
    const unsigned c1 = std_bit(n0, 1) ^ std_bit(n1, 2) ^ std_bit(n1, 0) ^ std_bit(n2, 2) ^ std_bit(n2, 1) ^ std_bit(n2, 0);
    const unsigned c0 = std_bit(n0, 0) ^ std_bit(n1, 1) ^ std_bit(n2, 3) ^ std_bit(n2, 2) ^ std_bit(n2, 1) ^ std_bit(n2, 0);
    return (uint8_t)(c4 << 4 | c3 << 3 | c2 << 2 | c1 << 1 | c0);
}

void orc_std_whitening(uint8_t* seq, size_t n)
{
    unsigned v = 0xFF;
    for (size_t i = 0; i < n; ++i) {
        seq[i] = (uint8_t)v;
        const unsigned fb = (v >> 7 ^ v >> 5 ^ v >> 4 ^ v >> 3) & 1u;
        v = (v << 1 | fb) & 0xFFu;
    }
}

uint16_t orc_std_crc(const uint8_t* p, size_t len)
{
    unsigned crc = 0;
    for (size_t i = 0; i + 2 < len; ++i)
        for (int b = 7; b >= 0; --b) {
            const unsigned top = (crc >> 15 & 1u) ^ (p[i] >> b & 1u);
            crc = (crc << 1) & 0xFFFFu;
            if (top) crc ^= 0x1021u;
        }
    if (len >= 1) crc ^= p[len - 1];
    if (len >= 2) crc ^= (unsigned)p[len - 2] << 8;
    return (uint16_t)crc;
}

static int std_cfg_ok(const orc_std_cfg* c)
{
    return c && c->sf >= 7 && c->sf <= 12 && c->cr >= 1 && c->cr <= 4 && c->payload_len <= 255;
}

static size_t std_nibbles(const orc_std_cfg* c, size_t len)
{
    return (c->implicit_header ? 0 : 5) + 2 * len + (c->has_crc ? 4 : 0);
}

size_t orc_std_symbol_count(const orc_std_cfg* c, size_t len)
{
    if (!std_cfg_ok(c) || len > 255) return 0;
    const size_t first = c->sf - 2, app = c->ldro ? c->sf - 2 : c->sf, nn = std_nibbles(c, len);
    const size_t rest = nn > first ? nn - first : 0;
    return 8 + (rest + app - 1) / app * (4 + c->cr);
}

/* one interleaver block: n_cw codewords of cw_len bits -> cw_len symbols */
static void std_interleave(const unsigned* cw, unsigned n_cw, unsigned cw_len, unsigned sf, int reduced, uint16_t* out)
{
    for (unsigned i = 0; i < cw_len; ++i) {
        unsigned v = 0;
        for (unsigned j = 0; j < n_cw; ++j) {
            const unsigned k = (i + n_cw * 2 - j - 1) % n_cw;
            v = v << 1 | std_bit(cw[k], cw_len - 1 - i);
        }
        unsigned b = v; /* Gray to binary */
        for (unsigned s = 1; s < sf; s <<= 1) b ^= b >> s;
        if (reduced) b <<= 2; /* reduced-rate symbols sit on every fourth bin (the capture's header symbols are 4k + 1) */
        out[i] = (uint16_t)((b + 1) & ((1u << sf) - 1));
    }
}

static void std_deinterleave(const uint16_t* sym, unsigned n_cw, unsigned cw_len, unsigned sf, int reduced, unsigned* cw)
{
    for (unsigned k = 0; k < n_cw; ++k) cw[k] = 0;
    for (unsigned i = 0; i < cw_len; ++i) {
        unsigned v = ((unsigned)sym[i] + (1u << sf) - 1) & ((1u << sf) - 1);
        if (reduced) v >>= 2;
        v ^= v >> 1;
        for (unsigned j = 0; j < n_cw; ++j) {
            const unsigned k = (i + n_cw * 2 - j - 1) % n_cw;
            cw[k] |= std_bit(v, n_cw - 1 - j) << (cw_len - 1 - i);
        }
    }
}

/* Symbols in the frame whose explicit header sits in the eight reduced-rate symbols `sym`; 0 when the header checksum
 * fails, the coding rate is not 1..4 or the length is zero. */
size_t orc_std_frame_symbols(unsigned sf, int ldro, const uint16_t* sym)
{
    unsigned cw[12];
    int fixed = 0, bad = 0;
    unsigned nib[5];
    if (sf < 7 || sf > 12) return 0;
    std_deinterleave(sym, sf - 2, 8, sf, 1, cw);
    for (unsigned k = 0; k < 5; ++k) nib[k] = std_hamming_decode(cw[k], 4, &fixed, &bad);
    const unsigned len = nib[0] << 4 | nib[1], cr = nib[2] >> 1;
    const unsigned chk = (nib[3] & 1u) << 4 | nib[4];
    if (chk != orc_std_header_checksum(nib[0], nib[1], nib[2]) || cr < 1 || cr > 4 || len == 0) return 0;
    orc_std_cfg c = {sf, cr, ldro, 0, (int)(nib[2] & 1u), 0};
    return orc_std_symbol_count(&c, len);
}

size_t orc_std_encode(const orc_std_cfg* c, const uint8_t* payload, size_t len, uint16_t* out, size_t cap)
{
    const size_t nsym = orc_std_symbol_count(c, len);
    if (nsym == 0 || nsym > cap || (!payload && len)) return 0;
    uint8_t nib[5 + 2 * 255 + 4 + 16];
    uint8_t white[255];
    size_t nn = 0;
    memset(nib, 0, sizeof(nib));
    orc_std_whitening(white, len);
    if (!c->implicit_header) {
        const unsigned n0 = (unsigned)len >> 4, n1 = (unsigned)len & 15u, n2 = c->cr << 1 | (c->has_crc ? 1u : 0u);
        const unsigned chk = orc_std_header_checksum(n0, n1, n2);
        nib[nn++] = (uint8_t)n0; nib[nn++] = (uint8_t)n1; nib[nn++] = (uint8_t)n2;
        nib[nn++] = (uint8_t)(chk >> 4); nib[nn++] = (uint8_t)(chk & 15u);
    }
    for (size_t i = 0; i < len; ++i) {
        const unsigned b = payload[i] ^ white[i];
        nib[nn++] = (uint8_t)(b & 15u); nib[nn++] = (uint8_t)(b >> 4);
    }
    if (c->has_crc) {
        const unsigned crc = orc_std_crc(payload, len);
        nib[nn++] = (uint8_t)(crc & 15u); nib[nn++] = (uint8_t)(crc >> 4 & 15u);
        nib[nn++] = (uint8_t)(crc >> 8 & 15u); nib[nn++] = (uint8_t)(crc >> 12 & 15u);
    }
    unsigned cw[12];
    size_t at = 0, o = 0;
    const unsigned first = c->sf - 2, app = c->ldro ? c->sf - 2 : c->sf;
    for (unsigned k = 0; k < first; ++k) cw[k] = std_hamming_encode(at < nn ? nib[at] : 0u, 4), ++at;
    std_interleave(cw, first, 8, c->sf, 1, out);
    o = 8;
    while (o < nsym) {
        for (unsigned k = 0; k < app; ++k) cw[k] = std_hamming_encode(at < nn ? nib[at] : 0u, c->cr), ++at;
        std_interleave(cw, app, 4 + c->cr, c->sf, c->ldro, out + o);
        o += 4 + c->cr;
    }
    return nsym;
}

long orc_std_decode(const orc_std_cfg* c, const uint16_t* symbols, size_t nsym, uint8_t* out, size_t cap, orc_std_info* info)
{
    orc_std_info r;
    memset(&r, 0, sizeof(r));
    r.n_bytes = -2;
    if (!std_cfg_ok(c) || !symbols || nsym < 8) { if (info) *info = r; return r.n_bytes; }
    uint8_t nib[16 + 2 * 255 + 4 + 16];
    size_t nn = 0;
    unsigned cw[12];
    int fixed = 0, bad = 0;
    const unsigned first = c->sf - 2, app = c->ldro ? c->sf - 2 : c->sf;
    std_deinterleave(symbols, first, 8, c->sf, 1, cw);
    for (unsigned k = 0; k < first; ++k) nib[nn++] = (uint8_t)std_hamming_decode(cw[k], 4, &fixed, &bad);
    unsigned len = c->payload_len, cr = c->cr;
    int has_crc = c->has_crc;
    size_t at = 0;
    if (!c->implicit_header) {
        len = (unsigned)nib[0] << 4 | nib[1];
        cr = nib[2] >> 1;
        has_crc = nib[2] & 1;
        const unsigned chk = (unsigned)(nib[3] & 1u) << 4 | nib[4];
        r.hdr_length = (int)len; r.hdr_cr = (int)cr; r.has_crc = has_crc;
        r.hdr_ok = chk == orc_std_header_checksum(nib[0], nib[1], nib[2]) && cr >= 1 && cr <= 4 && len > 0;
        if (!r.hdr_ok) { r.n_bytes = -1; r.corrected = fixed; r.uncorrectable = bad; if (info) *info = r; return -1; }
        at = 5;
    } else {
        r.hdr_ok = 1; r.hdr_length = (int)len; r.hdr_cr = (int)cr; r.has_crc = has_crc;
    }
    const size_t need = at + 2 * (size_t)len + (has_crc ? 4 : 0);
    size_t o = 8;
    while (nn < need) {
        if (o + 4 + cr > nsym) { r.n_bytes = -3; r.corrected = fixed; r.uncorrectable = bad; if (info) *info = r; return -3; }
        std_deinterleave(symbols + o, app, 4 + cr, c->sf, c->ldro, cw);
        for (unsigned k = 0; k < app; ++k) nib[nn++] = (uint8_t)std_hamming_decode(cw[k], cr, &fixed, &bad);
        o += 4 + cr;
    }
    uint8_t white[255], bytes[255];
    orc_std_whitening(white, len);
    for (unsigned i = 0; i < len; ++i) bytes[i] = (uint8_t)((nib[at + 2 * i] | nib[at + 2 * i + 1] << 4) ^ white[i]);
    if (has_crc) {
        const size_t q = at + 2 * (size_t)len;
        const unsigned rx = nib[q] | nib[q + 1] << 4 | nib[q + 2] << 8 | (unsigned)nib[q + 3] << 12;
        r.crc_ok = rx == orc_std_crc(bytes, len);
    }
    r.corrected = fixed; r.uncorrectable = bad;
    r.n_bytes = (int)len;
    r.symbols_used = (int)o;
    if (out) memcpy(out, bytes, len < cap ? len : cap);
    if (info) *info = r;
    return (long)len;
}
