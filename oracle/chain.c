/* chain.c -- the full LoRa coding chain: the Pothos blocks LoRa-SDR/LoRaEncoder.cpp and
 * LoRa-SDR/LoRaDecoder.cpp restated without Pothos, in plain C.
 * TEST INFRASTRUCTURE ONLY (see lora_oracle.h).
 *
 * Those two files cannot be compiled here (Pothos is absent), so this follows their work() bodies
 * line by line and calls the primitives of codes.c, each of which is checked one-to-one against the
 * reference's own LoRaCodes.hpp through oracle/_ref.  Parity of the orchestration itself is pinned by
 * the encoder->decoder identity over SF 7..12 x CR 4/4..4/8 (LoRa-SDR/TestLoopback.cpp:14-62). */
#include "lora_oracle.h"

#include <stdlib.h>
#include <string.h>

#define HEADER_RDD 4          /* LoRaCodes.hpp:16 */
#define N_HEADER_SYMBOLS 8    /* LoRaCodes.hpp:17 */
#define N_HEADER_CODEWORDS 5  /* LoRaCodes.hpp:18 */

/* LoRaEncoder.cpp:132-159.  The reference indexes bytes[dOfs >> 1] without a bound: when the
 * codeword count is rounded up to a multiple of PPM the last few reads run past the end of the
 * std::vector (undefined; the decoder never looks at those padding nibbles).  Here a read past the
 * end yields 0, and the CUDA encoder does the same. */
static void encode_fec(uint8_t* cw, unsigned rdd, size_t* c_ofs, size_t* d_ofs, const uint8_t* bytes,
                       size_t nbytes, size_t count)
{
    for (size_t i = 0; i < count; ++i, ++*d_ofs) {
        const size_t bi = *d_ofs >> 1;
        const uint8_t byte = bi < nbytes ? bytes[bi] : 0;
        const uint8_t nib = (*d_ofs & 1) ? (uint8_t)(byte >> 4) : (uint8_t)(byte & 0xf);
        uint8_t w;
        switch (rdd) {
        case 0: w = nib; break;
        case 1: w = orc_enc_p54(nib); break;
        case 2: w = orc_enc_p64(nib); break;
        case 3: w = orc_enc_h74(nib); break;
        default: w = orc_enc_h84(nib); break;
        }
        cw[(*c_ofs)++] = w;
    }
}

size_t orc_chain_encode(const orc_chain_cfg* c, const uint8_t* payload, size_t len, uint16_t* symbols,
                        size_t cap) /* LoRaEncoder.cpp:161-233 */
{
    const size_t ppm = c->ppm ? c->ppm : c->sf;
    if (ppm > c->sf) return 0; /* :166 throws */
    const size_t nbytes = len + (c->crc ? 2 : 0);
    const size_t ncw = orc_round_up((unsigned)(nbytes * 2 + (c->explicit_header ? N_HEADER_CODEWORDS : 0)),
                                    (unsigned)ppm);
    const size_t nsym = N_HEADER_SYMBOLS + (ncw / ppm - 1) * (4 + c->rdd); /* :175 */
    if (!symbols) return nsym;
    if (nsym > cap) return 0;

    uint8_t* bytes = (uint8_t*)calloc(nbytes + 1, 1);
    uint8_t* cw = (uint8_t*)calloc(ncw, 1);
    memcpy(bytes, payload, len);
    if (c->crc) { /* :181-185 */
        const uint16_t crc = orc_data_checksum(bytes, (int)len);
        bytes[len] = (uint8_t)(crc & 0xff);
        bytes[len + 1] = (uint8_t)((crc >> 8) & 0xff);
    }
    size_t c_ofs = 0, d_ofs = 0;
    if (c->explicit_header) { /* :187-199 */
        uint8_t hdr[3];
        hdr[0] = (uint8_t)len;
        hdr[1] = (uint8_t)((c->crc ? 1 : 0) | (c->rdd << 1));
        hdr[2] = orc_header_checksum(hdr);
        cw[c_ofs++] = orc_enc_h84((uint8_t)(hdr[0] >> 4));
        cw[c_ofs++] = orc_enc_h84((uint8_t)(hdr[0] & 0xf));
        cw[c_ofs++] = orc_enc_h84((uint8_t)(hdr[1] & 0xf));
        cw[c_ofs++] = orc_enc_h84((uint8_t)(hdr[2] >> 4));
        cw[c_ofs++] = orc_enc_h84((uint8_t)(hdr[2] & 0xf));
    }
    const size_t c_ofs1 = c_ofs; /* :200-204: the first block is always coded 4/8 */
    encode_fec(cw, 4, &c_ofs, &d_ofs, bytes, nbytes, ppm - c_ofs);
    if (c->whitening) orc_whiten_table(cw + c_ofs1, (uint16_t)(ppm - c_ofs1), 0, HEADER_RDD);
    if (ncw > ppm) { /* :206-212 */
        const size_t c_ofs2 = c_ofs;
        encode_fec(cw, c->rdd, &c_ofs, &d_ofs, bytes, nbytes, ncw - ppm);
        if (c->whitening)
            orc_whiten_table(cw + c_ofs2, (uint16_t)(ncw - ppm), (int)(ppm - c_ofs1), (int)c->rdd);
    }
    orc_interleave(cw, ppm, symbols, ppm, HEADER_RDD); /* :215-219 */
    if (ncw > ppm) orc_interleave(cw + ppm, ncw - ppm, symbols + N_HEADER_SYMBOLS, ppm, c->rdd);
    for (size_t i = 0; i < nsym; ++i) { /* :222-225 */
        uint16_t s = orc_gray_to_bin(symbols[i]);
        symbols[i] = (uint16_t)(s << (c->sf - ppm));
    }
    free(bytes);
    free(cw);
    return nsym;
}

/* Bounds-guarded codeword fetch: the reference reads codewords[cOfs++] unguarded (a std::vector);
 * within the limits it checks (:311) the index stays inside, the guard only makes the oracle safe. */
static uint8_t cw_at(const uint8_t* cw, size_t ncw, size_t i) { return i < ncw ? cw[i] : 0; }

static uint8_t fec_nibble(unsigned rdd, uint8_t w, int* error, int* bad)
{
    switch (rdd) {
    case 0: return w; /* raw: the reference does not mask here (:324, :347-348) */
    case 1: return orc_chk_p54(w, error);
    case 2: return orc_chk_p64(w, error);
    case 3: return orc_dec_h74(w, error);
    default: return orc_dec_h84(w, error, bad);
    }
}

#define DROP(line) do { if (info) { info->drop_line = (line); info->parity_error = error; \
                                    info->uncorrectable = bad; } goto dropped; } while (0)

long orc_chain_decode(const orc_chain_cfg* c, const uint16_t* in_symbols, size_t nin, uint8_t* out,
                      size_t cap, orc_chain_info* info) /* LoRaDecoder.cpp:196-397 */
{
    long result = -1;
    if (info) memset(info, 0, sizeof(*info));
    const size_t ppm = c->ppm ? c->ppm : c->sf;
    if (ppm > c->sf) return -2;            /* :202 throws */
    if (nin < N_HEADER_SYMBOLS) return -2; /* :208 */

    const size_t cfg_rdd = c->rdd;
    const size_t nsym = orc_round_up((unsigned)nin, (unsigned)(4 + cfg_rdd));   /* :210 */
    const size_t ncw = (nsym / (4 + cfg_rdd)) * ppm;                            /* :211 */
    uint16_t* sym = (uint16_t*)calloc(nsym, sizeof(uint16_t));
    uint8_t* cw = (uint8_t*)calloc(ncw ? ncw : 1, 1);
    const size_t nbytes = (ncw + 1) / 2;                                        /* :274 */
    uint8_t* bytes = (uint8_t*)calloc(nbytes + 8, 1);
    memcpy(sym, in_symbols, nin * sizeof(uint16_t));
    int error = 0, bad = 0;
    unsigned rdd = (unsigned)cfg_rdd; /* :215, replaced by the header's value in explicit mode */

    for (size_t i = 0; i < nsym; ++i) { /* :218-222 */
        uint16_t s = sym[i];
        s = (uint16_t)(s + (1 << (c->sf - ppm)) / 2);
        s = (uint16_t)(s >> (c->sf - ppm));
        sym[i] = orc_bin_to_gray(s);
    }

    if (!c->interleaving) { /* :264-270 */
        const size_t n = nsym * 2 <= cap ? nsym * 2 : cap;
        memcpy(out, sym, n);
        result = (long)(nsym * 2);
        goto done;
    }
    if (rdd != HEADER_RDD) { /* :227-246 */
        orc_deinterleave(sym, N_HEADER_SYMBOLS, cw, ppm, HEADER_RDD);
        if (c->explicit_header)
            orc_whiten_lfsr(cw + N_HEADER_CODEWORDS, (uint16_t)(ppm - N_HEADER_CODEWORDS), 0, HEADER_RDD);
        else
            orc_whiten_lfsr(cw, (uint16_t)ppm, 0, HEADER_RDD);
        const size_t c_ofs = ppm, s_ofs = N_HEADER_SYMBOLS;
        if (nsym - s_ofs > 0) {
            orc_deinterleave(sym + s_ofs, nsym - s_ofs, cw + c_ofs, ppm, rdd);
            orc_whiten_lfsr(cw + c_ofs, (uint16_t)(ncw - c_ofs),
                            c->explicit_header ? (int)(ppm - N_HEADER_CODEWORDS) : (int)ppm, rdd);
        }
    } else { /* :247-255 */
        orc_deinterleave(sym, nsym, cw, ppm, rdd);
        if (c->explicit_header)
            orc_whiten_lfsr(cw + N_HEADER_CODEWORDS, (uint16_t)(ncw - N_HEADER_CODEWORDS), 0, rdd);
        else
            orc_whiten_lfsr(cw, (uint16_t)ncw, 0, rdd);
    }

    size_t d_ofs = 0, c_ofs = 0, packet_len = 0, data_len = 0;
    int check_crc = c->crc;
    if (c->explicit_header) { /* :282-303 */
        bytes[0] = orc_dec_h84(cw_at(cw, ncw, 1), &error, &bad) & 0xf;
        bytes[0] |= (uint8_t)(orc_dec_h84(cw_at(cw, ncw, 0), &error, &bad) << 4);
        bytes[1] = orc_dec_h84(cw_at(cw, ncw, 2), &error, &bad) & 0xf;
        bytes[2] = orc_dec_h84(cw_at(cw, ncw, 4), &error, &bad) & 0xf;
        bytes[2] |= (uint8_t)(orc_dec_h84(cw_at(cw, ncw, 3), &error, &bad) << 4);
        bytes[2] ^= orc_header_checksum(bytes);
        if (info) {
            info->hdr_length = bytes[0];
            info->hdr_has_crc = bytes[1] & 1;
            info->hdr_rdd = (bytes[1] >> 1) & 7;
            info->hdr_checksum = bytes[2];
        }
        if (error && c->error_check) DROP(293);
        if ((bytes[1] & 1) == 0) check_crc = 0;
        rdd = (bytes[1] >> 1) & 0x7;
        if (rdd > 4) DROP(297);
        packet_len = bytes[0];
        data_len = packet_len + ((bytes[1] & 1) ? 5 : 3);
        c_ofs = N_HEADER_CODEWORDS;
        d_ofs = 6;
    } else { /* :304-311 */
        packet_len = c->data_length;
        data_len = c->crc ? packet_len + 2 : packet_len;
    }
    if (data_len > nbytes) DROP(313);

    for (; c_ofs < ppm; ++c_ofs, ++d_ofs) { /* :315-320: rest of the first block is always 4/8 */
        const uint8_t nib = orc_dec_h84(cw_at(cw, ncw, c_ofs), &error, &bad);
        if (d_ofs & 1) bytes[d_ofs >> 1] |= (uint8_t)(nib << 4);
        else bytes[d_ofs >> 1] = nib & 0xf;
    }
    if (d_ofs & 1) { /* :322-340 */
        bytes[d_ofs >> 1] |= (uint8_t)(fec_nibble(rdd, cw_at(cw, ncw, c_ofs++), &error, &bad) << 4);
        d_ofs++;
    }
    d_ofs >>= 1;
    if (error && c->error_check) DROP(342);

    for (size_t i = d_ofs; i < data_len; ++i) { /* :346-361 */
        uint8_t lo = fec_nibble(rdd, cw_at(cw, ncw, c_ofs++), &error, &bad);
        if (rdd == 0 || rdd == 3 || rdd == 4) lo &= 0xf; /* :347, :356, :359 mask the low nibble */
        const uint8_t hi = fec_nibble(rdd, cw_at(cw, ncw, c_ofs++), &error, &bad);
        bytes[i] = (uint8_t)(lo | (uint8_t)(hi << 4));
    }
    if (error && c->error_check) DROP(363);

    d_ofs = 0;
    if (c->explicit_header) { /* :367-384 */
        if (bytes[1] & 1) {
            const uint16_t crc = orc_data_checksum(bytes + 3, (int)packet_len);
            const uint16_t pkt = (uint16_t)(bytes[3 + packet_len] | (bytes[4 + packet_len] << 8));
            if (info) { info->crc_present = 1; info->crc_ok = crc == pkt; }
            if (crc != pkt && check_crc) DROP(371);
            bytes[3 + packet_len] ^= (uint8_t)crc;
            bytes[4 + packet_len] ^= (uint8_t)(crc >> 8);
        }
        if (!c->hdr_out) {
            d_ofs = 3;
            if (data_len < 5) { result = -3; goto done; } /* size_t wrap in the reference */
            data_len -= 5;
        }
    } else if (check_crc) { /* :386-393 */
        const uint16_t crc = orc_data_checksum(bytes, (int)c->data_length);
        const uint16_t pkt = (uint16_t)(bytes[c->data_length] | (bytes[c->data_length + 1] << 8));
        if (info) { info->crc_present = 1; info->crc_ok = crc == pkt; }
        if (crc != pkt) DROP(390);
        bytes[c->data_length] ^= (uint8_t)crc;
        bytes[c->data_length + 1] ^= (uint8_t)(crc >> 8);
    }
    if (info) { info->parity_error = error; info->uncorrectable = bad; }
    memcpy(out, bytes + d_ofs, data_len <= cap ? data_len : cap);
    result = (long)data_len;
    goto done;
dropped:
    result = -1;
done:
    free(sym);
    free(cw);
    free(bytes);
    return result;
}
