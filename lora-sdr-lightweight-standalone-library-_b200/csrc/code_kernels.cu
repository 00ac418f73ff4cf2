// code_kernels.cu -- the bit-exact integer back end.  One thread per packet: the work per packet is a few
// hundred integer ops on <= ~600 symbols, three orders of magnitude below the demod kernel, so the layout
// is chosen for clarity of parity with LoRa-SDR/LoRaDecoder.cpp rather than for throughput.
#include "code_kernels.h"
#include "stream_kernels.h" // StreamPacket / StreamResult records
#include "demod_kernels.h" // g_launches

namespace lora_b200 {

// ---- primitives (include/lora_phy/LoRaCodes.hpp) ---------------------------------------------------------
__device__ __forceinline__ unsigned bitof(unsigned v, unsigned i) { return (v >> i) & 1u; }

__device__ __forceinline__ unsigned enc_h84(unsigned x) // LoRaCodes.hpp:229-243
{
    const unsigned d0 = bitof(x, 0), d1 = bitof(x, 1), d2 = bitof(x, 2), d3 = bitof(x, 3);
    return (x & 0xf) | ((d0 ^ d1 ^ d2) << 4) | ((d1 ^ d2 ^ d3) << 5) | ((d0 ^ d1 ^ d3) << 6) | ((d0 ^ d2 ^ d3) << 7);
}
__device__ __forceinline__ unsigned enc_h74(unsigned x) { return enc_h84(x) & 0x7f; } // :287-299 (same low 7 bits)
__device__ __forceinline__ unsigned enc_p54(unsigned b) // :347-351
{
    unsigned x = b ^ (b >> 2);
    x = x ^ (x >> 1);
    return (b & 0xf) | ((x << 4) & 0x10);
}
__device__ __forceinline__ unsigned enc_p64(unsigned b) // :367-371
{
    const unsigned x = b ^ (b >> 1) ^ (b >> 2);
    const unsigned y = x ^ b ^ (b >> 3);
    return ((x & 1) << 4) | ((y & 1) << 5) | (b & 0xf);
}

__device__ __forceinline__ unsigned dec_h84(unsigned b, bool& error, bool& bad) // :251-281
{
    const unsigned p0 = bitof(b, 0) ^ bitof(b, 1) ^ bitof(b, 2) ^ bitof(b, 4);
    const unsigned p1 = bitof(b, 1) ^ bitof(b, 2) ^ bitof(b, 3) ^ bitof(b, 5);
    const unsigned p2 = bitof(b, 0) ^ bitof(b, 1) ^ bitof(b, 3) ^ bitof(b, 6);
    const unsigned p3 = bitof(b, 0) ^ bitof(b, 2) ^ bitof(b, 3) ^ bitof(b, 7);
    const unsigned syn = p0 | (p1 << 1) | (p2 << 2) | (p3 << 3);
    if (syn) error = true;
    switch (syn) {
    case 0xD: return (b ^ 1) & 0xf;
    case 0x7: return (b ^ 2) & 0xf;
    case 0xB: return (b ^ 4) & 0xf;
    case 0xE: return (b ^ 8) & 0xf;
    case 0x0: case 0x1: case 0x2: case 0x4: case 0x8: return b & 0xf;
    default: bad = true; return b & 0xf;
    }
}
__device__ __forceinline__ unsigned dec_h74(unsigned b, bool& error) // :306-334
{
    const unsigned p0 = bitof(b, 0) ^ bitof(b, 1) ^ bitof(b, 2) ^ bitof(b, 4);
    const unsigned p1 = bitof(b, 1) ^ bitof(b, 2) ^ bitof(b, 3) ^ bitof(b, 5);
    const unsigned p2 = bitof(b, 0) ^ bitof(b, 1) ^ bitof(b, 3) ^ bitof(b, 6);
    const unsigned syn = p0 | (p1 << 1) | (p2 << 2);
    if (syn) error = true;
    switch (syn) {
    case 0x5: return (b ^ 1) & 0xf;
    case 0x7: return (b ^ 2) & 0xf;
    case 0x3: return (b ^ 4) & 0xf;
    case 0x6: return (b ^ 8) & 0xf;
    default: return b & 0xf;
    }
}
__device__ __forceinline__ unsigned chk_p54(unsigned b, bool& error) // :340-345
{
    unsigned x = b ^ (b >> 2);
    x = x ^ (x >> 1) ^ (b >> 4);
    if (x & 1) error = true;
    return b & 0xf;
}
__device__ __forceinline__ unsigned chk_p64(unsigned b, bool& error) // :357-365
{
    unsigned x = b ^ (b >> 1) ^ (b >> 2);
    unsigned y = x ^ b ^ (b >> 3);
    x ^= b >> 4;
    y ^= b >> 5;
    if ((x | y) & 1) error = true;
    return b & 0xf;
}

__device__ __forceinline__ unsigned crc16sx(unsigned crc) // :69-79, poly 0x1021
{
#pragma unroll
    for (int i = 0; i < 8; ++i) crc = (crc & 0x8000u) ? (((crc << 1) ^ 0x1021u) & 0xffffu) : ((crc << 1) & 0xffffu);
    return crc;
}
__device__ __forceinline__ unsigned xsum8(unsigned t) { t ^= t >> 4; t ^= t >> 2; t ^= t >> 1; return t & 1u; } // :81-86

__device__ unsigned data_checksum(const uint8_t* data, int length) // :92-105
{
    unsigned res = 0, v = 0xff;
    for (int i = 0; i < length; ++i) {
        const unsigned crc = crc16sx(res);
        v = (xsum8(v & 0xB8) | (v << 1)) & 0xff;
        res = crc ^ data[i];
    }
    res ^= v;
    v = (xsum8(v & 0xB8) | (v << 1)) & 0xff;
    res ^= v << 8;
    return res & 0xffffu;
}

__device__ __forceinline__ unsigned header_checksum(unsigned h0, unsigned h1) // :43-67
{
    const unsigned a0 = bitof(h0, 4), a1 = bitof(h0, 5), a2 = bitof(h0, 6), a3 = bitof(h0, 7);
    const unsigned b0 = bitof(h0, 0), b1 = bitof(h0, 1), b2 = bitof(h0, 2), b3 = bitof(h0, 3);
    const unsigned c0 = bitof(h1, 0), c1 = bitof(h1, 1), c2 = bitof(h1, 2), c3 = bitof(h1, 3);
    return ((a0 ^ a1 ^ a2 ^ a3) << 4) | ((a3 ^ b1 ^ b2 ^ b3 ^ c0) << 3) | ((a2 ^ b0 ^ b3 ^ c1 ^ c3) << 2) |
           ((a1 ^ b0 ^ b2 ^ c0 ^ c1 ^ c2) << 1) | (a0 ^ b1 ^ c0 ^ c1 ^ c2 ^ c3);
}

__device__ __forceinline__ unsigned gray_to_bin(unsigned v) // :212-222
{
    v ^= v >> 8; v ^= v >> 4; v ^= v >> 2; v ^= v >> 1;
    return v & 0xffffu;
}

// ---- lora_decode / lora_encode ----------------------------------------------------------------------------
__global__ void simple_decode_kernel(SimpleDecodeArgs a)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n_packets) return;
    const uint16_t* s = a.symbols + (size_t)p * a.symbol_stride;
    uint8_t* out = a.bytes + (size_t)p * a.byte_stride;
    size_t k = 0;
    for (size_t i = 0; i + 1 < a.n_symbols; i += 2) { // src/phy/LoRaDecoder.cpp:9-17
        bool e = false, b = false;
        const unsigned hi = dec_h84(s[i] & 0xff, e, b), lo = dec_h84(s[i + 1] & 0xff, e, b);
        out[k++] = (uint8_t)((hi << 4) | lo);
    }
    if (a.crc_ok) { // phy.cpp:247-254
        bool ok = false;
        if (k >= 4) {
            const unsigned provided = out[k - 2] | (out[k - 1] << 8);
            ok = provided == data_checksum(out + 2, (int)(k - 4));
        }
        a.crc_ok[p] = ok;
    }
}

__global__ void simple_encode_kernel(SimpleEncodeArgs a)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n_packets) return;
    const uint8_t* in = a.bytes + (size_t)p * a.byte_stride;
    uint16_t* s = a.symbols + (size_t)p * a.symbol_stride;
    for (size_t i = 0; i < a.n_bytes; ++i) {
        s[2 * i] = (uint16_t)enc_h84(in[i] >> 4);
        s[2 * i + 1] = (uint16_t)enc_h84(in[i] & 0x0f);
    }
}

// ---- LoRaDecoder::work, streamed -------------------------------------------------------------------------
// The reference materialises every Gray'd symbol, every codeword and every byte in std::vectors.  Here the
// decoder walks its codeword cursor exactly as LoRaDecoder.cpp:282-361 does, and a block of PPM codewords is
// deinterleaved and dewhitened on demand when the cursor enters it, so a thread needs 12 bytes of codewords
// live instead of ~600.  The whitening sequence (two interleaved LFSRs, LoRaCodes.hpp:176-189) has period
// 510 for both seed sets, so its k-th byte is a table lookup at k mod 510; the reference's bitOfs
// pre-advance (PPM-5 or PPM for the second region, :241-245) makes the running index simply
// c-5 (explicit header) or c (implicit) for codeword c in every region.
constexpr int WHITEN_PERIOD = 510;

struct CodewordCursor {
    const uint16_t* sym;
    const uint8_t* whiten;
    int nin, nsym, ncw, ppm, cfg_rdd, sf, explicit_hdr;
    int cur_block;
    uint8_t cw[12];

    __device__ unsigned gray_symbol(int idx) const // LoRaDecoder.cpp:218-222 (uint16 arithmetic)
    {
        unsigned v = idx < nin ? sym[idx] : 0u; // symbols past the packet are zero padding (:212-213)
        v = (v + (unsigned)((1 << (sf - ppm)) / 2)) & 0xffffu;
        v >>= (sf - ppm);
        return (v ^ (v >> 1)) & 0xffffu;
    }
    __device__ void load_block(int b)
    {
#pragma unroll
        for (int c = 0; c < 12; ++c) cw[c] = 0;
        int rows, s_ofs;
        bool avail;
        if (cfg_rdd != 4) { // :227-246: header block is always 8 symbols of 4/8, the rest use the configured rate
            if (b == 0) { rows = 8; s_ofs = 0; avail = true; }
            else { rows = 4 + cfg_rdd; s_ofs = 8 + (b - 1) * rows; avail = (b - 1) < (nsym - 8) / rows; }
        } else { rows = 8; s_ofs = 8 * b; avail = b < nsym / 8; } // :247-255
        if (avail) {
            for (int row = 0; row < rows; ++row) { // diagonalDeterleaveSx, LoRaCodes.hpp:396-412
                unsigned v = gray_symbol(s_ofs + row);
                for (int col = 0; col < ppm; ++col, v >>= 1) {
                    const int dst = (col + row) % ppm; // row can exceed ppm (8 header rows, ppm >= 5)
                    cw[dst] |= (uint8_t)((v & 1u) << row);
                }
            }
        }
        const int blk_rdd = (cfg_rdd != 4 && b == 0) ? 4 : cfg_rdd;
        const uint8_t* seq = whiten + (blk_rdd == 1 ? WHITEN_LEN : 0);
        const unsigned mask = 0xffu >> (4 - blk_rdd);
        for (int c = 0; c < ppm; ++c) {
            const int wi = b * ppm + c - (explicit_hdr ? 5 : 0);
            if (wi >= 0) cw[c] ^= (uint8_t)(seq[wi % WHITEN_PERIOD] & mask);
        }
        cur_block = b;
    }
    __device__ unsigned get(int c)
    {
        if (c >= ncw) return 0; // guard only; the reference's own limit (:313) keeps the cursor inside
        const int b = c / ppm;
        if (b != cur_block) load_block(b);
        return cw[c - b * ppm];
    }
};

__device__ __forceinline__ unsigned fec_nibble(unsigned rdd, unsigned w, bool& error, bool& bad)
{
    switch (rdd) {
    case 0: return w;
    case 1: return chk_p54(w, error);
    case 2: return chk_p64(w, error);
    case 3: return dec_h74(w, error);
    default: return dec_h84(w, error, bad);
    }
}

__global__ void chain_decode_kernel(ChainDecodeArgs a)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n_packets) return;
    const lora_b200_chain_cfg& g = a.cfg;
    lora_b200_chain_result r;
    memset(&r, 0, sizeof(r));
    uint8_t* out = a.payload + (size_t)p * LORA_B200_PAYLOAD_STRIDE;
    const int ppm = g.ppm ? (int)g.ppm : (int)g.sf;
    int nin = (int)a.n_symbols;
    const uint16_t* sym_in = a.symbols + (size_t)p * a.symbol_stride;
    if (a.stream_packets) {
        const unsigned st = (unsigned)p / a.pkt_cap, k = (unsigned)p % a.pkt_cap;
        const StreamResult sr = static_cast<const StreamResult*>(a.stream_results)[st];
        if (k >= sr.n_packets) { r.n_bytes = -1; a.results[p] = r; return; } // no such packet in this stream
        const StreamPacket pk = static_cast<const StreamPacket*>(a.stream_packets)[p];
        const unsigned stored = pk.first_symbol < sr.n_symbols ? sr.n_symbols - pk.first_symbol : 0u;
        nin = (int)min(pk.n_symbols, stored); // symbols beyond the caller's capacity were counted but not kept
        sym_in = a.symbols + (size_t)st * a.symbol_stride + pk.first_symbol;
    }
    if (nin < 8) { r.n_bytes = -2; a.results[p] = r; return; } // :208
    const int rows = 4 + (int)g.rdd;
    const int nsym = ((nin + rows - 1) / rows) * rows; // :210
    const int ncw = (nsym / rows) * ppm;               // :211
    const int nbytes = (ncw + 1) / 2;                  // :274

    CodewordCursor cur;
    cur.sym = sym_in;
    cur.whiten = a.whiten;
    cur.nin = nin; cur.nsym = nsym; cur.ncw = ncw; cur.ppm = ppm; cur.cfg_rdd = (int)g.rdd; cur.sf = (int)g.sf;
    cur.explicit_hdr = g.explicit_header;
    cur.cur_block = -1;

    uint8_t bytes[LORA_B200_PAYLOAD_STRIDE + 8]; // header (3) + payload (<= 255) + crc (2), zero like the vector
    for (int i = 0; i < (int)sizeof(bytes); ++i) bytes[i] = 0;
    bool error = false, bad = false;
    unsigned rdd = g.rdd;
    int c_ofs = 0, d_ofs = 0, packet_len = 0, data_len = 0;
    bool check_crc = g.crc != 0;
    int drop = 0;

    if (g.explicit_header) { // :282-303
        bytes[0] = (uint8_t)(dec_h84(cur.get(1), error, bad) & 0xf);
        bytes[0] |= (uint8_t)(dec_h84(cur.get(0), error, bad) << 4);
        bytes[1] = (uint8_t)(dec_h84(cur.get(2), error, bad) & 0xf);
        bytes[2] = (uint8_t)(dec_h84(cur.get(4), error, bad) & 0xf);
        bytes[2] |= (uint8_t)(dec_h84(cur.get(3), error, bad) << 4);
        bytes[2] ^= (uint8_t)header_checksum(bytes[0], bytes[1]);
        r.hdr_length = bytes[0]; r.hdr_has_crc = bytes[1] & 1; r.hdr_rdd = (bytes[1] >> 1) & 7; r.hdr_checksum = bytes[2];
        if (error && g.error_check) drop = 293;
        if (!drop) {
            if ((bytes[1] & 1) == 0) check_crc = false;
            rdd = (bytes[1] >> 1) & 7;
            if (rdd > 4) drop = 297;
        }
        packet_len = bytes[0];
        data_len = packet_len + ((bytes[1] & 1) ? 5 : 3);
        c_ofs = 5;
        d_ofs = 6;
    } else { // :304-311
        packet_len = (int)g.data_length;
        data_len = g.crc ? packet_len + 2 : packet_len;
    }
    if (!drop && (data_len > nbytes || data_len > LORA_B200_PAYLOAD_STRIDE)) drop = 313;

    if (!drop) {
        for (; c_ofs < ppm; ++c_ofs, ++d_ofs) { // :315-320
            const unsigned nib = dec_h84(cur.get(c_ofs), error, bad);
            if (d_ofs & 1) bytes[d_ofs >> 1] |= (uint8_t)(nib << 4);
            else bytes[d_ofs >> 1] = (uint8_t)(nib & 0xf);
        }
        if (d_ofs & 1) { // :322-340
            bytes[d_ofs >> 1] |= (uint8_t)(fec_nibble(rdd, cur.get(c_ofs++), error, bad) << 4);
            d_ofs++;
        }
        d_ofs >>= 1;
        if (error && g.error_check) drop = 342;
    }
    if (!drop) {
        for (int i = d_ofs; i < data_len; ++i) { // :346-361
            unsigned lo = fec_nibble(rdd, cur.get(c_ofs++), error, bad);
            if (rdd == 0 || rdd == 3 || rdd == 4) lo &= 0xf;
            const unsigned hi = fec_nibble(rdd, cur.get(c_ofs++), error, bad);
            bytes[i] = (uint8_t)(lo | (hi << 4));
        }
        if (error && g.error_check) drop = 363;
    }
    int out_ofs = 0;
    if (!drop) {
        if (g.explicit_header) { // :367-384
            if (bytes[1] & 1) {
                const unsigned crc = data_checksum(bytes + 3, packet_len);
                const unsigned pkt = bytes[3 + packet_len] | (bytes[4 + packet_len] << 8);
                r.crc_present = 1; r.crc_ok = crc == pkt;
                if (crc != pkt && check_crc) drop = 371;
                else { bytes[3 + packet_len] ^= (uint8_t)crc; bytes[4 + packet_len] ^= (uint8_t)(crc >> 8); }
            }
            if (!drop && !g.hdr_out) {
                out_ofs = 3;
                if (data_len < 5) { r.n_bytes = -3; a.results[p] = r; return; } // size_t wrap in the reference
                data_len -= 5;
            }
        } else if (check_crc) { // :386-393
            const unsigned crc = data_checksum(bytes, packet_len);
            const unsigned pkt = bytes[packet_len] | (bytes[packet_len + 1] << 8);
            r.crc_present = 1; r.crc_ok = crc == pkt;
            if (crc != pkt) drop = 390;
            else { bytes[packet_len] ^= (uint8_t)crc; bytes[packet_len + 1] ^= (uint8_t)(crc >> 8); }
        }
    }
    r.parity_error = error; r.uncorrectable = bad;
    if (drop) { r.n_bytes = -1; r.drop_line = (int16_t)drop; }
    else {
        r.n_bytes = (int16_t)data_len;
        for (int i = 0; i < data_len; ++i) out[i] = bytes[out_ofs + i];
    }
    a.results[p] = r;
}

// ---- LoRaEncoder::work -----------------------------------------------------------------------------------
size_t chain_symbol_count(const lora_b200_chain_cfg& g, size_t len)
{
    const size_t ppm = g.ppm ? g.ppm : g.sf;
    if (ppm > g.sf || ppm < 5 || ppm > 12 || g.rdd > 4) return 0;
    const size_t nbytes = len + (g.crc ? 2 : 0);
    const size_t raw = nbytes * 2 + (g.explicit_header ? 5 : 0);
    const size_t ncw = ((raw + ppm - 1) / ppm) * ppm;
    if (ncw == 0) return 0;
    return 8 + (ncw / ppm - 1) * (4 + g.rdd); // LoRaEncoder.cpp:174-175
}

__global__ void chain_encode_kernel(ChainEncodeArgs a)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n_packets) return;
    const lora_b200_chain_cfg& g = a.cfg;
    const int ppm = g.ppm ? (int)g.ppm : (int)g.sf;
    const int len = (int)a.payload_len;
    const int nbytes = len + (g.crc ? 2 : 0);
    const int ncw = ((nbytes * 2 + (g.explicit_header ? 5 : 0) + ppm - 1) / ppm) * ppm;
    const uint8_t* pay = a.payload + (size_t)p * a.payload_stride;
    uint16_t* sym = a.symbols + (size_t)p * a.symbol_stride;

    unsigned crc = 0;
    if (g.crc) crc = data_checksum(pay, len); // :181-185
    unsigned hdr[3] = {0, 0, 0};
    if (g.explicit_header) { // :187-192
        hdr[0] = (unsigned)len & 0xff;
        hdr[1] = (g.crc ? 1u : 0u) | (g.rdd << 1);
        hdr[2] = header_checksum(hdr[0], hdr[1]);
    }
    const int c_ofs1 = g.explicit_header ? 5 : 0;
    auto byte_at = [&](int i) -> unsigned { // payload then little-endian CRC; 0 past the end (see oracle/chain.c)
        if (i < len) return pay[i];
        if (g.crc && i == len) return crc & 0xff;
        if (g.crc && i == len + 1) return (crc >> 8) & 0xff;
        return 0u;
    };
    for (int b = 0; b < ncw / ppm; ++b) {
        const unsigned blk_rdd = b == 0 ? 4u : g.rdd; // first block is always 4/8 (:200-204)
        const int rows = 4 + (int)blk_rdd;
        const uint8_t* seq = a.whiten + (blk_rdd == 1 ? WHITEN_LEN : 0);
        const unsigned mask = 0xffu >> (4 - blk_rdd);
        unsigned cw[12];
        for (int c = 0; c < ppm; ++c) {
            const int gc = b * ppm + c;
            unsigned w;
            if (gc < c_ofs1) { // :194-198
                const unsigned nib = gc == 0 ? hdr[0] >> 4 : gc == 1 ? hdr[0] & 0xf : gc == 2 ? hdr[1] & 0xf
                                   : gc == 3 ? hdr[2] >> 4 : hdr[2] & 0xf;
                w = enc_h84(nib);
            } else {
                const int d = gc - c_ofs1; // nibble cursor dOfs (:132-159)
                const unsigned byte = byte_at(d >> 1);
                const unsigned nib = (d & 1) ? (byte >> 4) : (byte & 0xf);
                switch (blk_rdd) {
                case 0: w = nib; break;
                case 1: w = enc_p54(nib); break;
                case 2: w = enc_p64(nib); break;
                case 3: w = enc_h74(nib); break;
                default: w = enc_h84(nib); break;
                }
                if (g.whitening) w ^= seq[d % WHITEN_PERIOD] & mask; // :203,210: running index c - cOfs1
            }
            cw[c] = w;
        }
        uint16_t* s = sym + (b == 0 ? 0 : 8 + (b - 1) * rows);
        for (int row = 0; row < rows; ++row) { // diagonalInterleaveSx, LoRaCodes.hpp:376-393
            unsigned v = 0;
            for (int col = 0; col < ppm; ++col) {
                const int src = (col + row) % ppm; // row can exceed ppm (8 header rows, ppm >= 5)
                v |= ((cw[src] >> row) & 1u) << col;
            }
            s[row] = (uint16_t)((gray_to_bin(v) << (g.sf - ppm)) & 0xffffu); // :222-225
        }
    }
}

// ---- launchers -----------------------------------------------------------------------------------------
static inline unsigned blocks_for(int n) { return (unsigned)((n + 127) / 128); }

cudaError_t launch_simple_decode(const SimpleDecodeArgs& a, cudaStream_t st)
{
    simple_decode_kernel<<<blocks_for(a.n_packets), 128, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}
cudaError_t launch_simple_encode(const SimpleEncodeArgs& a, cudaStream_t st)
{
    simple_encode_kernel<<<blocks_for(a.n_packets), 128, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}
cudaError_t launch_chain_decode(const ChainDecodeArgs& a, cudaStream_t st)
{
    chain_decode_kernel<<<blocks_for(a.n_packets), 128, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}
cudaError_t launch_chain_encode(const ChainEncodeArgs& a, cudaStream_t st)
{
    chain_encode_kernel<<<blocks_for(a.n_packets), 128, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

} // namespace lora_b200
