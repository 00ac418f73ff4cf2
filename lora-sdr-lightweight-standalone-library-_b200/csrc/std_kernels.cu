// std_kernels.cu -- deployed-LoRa coding chain, one thread per packet (integer work, orders of magnitude below the demod
// traffic).  Construction and what pins it: oracle/std_chain.c; tests/test_gpu_std.py compares the two bit for bit.
#include "std_kernels.h"

#include "std_chain.cuh"

#include "demod_kernels.h" // g_launches
#include "stream_kernels.h" // StreamPacket / StreamResult records

namespace lora_b200 {

namespace {

using namespace stdchain;

constexpr int kMaxNib = 16 + 2 * 255 + 4 + 16;

__global__ void std_decode_kernel(StdDecodeArgs a)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n_packets) return;
    const lora_b200_std_cfg& g = a.cfg;
    lora_b200_std_result r;
    memset(&r, 0, sizeof(r));
    uint8_t* out = a.payload + (size_t)p * LORA_B200_PAYLOAD_STRIDE;
    int nin = (int)a.n_symbols;
    const uint16_t* sym = a.symbols + (size_t)p * a.symbol_stride;
    if (a.stream_packets) {
        const unsigned st = (unsigned)p / a.pkt_cap, k = (unsigned)p % a.pkt_cap;
        const StreamResult sr = static_cast<const StreamResult*>(a.stream_results)[st];
        if (k >= sr.n_packets) { r.n_bytes = -1; a.results[p] = r; return; } // no such packet in this stream
        const StreamPacket pk = static_cast<const StreamPacket*>(a.stream_packets)[p];
        const unsigned stored = pk.first_symbol < sr.n_symbols ? sr.n_symbols - pk.first_symbol : 0u;
        nin = (int)min(pk.n_symbols, stored);
        sym = a.symbols + (size_t)st * a.symbol_stride + pk.first_symbol;
    }
    if (nin < 8) { r.n_bytes = -2; a.results[p] = r; return; }
    uint8_t nib[kMaxNib];
    unsigned cw[12];
    int nn = 0;
    bool fixed = false, bad = false;
    const unsigned first = g.sf - 2, app = g.ldro ? g.sf - 2 : g.sf;
    std_deinterleave(sym, first, 8, g.sf, true, cw);
    for (unsigned k = 0; k < first; ++k) nib[nn++] = (uint8_t)std_hamming_decode(cw[k], 4, fixed, bad);
    unsigned len = g.payload_len, cr = g.cr;
    bool has_crc = g.has_crc != 0;
    int at = 0;
    r.hdr_ok = 1;
    if (!g.implicit_header) {
        len = (unsigned)nib[0] << 4 | nib[1];
        cr = nib[2] >> 1;
        has_crc = (nib[2] & 1u) != 0;
        const unsigned chk = (unsigned)(nib[3] & 1u) << 4 | nib[4];
        r.hdr_ok = chk == std_header_checksum(nib[0], nib[1], nib[2]) && cr >= 1 && cr <= 4 && len > 0;
        at = 5;
    }
    r.hdr_length = (uint8_t)len; r.hdr_cr = (uint8_t)cr; r.has_crc = has_crc;
    r.corrected = fixed; r.uncorrectable = bad;
    if (!r.hdr_ok) { r.n_bytes = -1; a.results[p] = r; return; }
    const int need = at + 2 * (int)len + (has_crc ? 4 : 0);
    int o = 8;
    while (nn < need) {
        if (o + 4 + (int)cr > nin) { r.n_bytes = -3; r.corrected = fixed; r.uncorrectable = bad; a.results[p] = r; return; }
        std_deinterleave(sym + o, app, 4 + cr, g.sf, g.ldro != 0, cw);
        for (unsigned k = 0; k < app; ++k) nib[nn++] = (uint8_t)std_hamming_decode(cw[k], cr, fixed, bad);
        o += 4 + (int)cr;
    }
    unsigned w = 0xFF;
    for (unsigned i = 0; i < len; ++i) {
        out[i] = (uint8_t)((nib[at + 2 * i] | nib[at + 2 * i + 1] << 4) ^ w);
        w = whiten_next(w);
    }
    if (has_crc) {
        const int q = at + 2 * (int)len;
        const unsigned rx = nib[q] | nib[q + 1] << 4 | nib[q + 2] << 8 | (unsigned)nib[q + 3] << 12;
        r.crc_ok = rx == std_crc(out, len);
    }
    r.corrected = fixed; r.uncorrectable = bad;
    r.n_bytes = (int16_t)len;
    r.symbols_used = (uint16_t)o;
    a.results[p] = r;
}

__global__ void std_encode_kernel(StdEncodeArgs a)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n_packets) return;
    const lora_b200_std_cfg& g = a.cfg;
    const uint8_t* pay = a.payload + (size_t)p * a.payload_stride;
    uint16_t* out = a.symbols + (size_t)p * a.symbol_stride;
    const unsigned len = (unsigned)a.payload_len;
    uint8_t nib[kMaxNib];
    int nn = 0;
    if (!g.implicit_header) {
        const unsigned n0 = len >> 4, n1 = len & 15u, n2 = g.cr << 1 | (g.has_crc ? 1u : 0u);
        const unsigned chk = std_header_checksum(n0, n1, n2);
        nib[nn++] = (uint8_t)n0; nib[nn++] = (uint8_t)n1; nib[nn++] = (uint8_t)n2;
        nib[nn++] = (uint8_t)(chk >> 4); nib[nn++] = (uint8_t)(chk & 15u);
    }
    unsigned w = 0xFF;
    for (unsigned i = 0; i < len; ++i) {
        const unsigned b = pay[i] ^ w;
        w = whiten_next(w);
        nib[nn++] = (uint8_t)(b & 15u); nib[nn++] = (uint8_t)(b >> 4);
    }
    if (g.has_crc) {
        const unsigned crc = std_crc(pay, len);
        nib[nn++] = (uint8_t)(crc & 15u); nib[nn++] = (uint8_t)(crc >> 4 & 15u);
        nib[nn++] = (uint8_t)(crc >> 8 & 15u); nib[nn++] = (uint8_t)(crc >> 12 & 15u);
    }
    unsigned cw[12];
    int at = 0, o = 8;
    const unsigned first = g.sf - 2, app = g.ldro ? g.sf - 2 : g.sf;
    for (unsigned k = 0; k < first; ++k, ++at) cw[k] = std_hamming_encode(at < nn ? nib[at] : 0u, 4);
    std_interleave(cw, first, 8, g.sf, true, out);
    while (o < (int)a.n_symbols) {
        for (unsigned k = 0; k < app; ++k, ++at) cw[k] = std_hamming_encode(at < nn ? nib[at] : 0u, g.cr);
        std_interleave(cw, app, 4 + g.cr, g.sf, g.ldro != 0, out + o);
        o += 4 + (int)g.cr;
    }
}

inline unsigned blocks_for(int n) { return (unsigned)((n + 127) / 128); }

} // namespace

size_t std_symbol_count(const lora_b200_std_cfg& c, size_t len)
{
    if (c.sf < 7 || c.sf > 12 || c.cr < 1 || c.cr > 4 || len > 255) return 0;
    const size_t first = c.sf - 2, app = c.ldro ? c.sf - 2 : c.sf;
    const size_t nn = (c.implicit_header ? 0 : 5) + 2 * len + (c.has_crc ? 4 : 0);
    const size_t rest = nn > first ? nn - first : 0;
    return 8 + (rest + app - 1) / app * (4 + c.cr);
}

cudaError_t launch_std_decode(const StdDecodeArgs& a, cudaStream_t st)
{
    std_decode_kernel<<<blocks_for(a.n_packets), 128, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

cudaError_t launch_std_encode(const StdEncodeArgs& a, cudaStream_t st)
{
    std_encode_kernel<<<blocks_for(a.n_packets), 128, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

} // namespace lora_b200
