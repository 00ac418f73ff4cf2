// lora_rx_stream -- receive a continuous IQ capture: frame synchroniser + the LoRa coding chain as deployed on the air.
//
// The companion of the reference's rx_runner (runners/rx_runner.cpp) for captures that hold real frames -- preamble,
// sync word, down-chirps, explicit header -- such as the reference's own golden vector
//   vectors_binary/bw_125k_sf_7_cr_1_ldro_false_crc_true_implheader_false.unknown
// which rx_runner can only cut into aligned windows.  Same input format (raw float32 I,Q pairs, a file
// or stdin), same flag style.  Everything runs on the GPU through the C ABI (lora_b200_std_receive_stream_host);
// there is no CPU path.
//
//   lora_rx_stream --in=FILE|stdin [--sf=7] [--bw=125000] [--osr=1] [--sync=0x12] [--ldro] [--implicit=LEN --cr=1 --crc]
//                  [--mtu=SYMS] [--max-packets=64]
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <iterator>
#include <string>
#include <vector>

#include "lora_phy_b200.h"

int main(int argc, char** argv)
{
    std::string in;
    unsigned sf = 7, bw = 125000, osr = 1, sync = 0x12, cr = 1, mtu = 0, max_packets = 64, implicit_len = 0;
    bool ldro = false, implicit = false, crc = false;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto val = [&](const char* key) -> const char* {
            const size_t n = std::strlen(key);
            return a.compare(0, n, key) == 0 ? a.c_str() + n : nullptr;
        };
        if (const char* v = val("--in=")) in = v;
        else if (const char* v = val("--sf=")) sf = std::strtoul(v, nullptr, 0);
        else if (const char* v = val("--bw=")) bw = std::strtoul(v, nullptr, 0);
        else if (const char* v = val("--osr=")) osr = std::strtoul(v, nullptr, 0);
        else if (const char* v = val("--sync=")) sync = std::strtoul(v, nullptr, 0);
        else if (const char* v = val("--cr=")) cr = std::strtoul(v, nullptr, 0);
        else if (const char* v = val("--mtu=")) mtu = std::strtoul(v, nullptr, 0);
        else if (const char* v = val("--max-packets=")) max_packets = std::strtoul(v, nullptr, 0);
        else if (const char* v = val("--implicit=")) { implicit = true; implicit_len = std::strtoul(v, nullptr, 0); }
        else if (a == "--ldro") ldro = true;
        else if (a == "--crc") crc = true;
        else { std::cerr << "Unknown argument: " << a << "\n"; return 1; }
    }
    if (in.empty()) { std::cerr << "Usage: lora_rx_stream --in=FILE|stdin [--sf=7] [--bw=125000] [--osr=1] [--sync=0x12] [--ldro]\n"; return 1; }

    std::vector<char> raw;
    if (in == "stdin") raw.assign(std::istreambuf_iterator<char>(std::cin), std::istreambuf_iterator<char>());
    else {
        std::ifstream f(in, std::ios::binary);
        if (!f) { std::cerr << "Cannot open " << in << "\n"; return 1; }
        raw.assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
    }
    const size_t count = raw.size() / sizeof(lora_b200_c32);
    if (count == 0) { std::cerr << "No samples\n"; return 1; }

    lora_b200_config cfg{};
    cfg.sf = sf; cfg.bw_hz = bw; cfg.osr = osr; cfg.window = 0; cfg.device = -1;
    lora_b200_ctx* ctx = nullptr;
    if (lora_b200_ctx_create(&cfg, &ctx) != LORA_B200_OK) { std::cerr << "lora_b200: " << lora_b200_last_error() << "\n"; return 1; }

    lora_b200_std_cfg g{};
    g.sf = sf; g.cr = cr; g.ldro = ldro; g.implicit_header = implicit; g.has_crc = crc; g.payload_len = implicit_len;
    lora_b200_stream_cfg sc{};
    sc.sync = static_cast<uint8_t>(sync); sc.thresh_db = -30.0f;
    sc.mtu = mtu; // 0: frame length from the headers found in the capture
    sc.max_packets = max_packets;
    sc.max_symbols = static_cast<uint32_t>(count / ((size_t(1) << sf) * osr) + 8);

    std::vector<uint8_t> payload(size_t(LORA_B200_PAYLOAD_STRIDE) * max_packets);
    std::vector<lora_b200_std_result> res(max_packets);
    std::vector<lora_b200_stream_packet> pk(max_packets);
    uint32_t n = 0;
    if (lora_b200_std_receive_stream_host(ctx, &g, &sc, reinterpret_cast<const lora_b200_c32*>(raw.data()), count,
                                          payload.data(), res.data(), pk.data(), &n) != LORA_B200_OK) {
        std::cerr << "lora_b200: " << lora_b200_last_error() << "\n";
        return 1;
    }
    std::printf("Frames: %u\n", n);
    for (uint32_t k = 0; k < n; ++k) {
        const lora_b200_std_result& r = res[k];
        std::printf("Frame %u @%llu:", k, static_cast<unsigned long long>(pk[k].start));
        if (r.n_bytes < 0) { std::printf(" %s\n", r.n_bytes == -1 ? "header invalid" : r.n_bytes == -3 ? "truncated" : "too short"); continue; }
        std::printf(" len %d CR 4/%u CRC %s%s Payload: ", r.n_bytes, 4u + r.hdr_cr, !r.has_crc ? "none" : r.crc_ok ? "ok" : "BAD",
                    r.uncorrectable ? " (uncorrectable codewords)" : r.corrected ? " (corrected)" : "");
        const uint8_t* p = payload.data() + size_t(LORA_B200_PAYLOAD_STRIDE) * k;
        for (int i = 0; i < r.n_bytes; ++i) std::printf("%02x", p[i]);
        std::printf("  \"");
        for (int i = 0; i < r.n_bytes; ++i) std::putchar(p[i] >= 32 && p[i] < 127 ? p[i] : '.');
        std::printf("\"\n");
    }
    lora_b200_ctx_destroy(ctx);
    return 0;
}
