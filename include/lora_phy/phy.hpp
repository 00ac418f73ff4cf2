// lora_phy/phy.hpp -- the `lora_phy` C++ API (reference include/lora_phy/phy.hpp:29-215), B200 build.
//
// Same namespace, type names, public fields, defaults, function names, argument meaning and return conventions as
// the reference header, so runners/rx_runner.cpp and runners/tx_runner.cpp of the reference compile against this
// file unchanged and link against liblora_phy.so (host/phy.cpp), which forwards every call to the CUDA kernels
// through the C ABI of include/lora_phy_b200.h.  Differences a caller can observe:
//   * the FFT plans inside the workspaces are handles (lora_phy/kissfft.hh), not host tables;
//   * every entry point needs a CUDA device: there is no CPU path, a call without one returns -1 (or 0 symbols
//     for the size_t-returning legacy helpers);
//   * decode() checks payload_cap BEFORE writing (the reference writes first, phy.cpp:245-246).
// Everything else, including the reference's quirks, is kept: failures return -1 rather than -EINVAL/-ERANGE
// (phy.cpp:27,58,61,68,74,181-190), demodulate() estimates its offsets on the raw first two symbols and overwrites
// ws->sync_word (phy.cpp:192-193,235-237), init() neither validates nor stores sf (phy.cpp:13-18,26-49).
#pragma once

#include <sys/types.h>

#include <complex>
#include <cstddef>
#include <cstdint>

#include <lora_phy/LoRaDetector.hpp>
#include <lora_phy/kissfft.hh>

namespace lora_phy {

enum class window_type { window_none, window_hann };                                 // reference phy.hpp:29-32

enum class bandwidth : unsigned { bw_125 = 125000, bw_250 = 250000, bw_500 = 500000 }; // :37-41

constexpr float bw_to_hz(bandwidth bw) { return static_cast<float>(static_cast<unsigned>(bw)); }
constexpr float bw_scale(bandwidth bw) { return bw_to_hz(bw) / 125000.0f; }          // :47-49

// Modulation / coding options, copied by init() (:51-58).
struct lora_params {
    unsigned sf{};
    bandwidth bw{bandwidth::bw_125};
    unsigned cr{};       // parsed by the runners, never used by the library
    unsigned osr{1};
    window_type window{window_type::window_none};
    uint8_t sync_word{0x12};
};

// Results of the last demodulate / decode / estimate_offsets call (:65-69).
struct lora_metrics {
    bool crc_ok{};
    float cfo{};
    float time_offset{};
};

// Caller-owned state of the workspace API (:77-92).  The three buffers are kept for source compatibility; this
// build stages samples in device memory owned by the shared context and does not read or write them, except
// `window`, which init() fills exactly as the reference does when it is non-null.
struct lora_workspace {
    uint16_t* symbol_buf{};
    std::complex<float>* fft_in{};
    std::complex<float>* fft_out{};

    float* window{};
    window_type window_kind{window_type::window_none};

    kissfft_plan<float> plan_fwd{}; // nfft = 1 << sf; device = the context init() attached
    kissfft_plan<float> plan_inv{};

    lora_metrics metrics{};
    unsigned osr{1};
    bandwidth bw{bandwidth::bw_125};
    uint8_t sync_word{0x12};
};

// ---- workspace API (:102-156) --------------------------------------------------------------------------------
int init(lora_workspace* ws, const lora_params* cfg);
void reset(lora_workspace* ws);
ssize_t encode(lora_workspace* ws, const uint8_t* payload, size_t payload_len, uint16_t* symbols, size_t symbol_cap);
ssize_t decode(lora_workspace* ws, const uint16_t* symbols, size_t symbol_count, uint8_t* payload, size_t payload_cap);
ssize_t modulate(lora_workspace* ws, const uint16_t* symbols, size_t symbol_count, std::complex<float>* iq, size_t iq_cap);
ssize_t demodulate(lora_workspace* ws, const std::complex<float>* iq, size_t sample_count, uint16_t* symbols,
                   size_t symbol_cap);
void estimate_offsets(lora_workspace* ws, const std::complex<float>* samples, size_t sample_count);
void compensate_offsets(const lora_workspace* ws, std::complex<float>* samples, size_t sample_count);
const lora_metrics* get_last_metrics(const lora_workspace* ws);

// ---- legacy helpers (:170-215) ---------------------------------------------------------------------------------
// Same public fields as the reference's struct; the embedded FFT objects are replaced by the context handle.
struct lora_demod_workspace {
    static const size_t MAX_N = kissfft_utils::KISSFFT_MAX_N;
    size_t N{};
    float window[MAX_N];            // filled by lora_demod_init as LoRaDemod.cpp:14-25 fills it
    window_type window_kind{window_type::window_none};
    kissfft_plan<float> fft_plan{}; // handle
    lora_metrics metrics{};
    std::complex<float>* scratch{}; // the reference normalises into this buffer; here only its size is honoured
    size_t scratch_len{};
};

void lora_demod_init(lora_demod_workspace* ws, unsigned sf, window_type win = window_type::window_none,
                     std::complex<float>* scratch = nullptr, size_t max_samples = 0);
void lora_demod_free(lora_demod_workspace* ws);
size_t lora_modulate(const uint16_t* symbols, size_t symbol_count, std::complex<float>* out_samples, unsigned sf,
                     unsigned osr, bandwidth bw, float amplitude = 1.0f, uint8_t sync = 0x12);
size_t lora_demodulate(lora_demod_workspace* ws, const std::complex<float>* samples, size_t sample_count,
                       uint16_t* out_symbols, unsigned osr, uint8_t* out_sync = nullptr);
size_t lora_encode(const uint8_t* bytes, size_t byte_count, uint16_t* out_symbols, unsigned sf);
size_t lora_decode(const uint16_t* symbols, size_t symbol_count, uint8_t* out_bytes);

} // namespace lora_phy
