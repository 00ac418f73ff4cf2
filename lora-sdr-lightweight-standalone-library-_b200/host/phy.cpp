// host/phy.cpp -- the `lora_phy` C++ API (include/lora_phy/phy.hpp) over the C ABI of liblora_phy_b200.so.
//
// Host side of the drop-in: argument checks, return conventions and metric bookkeeping follow the reference's
// src/phy/phy.cpp, LoRaDemod.cpp, LoRaMod.cpp, LoRaEncoder.cpp and LoRaDecoder.cpp (cited per function); all
// arithmetic happens in the CUDA kernels.  Contexts (device tables + staging buffers) are shared: the reference API
// has no "destroy" for lora_workspace, so a process-wide registry keyed by (sf, bw, osr, window) owns them and
// releases them at exit.  A call that cannot reach a CUDA device fails the way the reference signals bad
// arguments (-1, or 0 items for the size_t helpers); it never computes on the CPU.
#include <lora_phy/phy.hpp>

#include <cmath>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "lora_phy_b200.h"

namespace lora_phy {

namespace {

struct Registry {
    std::mutex mu;
    std::map<std::tuple<unsigned, unsigned, unsigned, int>, lora_b200_ctx*> ctx;
    ~Registry()
    {
        for (auto& kv : ctx) lora_b200_ctx_destroy(kv.second);
    }
    lora_b200_ctx* get(unsigned sf, unsigned bw_hz, unsigned osr, bool hann)
    {
        std::lock_guard<std::mutex> lk(mu);
        const auto key = std::make_tuple(sf, bw_hz, osr ? osr : 1u, hann ? 1 : 0);
        auto it = ctx.find(key);
        if (it != ctx.end()) return it->second;
        lora_b200_config cfg{sf, bw_hz, osr ? osr : 1u, hann ? 1 : 0, -1};
        lora_b200_ctx* c = nullptr;
        if (lora_b200_ctx_create(&cfg, &c) != LORA_B200_OK) return nullptr;
        ctx.emplace(key, c);
        return c;
    }
};

Registry& registry()
{
    static Registry r;
    return r;
}

unsigned deduce_sf(size_t nfft) // phy.cpp:13-18: sf is not stored, it is recovered from the plan length
{
    unsigned sf = 0;
    while ((size_t(1) << sf) < nfft) ++sf;
    return sf;
}

void fill_window(float* w, size_t n, window_type kind) // phy.cpp:36-47, LoRaDemod.cpp:16-25 (host table for callers)
{
    if (kind == window_type::window_hann) {
        for (size_t i = 0; i < n; ++i)
            w[i] = 0.5f - 0.5f * std::cos(2.0f * float(M_PI) * static_cast<float>(i) / (static_cast<float>(n) - 1.0f));
    } else {
        for (size_t i = 0; i < n; ++i) w[i] = 1.0f;
    }
}

// The context a workspace call runs on.  demodulate() applies the window only when a window buffer was supplied
// AND a window kind was configured (phy.cpp:222-223), so that pair selects the Hann or the plain context.
lora_b200_ctx* ctx_of(const lora_workspace* ws)
{
    const unsigned sf = deduce_sf(ws->plan_fwd.nfft);
    const bool hann = ws->window && ws->window_kind == window_type::window_hann;
    return registry().get(sf, static_cast<unsigned>(ws->bw), ws->osr ? ws->osr : 1u, hann);
}

} // namespace

// ---- workspace API ------------------------------------------------------------------------------------------------
int init(lora_workspace* ws, const lora_params* cfg) // phy.cpp:26-49
{
    if (!ws || !cfg) return -1;
    const size_t n = size_t(1) << cfg->sf;
    ws->plan_fwd.nfft = n;
    ws->plan_fwd.inverse = false;
    ws->plan_inv.nfft = n;
    ws->plan_inv.inverse = true;
    ws->metrics = {};
    ws->osr = cfg->osr ? cfg->osr : 1u;
    ws->bw = cfg->bw;
    ws->sync_word = cfg->sync_word;
    ws->window_kind = cfg->window;
    if (ws->window) fill_window(ws->window, n, ws->window_kind);
    // the reference cannot fail here; a missing GPU (or an sf this build has no kernel for) can
    lora_b200_ctx* c = ctx_of(ws);
    ws->plan_fwd.device = c;
    ws->plan_inv.device = c;
    return c ? 0 : -1;
}

void reset(lora_workspace* ws) // phy.cpp:51-53
{
    if (ws) ws->metrics = {};
}

ssize_t encode(lora_workspace* ws, const uint8_t* payload, size_t payload_len, uint16_t* symbols, size_t symbol_cap)
{ // phy.cpp:55-63
    if (!ws || !payload || !symbols) return -1;
    lora_b200_ctx* c = ctx_of(ws);
    size_t produced = 0;
    if (!c || lora_b200_encode_host(c, payload, payload_len, symbols, symbol_cap, &produced) != LORA_B200_OK) return -1;
    return static_cast<ssize_t>(produced);
}

ssize_t decode(lora_workspace* ws, const uint16_t* symbols, size_t symbol_count, uint8_t* payload, size_t payload_cap)
{ // phy.cpp:241-256
    if (!ws || !symbols || !payload) return -1;
    lora_b200_ctx* c = ctx_of(ws);
    size_t produced = 0;
    int crc = 0;
    if (!c || lora_b200_decode_host(c, symbols, symbol_count, payload, payload_cap, &produced, &crc) != LORA_B200_OK)
        return -1;
    ws->metrics.crc_ok = crc != 0;
    return static_cast<ssize_t>(produced);
}

ssize_t modulate(lora_workspace* ws, const uint16_t* symbols, size_t symbol_count, std::complex<float>* iq, size_t iq_cap)
{ // phy.cpp:65-76 (the reference would overrun iq before noticing iq_cap; here nothing is written on failure)
    if (!ws || !symbols || !iq) return -1;
    lora_b200_ctx* c = ctx_of(ws);
    size_t produced = 0;
    if (!c || lora_b200_modulate_host(c, symbols, symbol_count, 1.0f, ws->sync_word,
                                      reinterpret_cast<lora_b200_c32*>(iq), iq_cap, &produced) != LORA_B200_OK)
        return -1;
    return static_cast<ssize_t>(produced);
}

void estimate_offsets(lora_workspace* ws, const std::complex<float>* samples, size_t sample_count)
{ // phy.cpp:78-145; writes metrics only when at least one whole symbol was offered (:83-88)
    if (!ws || !samples || sample_count == 0) return;
    lora_b200_ctx* c = ctx_of(ws);
    if (!c) return;
    const size_t step = ws->plan_fwd.nfft * (ws->osr ? ws->osr : 1u);
    if (step == 0 || sample_count / step == 0) return;
    float cfo = ws->metrics.cfo, to = ws->metrics.time_offset;
    if (lora_b200_estimate_host(c, reinterpret_cast<const lora_b200_c32*>(samples), sample_count, &cfo, &to) != LORA_B200_OK)
        return;
    ws->metrics.cfo = cfo;
    ws->metrics.time_offset = to;
}

void compensate_offsets(const lora_workspace* ws, std::complex<float>* samples, size_t sample_count)
{ // phy.cpp:147-176
    if (!ws || !samples || sample_count == 0) return;
    lora_b200_ctx* c = ctx_of(ws);
    if (!c) return;
    lora_b200_compensate_host(c, ws->metrics.cfo, ws->metrics.time_offset, reinterpret_cast<lora_b200_c32*>(samples),
                              sample_count);
}

ssize_t demodulate(lora_workspace* ws, const std::complex<float>* iq, size_t sample_count, uint16_t* symbols,
                   size_t symbol_cap)
{ // phy.cpp:178-239
    if (!ws || !iq || !symbols) return -1;
    const size_t step = ws->plan_fwd.nfft * (ws->osr ? ws->osr : 1u);
    if (step == 0 || sample_count % step != 0) return -1;
    const size_t total = sample_count / step;
    if (total < 2) return -1;
    if (total - 2 > symbol_cap) return -1;
    lora_b200_ctx* c = ctx_of(ws);
    if (!c) return -1;
    size_t n = 0;
    lora_b200_pkt_metrics m{};
    if (lora_b200_demodulate_host(c, LORA_B200_DEMOD_WORKSPACE, reinterpret_cast<const lora_b200_c32*>(iq), sample_count,
                                  symbols, symbol_cap, &n, &m) != LORA_B200_OK)
        return -1;
    ws->metrics.cfo = m.cfo;                 // estimate_offsets ran on the first two symbols (:192-193)
    ws->metrics.time_offset = m.time_offset;
    ws->sync_word = m.sync_word;             // :235-237 overwrite the configured sync word
    return static_cast<ssize_t>(n);
}

const lora_metrics* get_last_metrics(const lora_workspace* ws) // phy.cpp:258-261
{
    return ws ? &ws->metrics : nullptr;
}

// ---- legacy helpers -------------------------------------------------------------------------------------------------
void lora_demod_init(lora_demod_workspace* ws, unsigned sf, window_type win, std::complex<float>* scratch, size_t max_samples)
{ // LoRaDemod.cpp:10-32
    ws->N = size_t(1) << sf;
    ws->window_kind = win;
    if (ws->N <= lora_demod_workspace::MAX_N) fill_window(ws->window, ws->N, win);
    ws->fft_plan.nfft = ws->N;
    ws->fft_plan.inverse = false;
    ws->fft_plan.device = nullptr; // attached at the first lora_demodulate: osr is only known there
    ws->scratch = scratch;
    ws->scratch_len = max_samples;
}

void lora_demod_free(lora_demod_workspace* ws) // LoRaDemod.cpp:34-47 (contexts stay with the registry)
{
    ws->fft_plan.device = nullptr;
    ws->N = 0;
    ws->scratch = nullptr;
    ws->scratch_len = 0;
}

size_t lora_modulate(const uint16_t* symbols, size_t symbol_count, std::complex<float>* out_samples, unsigned sf,
                     unsigned osr, bandwidth bw, float amplitude, uint8_t sync)
{ // LoRaMod.cpp:8-43; `symbols == nullptr` with a zero count is legal (tests/sync_word_test.cpp:47)
    lora_b200_ctx* c = registry().get(sf, static_cast<unsigned>(bw), osr, false);
    const size_t produced = (symbol_count + 2) * (size_t(1) << sf) * (osr ? osr : 1u);
    size_t n = 0;
    if (!c || lora_b200_modulate_host(c, symbols, symbol_count, amplitude, sync,
                                      reinterpret_cast<lora_b200_c32*>(out_samples), produced, &n) != LORA_B200_OK)
        return 0;
    return n;
}

size_t lora_demodulate(lora_demod_workspace* ws, const std::complex<float>* samples, size_t sample_count,
                       uint16_t* out_symbols, unsigned osr, uint8_t* out_sync)
{ // LoRaDemod.cpp:49-195
    const unsigned sf = deduce_sf(ws->N);
    lora_b200_ctx* c = registry().get(sf, 125000u, osr, ws->window_kind == window_type::window_hann);
    if (!c) return 0;
    ws->fft_plan.device = c;
    const size_t step = ws->N * (osr ? osr : 1u);
    const size_t total = step ? sample_count / step : 0;
    // grow-only, per calling thread: no allocation once a packet length has been seen (tests/no_alloc_test.cpp's rule
    // for the steady state)
    static thread_local std::vector<uint16_t> tmp;
    if (tmp.size() < total + 1) tmp.resize(total + 1);
    size_t n = 0;
    lora_b200_pkt_metrics m{};
    if (lora_b200_demodulate_host(c, LORA_B200_DEMOD_LEGACY, reinterpret_cast<const lora_b200_c32*>(samples), sample_count,
                                  tmp.data(), tmp.size(), &n, &m) != LORA_B200_OK)
        return 0;
    // :68-71 samples beyond [-1, 1] are normalised into the scratch buffer; without one the call gives up
    if (m.normalised && (!ws->scratch || ws->scratch_len < sample_count)) return 0;
    for (size_t i = 0; i < n; ++i) out_symbols[i] = tmp[i];
    ws->metrics.cfo = m.cfo;                 // :133-135
    ws->metrics.time_offset = m.time_offset;
    if (out_sync) *out_sync = total >= 2 ? m.sync_word : 0; // :177-192
    return total >= 2 ? n : total;           // :194
}

size_t lora_encode(const uint8_t* bytes, size_t byte_count, uint16_t* out_symbols, unsigned sf)
{ // src/phy/LoRaEncoder.cpp:6-18 (sf is ignored there too, beyond selecting a context here)
    lora_b200_ctx* c = registry().get(sf >= 2 && sf <= 12 ? sf : 7u, 125000u, 1u, false);
    size_t produced = 0;
    if (!c || lora_b200_encode_host(c, bytes, byte_count, out_symbols, 2 * byte_count, &produced) != LORA_B200_OK) return 0;
    return produced;
}

size_t lora_decode(const uint16_t* symbols, size_t symbol_count, uint8_t* out_bytes)
{ // src/phy/LoRaDecoder.cpp:6-19
    lora_b200_ctx* c = registry().get(7u, 125000u, 1u, false);
    size_t produced = 0;
    int crc = 0;
    if (!c || lora_b200_decode_host(c, symbols, symbol_count, out_bytes, symbol_count / 2, &produced, &crc) != LORA_B200_OK)
        return 0;
    return produced;
}

// ---- kissfft<float>::transform / LoRaDetector<float>::detect (include/lora_phy/kissfft.hh, LoRaDetector.hpp) ----------
namespace detail {

static lora_b200_ctx* ctx_for_length(std::size_t nfft)
{
    const unsigned sf = deduce_sf(nfft);
    if (sf < 2 || sf > 12 || (std::size_t(1) << sf) != nfft) return nullptr;
    return registry().get(sf, 125000u, 1u, false);
}

int device_fft(std::size_t nfft, bool inverse, const std::complex<float>* src, std::complex<float>* dst)
{ // kissfft.hh:100-103
    lora_b200_ctx* c = ctx_for_length(nfft);
    if (!c || !src || !dst) return -1;
    return lora_b200_fft_detect_host(c, reinterpret_cast<const lora_b200_c32*>(src), 1, inverse ? 1 : 0,
                                     reinterpret_cast<lora_b200_c32*>(dst), nullptr) == LORA_B200_OK ? 0 : -1;
}

int device_detect(std::size_t nfft, const std::complex<float>* src, std::complex<float>* spectrum, std::size_t* index,
                  float* power, float* power_avg, float* f_index)
{ // LoRaDetector.hpp:39-74
    lora_b200_ctx* c = ctx_for_length(nfft);
    if (!c || !src) return -1;
    lora_b200_detect_result r{};
    if (lora_b200_fft_detect_host(c, reinterpret_cast<const lora_b200_c32*>(src), 1, 0,
                                  reinterpret_cast<lora_b200_c32*>(spectrum), &r) != LORA_B200_OK)
        return -1;
    if (index) *index = r.index;
    if (power) *power = r.power;
    if (power_avg) *power_avg = r.power_avg;
    if (f_index) *f_index = r.f_index;
    return 0;
}

} // namespace detail

} // namespace lora_phy
