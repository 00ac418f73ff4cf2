// ref_shim.cpp -- extern "C" doors into the UNMODIFIED reference, for oracle validation only.
//
// TEST INFRASTRUCTURE ONLY.  This file contains no algorithm: every function forwards to the
// reference's own code, which oracle/Makefile compiles where it lies under /root/reference
// (src/phy/*.cpp + include/lora_phy/*.hpp) into oracle/_ref/liblora_ref.so.  No reference source
// is copied into this repository.  The signatures mirror lora_oracle.h (orc_X <-> ref_X) so the
// tests can call both with the same arguments and compare bits.
#include <lora_phy/phy.hpp>
#include <lora_phy/ChirpGenerator.hpp>
#include <lora_phy/LoRaCodes.hpp>
#include <lora_phy/LoRaDetector.hpp>
#include <lora_phy/kissfft.hh>

#include <chrono>
#include <complex>
#include <cstring>
#include <thread>
#include <vector>

typedef std::complex<float> cf;

static lora_phy::bandwidth bw_of(unsigned hz) { return static_cast<lora_phy::bandwidth>(hz); }

extern "C" {

// ---- LoRaCodes.hpp ---------------------------------------------------------------------------------
unsigned ref_round_up(unsigned a, unsigned b) { return roundUp(a, b); }
uint8_t ref_checksum8(const uint8_t* p, size_t n) { return checksum8(p, n); }
uint8_t ref_header_checksum(const uint8_t* h) { return headerChecksum(h); }
uint16_t ref_crc16sx(uint16_t crc, uint16_t poly) { return crc16sx(crc, poly); }
uint8_t ref_xsum8(uint8_t t) { return xsum8(t); }
uint16_t ref_data_checksum(const uint8_t* d, int n) { return sx1272DataChecksum(d, n); }
void ref_whiten_sx1232(uint8_t* b, uint16_t n) { SX1232RadioComputeWhitening(b, n); }
void ref_whiten_table(uint8_t* b, uint16_t n, int ofs, int rdd) { Sx1272ComputeWhitening(b, n, ofs, rdd); }
void ref_whiten_lfsr(uint8_t* b, uint16_t n, int ofs, size_t rdd) { Sx1272ComputeWhiteningLfsr(b, n, ofs, rdd); }
uint16_t ref_bin_to_gray(uint16_t v) { return binaryToGray16(v); }
uint16_t ref_gray_to_bin(uint16_t v) { return grayToBinary16(v); }
uint8_t ref_enc_h84(uint8_t x) { return encodeHamming84sx(x); }
uint8_t ref_dec_h84(uint8_t b, int* error, int* bad)
{
    bool e = *error != 0, d = *bad != 0;
    const uint8_t r = decodeHamming84sx(b, e, d);
    *error = e; *bad = d;
    return r;
}
uint8_t ref_enc_h74(uint8_t x) { return encodeHamming74sx(x); }
uint8_t ref_dec_h74(uint8_t b, int* error)
{
    bool e = *error != 0;
    const uint8_t r = decodeHamming74sx(b, e);
    *error = e;
    return r;
}
uint8_t ref_chk_p54(uint8_t b, int* error) { bool e = *error != 0; const uint8_t r = checkParity54(b, e); *error = e; return r; }
uint8_t ref_enc_p54(uint8_t b) { return encodeParity54(b); }
uint8_t ref_chk_p64(uint8_t b, int* error) { bool e = *error != 0; const uint8_t r = checkParity64(b, e); *error = e; return r; }
uint8_t ref_enc_p64(uint8_t b) { return encodeParity64(b); }
void ref_interleave(const uint8_t* cw, size_t ncw, uint16_t* sym, size_t ppm, size_t rdd) { diagonalInterleaveSx(cw, ncw, sym, ppm, rdd); }
void ref_deinterleave(const uint16_t* sym, size_t nsym, uint8_t* cw, size_t ppm, size_t rdd) { diagonalDeterleaveSx(sym, nsym, cw, ppm, rdd); }
void ref_deinterleave2(const uint16_t* sym, size_t nsym, uint8_t* cw, size_t ppm, size_t rdd) { diagonalDeterleaveSx2(sym, nsym, cw, ppm, rdd); }

// ---- ChirpGenerator.hpp / kissfft.hh / LoRaDetector.hpp ---------------------------------------------------
int ref_gen_chirp(cf* out, int N, int osr, int NN, float f0, int down, float ampl, float* phase, float bw_scale)
{
    return genChirp(out, N, osr, NN, f0, down != 0, ampl, *phase, bw_scale);
}

void ref_fft_twiddles(cf* tw, int nfft, int inverse)
{
    static kissfft_plan<float> plan;
    kissfft<float>::init(plan, nfft, inverse != 0);
    std::memcpy(tw, plan.twiddles, sizeof(cf) * nfft);
}

int ref_fft_factor(int nfft, int* radix, int* remainder)
{
    static kissfft_plan<float> plan;
    kissfft<float>::init(plan, nfft, false);
    for (int i = 0; i < plan.stages; ++i) { radix[i] = plan.stageRadix[i]; remainder[i] = plan.stageRemainder[i]; }
    return plan.stages;
}

void ref_fft(int nfft, int inverse, const cf* src, cf* dst)
{
    static kissfft_plan<float> plan;
    kissfft<float>::init(plan, nfft, inverse != 0);
    kissfft<float> fft(plan);
    fft.transform(src, dst);
}

size_t ref_detect(size_t N, const cf* in, cf* spectrum, float* power, float* power_avg, float* f_index)
{
    static kissfft_plan<float> plan;
    kissfft<float>::init(plan, static_cast<int>(N), false);
    kissfft<float> fft(plan);
    std::vector<cf> fin(in, in + N);
    LoRaDetector<float> det(N, fin.data(), spectrum, fft);
    return det.detect(*power, *power_avg, *f_index);
}

// ---- src/phy ----------------------------------------------------------------------------------------
size_t ref_lora_encode(const uint8_t* b, size_t n, uint16_t* s) { return lora_phy::lora_encode(b, n, s, 7); }
size_t ref_lora_decode(const uint16_t* s, size_t n, uint8_t* b) { return lora_phy::lora_decode(s, n, b); }

size_t ref_lora_modulate(const uint16_t* symbols, size_t n, cf* out, unsigned sf, unsigned osr,
                         unsigned bw_hz, float amplitude, uint8_t sync)
{
    return lora_phy::lora_modulate(symbols, n, out, sf, osr, bw_of(bw_hz), amplitude, sync);
}

struct ref_params { unsigned sf, bw_hz, osr; int hann, has_window; uint8_t sync_word; };
struct ref_metrics { int crc_ok; float cfo, time_offset; uint8_t sync_word; };

// A workspace set up exactly as runners/rx_runner.cpp:93-105 does (plus the optional window buffer).
struct ref_ws {
    lora_phy::lora_workspace ws;
    std::vector<uint16_t> symbuf;
    std::vector<cf> fin, fout;
    std::vector<float> window;
    ref_ws(const ref_params* p) : ws()
    {
        const size_t N = size_t(1) << p->sf;
        symbuf.resize(N); fin.resize(N); fout.resize(N);
        ws.symbol_buf = symbuf.data(); ws.fft_in = fin.data(); ws.fft_out = fout.data();
        if (p->has_window) { window.resize(N); ws.window = window.data(); }
        lora_phy::lora_params cfg{};
        cfg.sf = p->sf; cfg.bw = bw_of(p->bw_hz); cfg.osr = p->osr;
        cfg.window = p->hann ? lora_phy::window_type::window_hann : lora_phy::window_type::window_none;
        cfg.sync_word = p->sync_word;
        lora_phy::init(&ws, &cfg);
    }
};

long ref_decode(const uint16_t* symbols, size_t n, uint8_t* payload, size_t cap, int* crc_ok)
{
    ref_params p = {7, 125000, 1, 0, 0, 0x12};
    ref_ws w(&p);
    const long r = lora_phy::decode(&w.ws, symbols, n, payload, cap);
    *crc_ok = lora_phy::get_last_metrics(&w.ws)->crc_ok;
    return r;
}

void ref_estimate_offsets(const ref_params* p, const cf* samples, size_t count, float* cfo, float* time_offset)
{
    ref_ws w(p);
    w.ws.metrics.cfo = *cfo; w.ws.metrics.time_offset = *time_offset;
    lora_phy::estimate_offsets(&w.ws, samples, count);
    *cfo = w.ws.metrics.cfo; *time_offset = w.ws.metrics.time_offset;
}

void ref_compensate_offsets(const ref_params* p, float cfo, float time_offset, cf* samples, size_t count)
{
    ref_ws w(p);
    w.ws.metrics.cfo = cfo; w.ws.metrics.time_offset = time_offset;
    lora_phy::compensate_offsets(&w.ws, samples, count);
}

long ref_demodulate(const ref_params* p, const cf* iq, size_t count, uint16_t* symbols, size_t cap, ref_metrics* m)
{
    ref_ws w(p);
    const long r = lora_phy::demodulate(&w.ws, iq, count, symbols, cap);
    m->cfo = w.ws.metrics.cfo; m->time_offset = w.ws.metrics.time_offset;
    m->crc_ok = w.ws.metrics.crc_ok; m->sync_word = w.ws.sync_word;
    return r;
}

size_t ref_lora_demodulate(unsigned sf, int hann, const cf* samples, size_t count, uint16_t* out_symbols,
                           unsigned osr, uint8_t* out_sync, cf* scratch, size_t scratch_len, ref_metrics* m)
{
    static lora_phy::lora_demod_workspace ws; // 115 KB: keep it off the stack
    lora_phy::lora_demod_init(&ws, sf, hann ? lora_phy::window_type::window_hann : lora_phy::window_type::window_none,
                              scratch, scratch_len);
    const size_t r = lora_phy::lora_demodulate(&ws, samples, count, out_symbols, osr, out_sync);
    if (m) { m->cfo = ws.metrics.cfo; m->time_offset = ws.metrics.time_offset; m->crc_ok = 0; m->sync_word = out_sync ? *out_sync : 0; }
    lora_phy::lora_demod_free(&ws);
    return r;
}

// ---- persistent workspaces, for the timed CPU baseline (init once, call many times) -------------------
void* ref_ws_create(const ref_params* p) { return new ref_ws(p); }
void ref_ws_destroy(void* h) { delete static_cast<ref_ws*>(h); }
long ref_ws_demodulate(void* h, const cf* iq, size_t count, uint16_t* symbols, size_t cap, ref_metrics* m)
{
    ref_ws* w = static_cast<ref_ws*>(h);
    const long r = lora_phy::demodulate(&w->ws, iq, count, symbols, cap);
    if (m) { m->cfo = w->ws.metrics.cfo; m->time_offset = w->ws.metrics.time_offset; m->crc_ok = 0; m->sync_word = w->ws.sync_word; }
    return r;
}

struct ref_legacy { lora_phy::lora_demod_workspace ws; std::vector<cf> scratch; std::vector<cf> down, dechirped; unsigned sf; };
void* ref_legacy_create(unsigned sf, int hann, size_t max_samples, unsigned bw_hz)
{
    ref_legacy* l = new ref_legacy();
    l->sf = sf;
    l->scratch.resize(max_samples);
    l->dechirped.resize(max_samples);
    const size_t N = size_t(1) << sf;
    l->down.resize(N);
    float phase = 0.0f; // tests/performance_test.cpp:92-98
    genChirp(l->down.data(), static_cast<int>(N), 1, static_cast<int>(N), 0.0f, true, 1.0f, phase,
             lora_phy::bw_scale(bw_of(bw_hz)));
    lora_phy::lora_demod_init(&l->ws, sf, hann ? lora_phy::window_type::window_hann : lora_phy::window_type::window_none,
                              l->scratch.data(), l->scratch.size());
    return l;
}
void ref_legacy_destroy(void* h) { ref_legacy* l = static_cast<ref_legacy*>(h); lora_phy::lora_demod_free(&l->ws); delete l; }
size_t ref_legacy_demodulate(void* h, const cf* samples, size_t count, uint16_t* out, unsigned osr, uint8_t* out_sync, ref_metrics* m)
{
    ref_legacy* l = static_cast<ref_legacy*>(h);
    const size_t r = lora_phy::lora_demodulate(&l->ws, samples, count, out, osr, out_sync);
    if (m) { m->cfo = l->ws.metrics.cfo; m->time_offset = l->ws.metrics.time_offset; m->crc_ok = 0; m->sync_word = out_sync ? *out_sync : 0; }
    return r;
}
// Caller-side dechirp exactly as tests/performance_test.cpp:109-114 / tests/e2e_chain_test.cpp:88-93 do it,
// followed by lora_demodulate (osr = 1).
size_t ref_legacy_dechirp_demodulate(void* h, const cf* samples, size_t count, uint16_t* out, uint8_t* out_sync, ref_metrics* m)
{
    ref_legacy* l = static_cast<ref_legacy*>(h);
    const size_t N = size_t(1) << l->sf;
    for (size_t s = 0; s < count / N; ++s)
        for (size_t i = 0; i < N; ++i) l->dechirped[s * N + i] = samples[s * N + i] * l->down[i];
    return ref_legacy_demodulate(h, l->dechirped.data(), count, out, 1, out_sync, m);
}


// The timed CPU baseline of bench.py: the pipeline of tests/performance_test.cpp:92-116 (down-chirp and workspace built
// once, OUTSIDE the timed loop; per packet: caller dechirp + lora_demodulate) over `packets` packets of `count` samples
// each, shared out in contiguous ranges over `threads` host threads, one workspace per thread.  Writes the data symbols
// (symbol_stride uint16 per packet) and the sync byte per packet; returns the seconds the timed loops took (threads
// started after every workspace exists, steady_clock around start..join), or -1 on a bad argument.
double ref_legacy_batch(unsigned sf, unsigned bw_hz, const cf* iq, size_t packets, size_t count, unsigned threads,
                        uint16_t* out_symbols, size_t symbol_stride, uint8_t* out_sync, size_t* out_counts)
{
    if (!iq || !out_symbols || threads == 0 || sf < 2 || sf > 12) return -1.0;
    if (threads > packets) threads = packets ? unsigned(packets) : 1u;
    std::vector<void*> ws(threads);
    for (unsigned t = 0; t < threads; ++t) ws[t] = ref_legacy_create(sf, 0, count, bw_hz);
    const size_t total = count >> sf;
    auto body = [&](unsigned t) {
        const size_t a = packets * t / threads, b = packets * (t + 1) / threads;
        std::vector<uint16_t> sym(total + 2);
        for (size_t p = a; p < b; ++p) {
            uint8_t sync = 0;
            const size_t k = ref_legacy_dechirp_demodulate(ws[t], iq + p * count, count, sym.data(), &sync, nullptr);
            std::memcpy(out_symbols + p * symbol_stride, sym.data(), sizeof(uint16_t) * (k < symbol_stride ? k : symbol_stride));
            if (out_sync) out_sync[p] = sync;
            if (out_counts) out_counts[p] = k;
        }
    };
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < threads; ++t) pool.emplace_back(body, t);
    body(0);
    for (auto& th : pool) th.join();
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    for (unsigned t = 0; t < threads; ++t) ref_legacy_destroy(ws[t]);
    return dt;
}

// The same for the workspace API (phy.cpp demodulate(), which dechirps internally).
double ref_ws_batch(const ref_params* prm, const cf* iq, size_t packets, size_t count, unsigned threads,
                    uint16_t* out_symbols, size_t symbol_stride)
{
    if (!iq || !out_symbols || threads == 0) return -1.0;
    if (threads > packets) threads = packets ? unsigned(packets) : 1u;
    std::vector<void*> ws(threads);
    for (unsigned t = 0; t < threads; ++t) ws[t] = ref_ws_create(prm);
    auto body = [&](unsigned t) {
        const size_t a = packets * t / threads, b = packets * (t + 1) / threads;
        for (size_t p = a; p < b; ++p) ref_ws_demodulate(ws[t], iq + p * count, count, out_symbols + p * symbol_stride, symbol_stride, nullptr);
    };
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < threads; ++t) pool.emplace_back(body, t);
    body(0);
    for (auto& th : pool) th.join();
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    for (unsigned t = 0; t < threads; ++t) ref_ws_destroy(ws[t]);
    return dt;
}

// lora_modulate over a batch (CPU-side synthesis for bench.py --impl reference; not timed).
void ref_modulate_batch(const uint16_t* symbols, size_t packets, size_t nsym, unsigned sf, unsigned bw_hz, float amplitude,
                        uint8_t sync, cf* out, unsigned threads)
{
    const size_t count = (nsym + 2) << sf;
    if (threads == 0) threads = 1;
    auto body = [&](unsigned t) {
        for (size_t p = packets * t / threads; p < packets * (t + 1) / threads; ++p)
            lora_phy::lora_modulate(symbols + p * nsym, nsym, out + p * count, sf, 1, bw_of(bw_hz), amplitude, sync);
    };
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < threads; ++t) pool.emplace_back(body, t);
    body(0);
    for (auto& th : pool) th.join();
}

} // extern "C"
