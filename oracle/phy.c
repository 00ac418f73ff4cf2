/* phy.c -- the reference's lightweight PHY entry points (src/phy/), restated in plain C.
 * TEST INFRASTRUCTURE ONLY (see lora_oracle.h).  Same build flags as dsp.c. */
#include "lora_oracle.h"

#include <math.h>
#include <string.h>

#define ORC_MAX_N 4096
static const float PI_F = (float)M_PI; /* float(M_PI) as the reference writes it */

static orc_c32 cmul(orc_c32 a, orc_c32 b)
{
    orc_c32 r;
    r.re = a.re * b.re - a.im * b.im;
    r.im = a.re * b.im + a.im * b.re;
    return r;
}

/* The reference builds its twiddle table once, in init (phy.cpp:30, LoRaDemod.cpp:26).  Cache it
 * per thread and per N so a timed loop over packets does not pay cexpf N times per call. */
static const orc_c32* twiddles_for(size_t N)
{
    static _Thread_local orc_c32 tw[ORC_MAX_N];
    static _Thread_local size_t have = 0;
    if (have != N) { orc_fft_twiddles(tw, (int)N, 0); have = N; }
    return tw;
}

static float bw_scale_of(unsigned bw_hz) { return (float)bw_hz / 125000.0f; } /* phy.hpp:43-49 */

static void make_window(float* w, size_t N, int hann) /* phy.cpp:36-47, LoRaDemod.cpp:17-25 */
{
    for (size_t i = 0; i < N; ++i)
        w[i] = hann ? 0.5f - 0.5f * cosf(2.0f * PI_F * (float)i / ((float)N - 1.0f)) : 1.0f;
}

size_t orc_lora_encode(const uint8_t* bytes, size_t n, uint16_t* symbols) /* LoRaEncoder.cpp:6-18 */
{
    size_t k = 0;
    for (size_t i = 0; i < n; ++i) {
        symbols[k++] = orc_enc_h84((uint8_t)(bytes[i] >> 4));
        symbols[k++] = orc_enc_h84((uint8_t)(bytes[i] & 0x0f));
    }
    return k;
}

size_t orc_lora_decode(const uint16_t* symbols, size_t n, uint8_t* bytes) /* LoRaDecoder.cpp:6-19 */
{
    size_t k = 0;
    for (size_t i = 0; i + 1 < n; i += 2) {
        int e = 0, b = 0;
        const uint8_t hi = orc_dec_h84((uint8_t)symbols[i], &e, &b) & 0x0f;
        const uint8_t lo = orc_dec_h84((uint8_t)symbols[i + 1], &e, &b) & 0x0f;
        bytes[k++] = (uint8_t)((hi << 4) | lo);
    }
    return k;
}

long orc_decode(const uint16_t* symbols, size_t n, uint8_t* payload, size_t cap, int* crc_ok)
{ /* phy.cpp:241-256.  Like the reference, bytes are written before the capacity check. */
    const size_t produced = orc_lora_decode(symbols, n, payload);
    if (produced > cap) return -1;
    if (produced >= 4) {
        const uint16_t provided = (uint16_t)(payload[produced - 2] | (payload[produced - 1] << 8));
        *crc_ok = provided == orc_data_checksum(payload + 2, (int)(produced - 4));
    } else {
        *crc_ok = 0;
    }
    return (long)produced;
}

size_t orc_lora_modulate(const uint16_t* symbols, size_t n, orc_c32* out, unsigned sf, unsigned osr,
                         unsigned bw_hz, float amplitude, uint8_t sync) /* LoRaMod.cpp:8-43 */
{
    const size_t N = (size_t)1 << sf, step = N * osr;
    const float scale = bw_scale_of(bw_hz);
    float phase = 0.0f;
    amplitude = fmaxf(-1.0f, fminf(1.0f, amplitude)); /* :18 */
    const unsigned shift = sf > 4 ? sf - 4 : 0;
    const uint16_t sw[2] = {(uint16_t)((sync >> 4) << shift), (uint16_t)((sync & 0x0f) << shift)};
    for (size_t s = 0; s < n + 2; ++s) {
        const uint16_t v = s < 2 ? sw[s] : symbols[s - 2];
        /* :24,29,36 -- ((2*pi_f * v) * scale) / (N * osr), all in float */
        const float f0 = (2.0f * PI_F * (float)v * scale) / ((float)N * (float)osr);
        orc_gen_chirp(out + s * step, (int)N, (int)osr, (int)step, f0, 0, amplitude, &phase, scale);
    }
    return (n + 2) * step;
}

/* Shared by estimate_offsets (phy.cpp:78-145) and the copy inlined in lora_demodulate
 * (LoRaDemod.cpp:79-135); `legacy_tie` selects the extra lowest-index tie rule of the latter. */
static void estimate_core(const orc_c32* tw, size_t N, unsigned osr, const float* window,
                          const orc_c32* samples, size_t symbols, int legacy_tie, float* cfo,
                          float* time_offset)
{
    static _Thread_local orc_c32 in[ORC_MAX_N], spec[ORC_MAX_N];
    const size_t step = N * osr;
    float sum_index = 0.0f, phase_diff = 0.0f, prev_phase = 0.0f;
    int have_prev = 0;
    unsigned sum_t = 0;
    for (size_t s = 0; s < symbols; ++s) {
        const orc_c32* sym = samples + s * step;
        float best_p = -1e30f, best_f = 0.0f;
        size_t best_idx = 0;
        unsigned best_t = 0;
        orc_c32 best_bin = {0.0f, 0.0f};
        for (unsigned t = 0; t < osr; ++t) {
            for (size_t i = 0; i < N; ++i) {
                orc_c32 v = sym[t + i * osr];
                if (window) { v.re *= window[i]; v.im *= window[i]; }
                in[i] = v;
            }
            float p, pav, fi;
            const size_t idx = orc_detect(tw, N, in, spec, &p, &pav, &fi, NULL);
            if (p > best_p || (legacy_tie && p == best_p && idx < best_idx)) {
                best_p = p; best_idx = idx; best_f = fi; best_t = t; best_bin = spec[idx];
            }
        }
        sum_t += best_t;
        sum_index += (float)best_idx + best_f;
        const float phase = atan2f(best_bin.im, best_bin.re); /* std::arg */
        if (have_prev) {
            float d = phase - prev_phase;
            while (d > PI_F) d -= 2.0f * PI_F;
            while (d < -PI_F) d += 2.0f * PI_F;
            phase_diff += d;
        }
        prev_phase = phase;
        have_prev = 1;
    }
    const float avg_index = sum_index / (float)symbols;
    const float cfo_coarse = avg_index / (float)N;
    float cfo_fine = 0.0f;
    if (symbols > 1)
        cfo_fine = (phase_diff / (float)(symbols - 1)) / (2.0f * PI_F * (float)N);
    *cfo = cfo_coarse + cfo_fine;
    const float frac = avg_index - floorf(avg_index + 0.5f);
    const float avg_t = (float)sum_t / (float)symbols;
    *time_offset = avg_t - frac * (float)N * (float)osr;
}

void orc_estimate_offsets(const orc_params* p, const orc_c32* samples, size_t count, float* cfo,
                          float* time_offset)
{ /* phy.cpp:78-145; leaves *cfo / *time_offset untouched on the early returns */
    static _Thread_local float win[ORC_MAX_N];
    if (!samples || count == 0) return;
    const unsigned osr = p->osr ? p->osr : 1u;
    const size_t N = (size_t)1 << p->sf;
    const size_t symbols = count / (N * osr);
    if (symbols == 0) return;
    const orc_c32* tw = twiddles_for(N);
    const int use_win = p->hann && p->has_window; /* window_kind != none && ws->window */
    if (use_win) make_window(win, N, 1);
    estimate_core(tw, N, osr, use_win ? win : NULL, samples, symbols, 0, cfo, time_offset);
}

void orc_compensate_offsets(const orc_params* p, float cfo, float time_offset, orc_c32* samples,
                            size_t count) /* phy.cpp:147-176 */
{
    if (!samples || count == 0) return;
    const unsigned osr = p->osr ? p->osr : 1u;
    const size_t N = (size_t)1 << p->sf;
    const float rate = -2.0f * PI_F * cfo / ((float)N * (float)osr);
    for (size_t n = 0; n < count; ++n) {
        const float ph = rate * (float)n;
        const orc_c32 rot = {cosf(ph), sinf(ph)};
        samples[n] = cmul(samples[n], rot);
    }
    const int offset = (int)roundf(time_offset);
    const orc_c32 zero = {0.0f, 0.0f};
    if (offset > 0 && (size_t)offset < count) {
        for (size_t n = count; n-- > (size_t)offset;) samples[n] = samples[n - (size_t)offset];
        for (size_t n = 0; n < (size_t)offset; ++n) samples[n] = zero;
    } else if (offset < 0 && (size_t)(-offset) < count) {
        const size_t off = (size_t)(-offset);
        for (size_t n = 0; n + off < count; ++n) samples[n] = samples[n + off];
        for (size_t n = count - off; n < count; ++n) samples[n] = zero;
    }
}

/* The per-symbol loop common to demodulate (phy.cpp:200-234) and lora_demodulate
 * (LoRaDemod.cpp:141-175): window base with the t_off shift, CFO rotation, optional window, detect. */
static void symbol_pass(const orc_c32* tw, size_t N, unsigned osr, const float* window,
                        const orc_c32* down, const orc_c32* iq, size_t count, size_t total,
                        float cfo, float time_offset, uint16_t* raw_idx, float* peaks)
{
    static _Thread_local orc_c32 in[ORC_MAX_N], spec[ORC_MAX_N];
    const size_t step = N * osr;
    const int t_off = (int)roundf(time_offset);
    const float rate = -2.0f * PI_F * cfo / (float)N;
    for (size_t s = 0; s < total; ++s) {
        size_t base = s * step;
        if (t_off > 0) {
            if (base + (size_t)t_off + step <= count) base += (size_t)t_off;
        } else if (t_off < 0) {
            const size_t off = (size_t)(-t_off);
            if (off <= base) base -= off;
        }
        const orc_c32* sym = iq + base;
        const float start = rate * ((float)(s * N) + (float)t_off / (float)osr);
        for (size_t i = 0; i < N; ++i) {
            const float ph = start + rate * (float)i;
            const orc_c32 rot = {cosf(ph), sinf(ph)};
            orc_c32 v = sym[i * osr];
            if (down) v = cmul(v, down[i]); /* (iq * down) * rot, left to right */
            v = cmul(v, rot);
            if (window) { v.re *= window[i]; v.im *= window[i]; }
            in[i] = v;
        }
        float p, pav, fi, mx;
        raw_idx[s] = (uint16_t)orc_detect(tw, N, in, spec, &p, &pav, &fi, &mx);
        if (peaks) peaks[s] = mx;
    }
}

long orc_demodulate(const orc_params* p, const orc_c32* iq, size_t count, uint16_t* symbols,
                    size_t cap, orc_metrics* m, float* peaks) /* phy.cpp:178-239 */
{
    static _Thread_local orc_c32 down[ORC_MAX_N];
    static _Thread_local float win[ORC_MAX_N];
    static _Thread_local uint16_t raw[1 << 20];
    if (!iq || !symbols) return -1;
    const unsigned osr = p->osr ? p->osr : 1u;
    const size_t N = (size_t)1 << p->sf, step = N * osr;
    const orc_c32* tw = twiddles_for(N);
    if (count % step != 0) return -1;
    const size_t total = count / step;
    if (total < 2) return -1;
    if (total - 2 > cap) return -1;
    if (total > sizeof(raw) / sizeof(raw[0])) return -1; /* oracle buffer limit, not a reference rule */

    const int use_win = p->hann && p->has_window;
    if (use_win) make_window(win, N, 1);
    const float* w = use_win ? win : NULL;

    /* :192-193 -- the estimate sees the RAW (not dechirped) first two symbols */
    const size_t est = count < 2 * step ? count : 2 * step;
    estimate_core(tw, N, osr, w, iq, est / step, 0, &m->cfo, &m->time_offset);

    float tmp = 0.0f; /* :202-204 -- the same table every symbol */
    orc_gen_chirp(down, (int)N, 1, (int)N, 0.0f, 1, 1.0f, &tmp, bw_scale_of(p->bw_hz));

    symbol_pass(tw, N, osr, w, down, iq, count, total, m->cfo, m->time_offset, raw, peaks);
    for (size_t s = 2; s < total; ++s) symbols[s - 2] = raw[s];
    const unsigned shift = p->sf > 4 ? p->sf - 4 : 0;
    m->sync_word = (uint8_t)((((raw[0] >> shift) & 0x0f) << 4) | ((raw[1] >> shift) & 0x0f));
    return (long)(total - 2);
}

size_t orc_lora_demodulate(unsigned sf, int hann, const orc_c32* samples, size_t count,
                           uint16_t* out_symbols, unsigned osr, uint8_t* out_sync,
                           orc_c32* scratch, size_t scratch_len, orc_metrics* m, float* peaks)
{ /* LoRaDemod.cpp:49-195 */
    static _Thread_local float win[ORC_MAX_N];
    static _Thread_local uint16_t raw[1 << 20];
    const size_t N = (size_t)1 << sf, step = N * osr;
    const size_t total = count / step;
    const int have_sync = total >= 2;
    const orc_c32* tw = twiddles_for(N);
    if (total > sizeof(raw) / sizeof(raw[0])) return 0;

    const orc_c32* norm = samples; /* :59-77 */
    float max_amp = 0.0f;
    for (size_t i = 0; i < count; ++i) {
        const float a = fabsf(samples[i].re), b = fabsf(samples[i].im);
        const float mx = a < b ? b : a; /* std::max(r, im) */
        if (mx > max_amp) max_amp = mx;
    }
    if (max_amp > 1.0f) {
        if (!scratch || scratch_len < count) return 0;
        const float scale = 1.0f / max_amp;
        for (size_t i = 0; i < count; ++i) {
            scratch[i].re = samples[i].re * scale;
            scratch[i].im = samples[i].im * scale;
        }
        norm = scratch;
    }

    if (hann) make_window(win, N, 1);
    const float* w = hann ? win : NULL; /* an all-ones window multiply is skipped by the reference too */

    const size_t est = total < 2 ? total : 2;
    orc_metrics local;
    if (!m) m = &local;
    if (est == 0) { /* 0/0 in the reference: cfo and time_offset become NaN; nothing else to do */
        m->cfo = NAN; m->time_offset = NAN;
        if (out_sync) *out_sync = 0;
        return 0;
    }
    estimate_core(tw, N, osr, w, norm, est, 1, &m->cfo, &m->time_offset);
    symbol_pass(tw, N, osr, w, NULL, norm, count, total, m->cfo, m->time_offset, raw, peaks);

    size_t out = 0;
    for (size_t s = have_sync ? 2 : 0; s < total; ++s) out_symbols[out++] = raw[s];
    uint8_t sync = 0;
    if (have_sync) {
        const unsigned shift = sf > 4 ? sf - 4 : 0;
        sync = (uint8_t)((((uint8_t)(raw[0] >> shift) & 0x0f) << 4) | ((uint8_t)(raw[1] >> shift) & 0x0f));
    }
    m->sync_word = sync;
    if (out_sync) *out_sync = sync;
    return have_sync ? out : total;
}
