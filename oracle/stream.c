/* stream.c -- the streaming frame synchroniser and the full frame generator of the upstream Pothos blocks, restated
 * in plain C.  TEST INFRASTRUCTURE ONLY (see lora_oracle.h).
 *
 * The two blocks cannot be built here (they need the Pothos runtime); their arithmetic is entirely in headers that
 * are present (LoRa-SDR/ChirpGenerator.hpp, LoRaDetector.hpp, kissfft.hh -- the same expressions as the lightweight
 * copies restated in dsp.c).  What is restated below is the orchestration:
 *   orc_frame_modulate   LoRa-SDR/LoRaMod.cpp:109-238   (10 up-chirps, 2 sync chirps, 2 1/4 down-chirps, data, padding)
 *   orc_stream_demod     LoRa-SDR/LoRaDemod.cpp:98-117 (tables), :145-327 (work() state machine)
 * A Pothos scheduler calls work() repeatedly while at least 2N samples are buffered (:148) and consumes `total`
 * samples per call (:321); here that loop runs over one finite buffer.  No committed expected output exists for
 * these blocks: parity of this restatement is pinned by the loopback property the Pothos TestLoopback states
 * (generated frames come back as the symbols that went in) -- "parity unpinned" against reference outputs. */
#include "lora_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define FINE_STEPS 128 /* LoRaDemod.cpp:70 */

static orc_c32 cmulf(orc_c32 a, orc_c32 b)
{
    orc_c32 r;
    r.re = a.re * b.re - a.im * b.im;
    r.im = a.re * b.im + a.im * b.re;
    return r;
}

/* LoRaMod.cpp:109-238.  symbols: n_sym payload symbols.  Returns samples written (out may be NULL to size). */
size_t orc_frame_modulate(unsigned sf, unsigned ovs, const uint16_t* symbols, size_t n_sym, unsigned char sync,
                          unsigned padding, float ampl, orc_c32* out)
{
    const int N = 1 << sf;
    const size_t NN = (size_t)N * ovs;
    /* :217-228: STATE_PADSYMBOLS increments its counter BEFORE comparing it with _padding, so a padding of 0 still
     * produces one zero symbol (pinned against the compiled block, tests/test_oracle_vs_pothos.py) */
    const size_t n_pad = padding ? padding : 1;
    const size_t total = (10 + 2 + 2) * NN + NN / 4 + n_sym * NN + n_pad * NN;
    if (!out) return total;
    float phase = 0.0f; /* :131 _phaseAccum = 0 at every packet */
    size_t pos = 0;
    for (int k = 0; k < 10; ++k) /* :138-143 */
        pos += (size_t)orc_gen_chirp(out + pos, N, (int)ovs, (int)NN, 0.0f, 0, ampl, &phase, 1.0f);
    {   /* :150-156: freq = (2*M_PI*sw)/NN evaluated in double, stored in a float */
        const int sw0 = (sync >> 4) * 8;
        const float f0 = (float)((2 * M_PI * sw0) / (double)NN);
        pos += (size_t)orc_gen_chirp(out + pos, N, (int)ovs, (int)NN, f0, 0, ampl, &phase, 1.0f);
        const int sw1 = (sync & 0xf) * 8; /* :162-167 */
        const float f1 = (float)((2 * M_PI * sw1) / (double)NN);
        pos += (size_t)orc_gen_chirp(out + pos, N, (int)ovs, (int)NN, f1, 0, ampl, &phase, 1.0f);
    }
    pos += (size_t)orc_gen_chirp(out + pos, N, (int)ovs, (int)NN, 0.0f, 1, ampl, &phase, 1.0f);       /* :174 */
    pos += (size_t)orc_gen_chirp(out + pos, N, (int)ovs, (int)NN, 0.0f, 1, ampl, &phase, 1.0f);       /* :184 */
    pos += (size_t)orc_gen_chirp(out + pos, N, (int)ovs, (int)(NN / 4), 0.0f, 1, ampl, &phase, 1.0f); /* :194 */
    for (size_t s = 0; s < n_sym; ++s) { /* :205-208 */
        const int sym = symbols[s];
        const float f = (float)((2 * M_PI * sym) / (double)NN);
        pos += (size_t)orc_gen_chirp(out + pos, N, (int)ovs, (int)NN, f, 0, ampl, &phase, 1.0f);
    }
    for (size_t i = 0; i < n_pad * NN; ++i) { out[pos].re = 0.0f; out[pos].im = 0.0f; ++pos; } /* :224-226 */
    return pos;
}

/* Tables of LoRaDemod.cpp:98-117: phase accumulated in double from a float frequency, std::polar in double, stored
 * as complex<float>.  up = conj(entry), down = entry; fine[i] = polar(1, (i+1) * 2*pi/(N*128)) accumulated. */
void orc_stream_tables(unsigned sf, orc_c32* up, orc_c32* down, orc_c32* fine)
{
    const size_t N = (size_t)1 << sf;
    float phase = (float)(-M_PI);
    double accum = 0.0;
    for (size_t i = 0; i < N; ++i) {
        accum += phase;
        const double c = cos(accum), s = sin(accum);
        up[i].re = (float)c;   up[i].im = (float)(-s);
        down[i].re = (float)c; down[i].im = (float)s;
        phase += (float)((2 * M_PI) / (double)N); /* float += double: rounded to float each step */
    }
    accum = 0.0;
    phase = (float)(2.0 * M_PI / (double)(N * FINE_STEPS));
    for (size_t i = 0; i < N * FINE_STEPS; ++i) {
        accum += phase;
        fine[i].re = (float)cos(accum);
        fine[i].im = (float)sin(accum);
    }
}

/* One dechirped window through the detector.  `ft` / the fine-tune walk follow :157-165 exactly: the index is an int
 * that is decremented by a float product (int -> float, subtract, truncate back) once per sample. */
static size_t window_detect(const orc_c32* tw, size_t N, const orc_c32* in, const orc_c32* chirp, const orc_c32* fine,
                            int* ft, float fine_err, orc_c32* buf, orc_c32* spec, float* power, float* power_avg,
                            float* f_index)
{
    const int wrap = (int)(N * FINE_STEPS);
    int idx = *ft;
    for (size_t i = 0; i < N; ++i) {
        buf[i] = cmulf(cmulf(in[i], chirp[i]), fine[idx]);
        idx = (int)((float)idx - fine_err * (float)FINE_STEPS);
        if (idx < 0) idx += wrap;
        else if (idx >= wrap) idx -= wrap;
    }
    *ft = idx;
    return orc_detect(tw, N, buf, spec, power, power_avg, f_index, NULL);
}

/* LoRaDemod.cpp:145-327 over a finite buffer.  Outputs: packets of int16 symbols, written back to back into
 * out_symbols (capacity sym_cap); per packet pkt_len, pkt_start (sample index of the first data window),
 * pkt_err (the averaged down-chirp frequency error).  Returns the number of packets emitted. */
size_t orc_stream_demod(unsigned sf, unsigned char sync, float thresh, size_t mtu, const orc_c32* samples, size_t count,
                        int16_t* out_symbols, size_t sym_cap, uint32_t* pkt_len, uint64_t* pkt_start, int32_t* pkt_err,
                        size_t pkt_cap, uint64_t* consumed)
{
    return orc_stream_demod_hdr(sf, sync, thresh, mtu, 0, samples, count, out_symbols, sym_cap, pkt_len, pkt_start, pkt_err,
                                pkt_cap, consumed);
}

size_t orc_stream_demod_hdr(unsigned sf, unsigned char sync, float thresh, size_t mtu, int frame_header,
                            const orc_c32* samples, size_t count, int16_t* out_symbols, size_t sym_cap, uint32_t* pkt_len,
                            uint64_t* pkt_start, int32_t* pkt_err, size_t pkt_cap, uint64_t* consumed)
{
    enum { FRAMESYNC, DOWNCHIRP0, DOWNCHIRP1, QUARTERCHIRP, DATASYMBOLS };
    const size_t N = (size_t)1 << sf;
    orc_c32* tw = malloc(sizeof(orc_c32) * N);
    orc_c32* up = malloc(sizeof(orc_c32) * N);
    orc_c32* down = malloc(sizeof(orc_c32) * N);
    orc_c32* fine = malloc(sizeof(orc_c32) * N * FINE_STEPS);
    orc_c32* buf = malloc(sizeof(orc_c32) * N);
    orc_c32* spec = malloc(sizeof(orc_c32) * N);
    orc_fft_twiddles(tw, (int)N, 0);
    orc_stream_tables(sf, up, down, fine);

    int state = FRAMESYNC;                 /* activate(), :138-142 */
    const orc_c32* chirp = up;
    int fine_idx = 0, freq_err = 0;
    float fine_err = 0.0f;                 /* _finefreqError: see the note on prev_value */
    short prev_value = 0;                  /* _prevValue, _freqError, _finefreqError are left uninitialised by the
                                              reference's constructor (:68-75); zero is assumed here */
    size_t sym_count = 0, n_pkt = 0, sym_used = 0, pos = 0, frame_limit = mtu;
    uint16_t hdr_sym[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint64_t cur_start = 0;

    while (pos + 2 * N <= count) {         /* :148 */
        const orc_c32* in = samples + pos;
        float power = 0, power_avg = 0, f_index = 0;
        const size_t value = window_detect(tw, N, in, chirp, fine, &fine_idx, fine_err, buf, spec, &power, &power_avg, &f_index);
        const float snr = power - power_avg;
        const int squelched = snr < thresh; /* :174 */
        size_t total = 0;
        switch (state) {
        case FRAMESYNC: {                   /* :180-237 */
            const int syncd = !squelched && (prev_value + 4) / 8 == 0;
            const int match0 = (value + 4) / 8 == (unsigned)(sync >> 4);
            int match1 = 0;
            if (syncd && match0) {
                int ft = fine_idx;          /* a copy: the member is not advanced by the look-ahead (:191-203) */
                /* the look-ahead writes into the SAME power / powerAvg / fIndex variables (:204): when sync word 1 does
                 * not match, the frequency-error branch below adds the look-ahead window's fIndex, not this one's */
                const size_t value1 = window_detect(tw, N, in + N, chirp, fine, &ft, fine_err, buf, spec, &power, &power_avg, &f_index);
                match1 = (value1 + 4) / 8 == (unsigned)(sync & 0xf);
            }
            if (syncd && match0 && match1) {
                total = 2 * N;
                state = DOWNCHIRP0;
                chirp = down;
            } else if (!squelched) {
                total = N - value;
                fine_err += f_index;
            } else {
                total = N;
                fine_err = 0;
                fine_idx = 0;
            }
        } break;
        case DOWNCHIRP0: {                  /* :242-250 */
            state = DOWNCHIRP1;
            total = N;
            int error = (int)value;
            if (value > N / 2) error -= (int)N;
            freq_err = error;
        } break;
        case DOWNCHIRP1: {                  /* :255-270 */
            state = QUARTERCHIRP;
            total = N;
            chirp = up;
            int error = (int)value;
            if (value > N / 2) error -= (int)N;
            freq_err = (freq_err + error) / 2;
        } break;
        case QUARTERCHIRP: {                /* :275-284 */
            state = DATASYMBOLS;
            total = N / 4 + (size_t)(freq_err / 2); /* size_t + int: wraps modulo 2^64 exactly as the reference does */
            fine_err += (float)(freq_err / 2);
            sym_count = 0;
            frame_limit = mtu;
            cur_start = pos + total;
        } break;
        case DATASYMBOLS: {                 /* :289-310 */
            total = N;
            if (sym_used + sym_count < sym_cap) out_symbols[sym_used + sym_count] = (int16_t)value;
            if (frame_header && sym_count < 8) hdr_sym[sym_count] = (uint16_t)value;
            ++sym_count;
            if (frame_header && sym_count == 8) { /* not in the reference: frame length from the explicit header */
                const size_t want = orc_std_frame_symbols(sf, frame_header == 2, hdr_sym);
                frame_limit = want ? (want < mtu ? want : mtu) : 8;
            }
            if (sym_count >= frame_limit || squelched) {
                if (n_pkt < pkt_cap) {
                    pkt_len[n_pkt] = (uint32_t)sym_count;
                    pkt_start[n_pkt] = cur_start;
                    pkt_err[n_pkt] = freq_err;
                }
                ++n_pkt;
                sym_used += sym_count;
                if (sym_used > sym_cap) sym_used = sym_cap;
                fine_err = 0;
                state = FRAMESYNC;
            }
        } break;
        }
        prev_value = (short)value;          /* :326 */
        pos += total;                       /* :321 inPort->consume(total) */
    }
    if (consumed) *consumed = pos;
    free(tw); free(up); free(down); free(fine); free(buf); free(spec);
    return n_pkt;
}
