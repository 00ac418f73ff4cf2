/* dsp.c -- chirp generator, FFT and peak detector of the reference, restated in plain C.
 * TEST INFRASTRUCTURE ONLY (see lora_oracle.h).
 *
 * Build with -O2 -ffp-contract=off and no -ffast-math: the reference is built by its CMake with
 * plain -O2 for baseline x86-64 (CMakeLists.txt:7), so none of its float expressions are fused.
 * Transcendentals come from the host libm exactly as in the reference (std::polar/std::cos ->
 * cosf/sinf, std::exp(complex<float>) -> cexpf, std::arg -> atan2f, std::abs -> hypotf). */
#include "lora_oracle.h"

#include <complex.h>
#include <math.h>
#include <string.h>

#define ORC_MAX_N 4096
#define ORC_MAX_STAGES 32

static orc_c32 cmul(orc_c32 a, orc_c32 b) /* std::complex<float> operator*, finite operands */
{
    orc_c32 r;
    r.re = a.re * b.re - a.im * b.im;
    r.im = a.re * b.im + a.im * b.re;
    return r;
}
static orc_c32 cadd(orc_c32 a, orc_c32 b) { orc_c32 r = {a.re + b.re, a.im + b.im}; return r; }
static orc_c32 csub(orc_c32 a, orc_c32 b) { orc_c32 r = {a.re - b.re, a.im - b.im}; return r; }

/* include/lora_phy/ChirpGenerator.hpp:24-50, Type = float */
int orc_gen_chirp(orc_c32* out, int N, int osr, int NN, float f0, int down, float ampl,
                  float* phase_accum, float bw_scale)
{
    /* :27-29 -- each bound is evaluated in double (M_PI is double) and then stored as float */
    const float f_min = (float)(-M_PI * (double)bw_scale / (double)osr);
    const float f_max = (float)(M_PI * (double)bw_scale / (double)osr);
    const float f_step = (float)((2 * M_PI * (double)bw_scale) / (double)(N * osr * osr));
    float f = f_min + f0;         /* :30 */
    float ph = *phase_accum;
    int i;
    for (i = 0; i < NN; ++i) {    /* :33-46 */
        f += f_step;
        if (f > f_max) f -= (f_max - f_min);
        if (down) ph -= f; else ph += f;
        out[i].re = ampl * cosf(ph); /* std::polar(rho, theta) = (rho*cos, rho*sin) */
        out[i].im = ampl * sinf(ph);
    }
    /* :48 -- wrap once per call, in double, then narrow back into the float accumulator */
    ph = (float)((double)ph - floor((double)ph / (2 * M_PI)) * 2 * M_PI);
    *phase_accum = ph;
    return i;
}

/* include/lora_phy/kissfft.hh:24-29 */
void orc_fft_twiddles(orc_c32* tw, int nfft, int inverse)
{
    const float phinc = (float)(inverse ? 2 : -2) * acosf(-1.0f) / (float)nfft;
    for (int i = 0; i < nfft; ++i) {
        const float complex e = cexpf(CMPLXF(0.0f, (float)i * phinc));
        tw[i].re = crealf(e);
        tw[i].im = cimagf(e);
    }
}

/* include/lora_phy/kissfft.hh:80-97 */
int orc_fft_factor(int nfft, int* radix, int* remainder)
{
    int n = nfft, p = 4, stages = 0;
    do {
        while (n % p) {
            switch (p) {
            case 4: p = 2; break;
            case 2: p = 3; break;
            default: p += 2; break;
            }
            if (p * p > n) p = n;
        }
        n /= p;
        radix[stages] = p;
        remainder[stages] = n;
        ++stages;
    } while (n > 1);
    return stages;
}

/* include/lora_phy/kissfft.hh:100-185 for nfft = 2^k (radix list 4,4,...[,2]).
 *
 * The reference recurses (kf_work).  Unrolled, that recursion is: (1) copy the input into the output
 * in mixed-radix digit-reversed order (the m==1 leaf, :114-120), (2) run the butterfly passes from the
 * innermost stage (last in the radix list) outwards.  Each butterfly is a pure function of its
 * operands, so this order computes bit-identical values; tests/test_oracle_vs_ref.py checks it. */
void orc_fft(const orc_c32* tw, int nfft, int inverse, const orc_c32* src, orc_c32* dst)
{
    int radix[ORC_MAX_STAGES], rem[ORC_MAX_STAGES];
    const int stages = orc_fft_factor(nfft, radix, rem);

    /* digit-reversed load: output position sum q_s*rem[s]  <-  input index sum q_s*fstride[s] */
    for (int pos = 0; pos < nfft; ++pos) {
        int in = 0, fstride = 1, r = pos;
        for (int s = 0; s < stages; ++s) {
            const int q = r / rem[s];
            r -= q * rem[s];
            in += q * fstride;
            fstride *= radix[s];
        }
        dst[pos] = src[in];
    }

    for (int s = stages - 1; s >= 0; --s) {
        const int p = radix[s], m = rem[s];
        const int fstride = nfft / (p * m);
        for (int base = 0; base < nfft; base += p * m) {
            orc_c32* F = dst + base;
            if (p == 2) { /* kf_bfly2, :155-162 */
                for (int k = 0; k < m; ++k) {
                    const orc_c32 t = cmul(F[m + k], tw[k * fstride]);
                    F[m + k] = csub(F[k], t);
                    F[k] = cadd(F[k], t);
                }
            } else { /* kf_bfly4, :164-185 */
                for (int k = 0; k < m; ++k) {
                    const orc_c32 s0 = cmul(F[k + m], tw[k * fstride]);
                    const orc_c32 s1 = cmul(F[k + 2 * m], tw[k * fstride * 2]);
                    const orc_c32 s2 = cmul(F[k + 3 * m], tw[k * fstride * 3]);
                    const orc_c32 s5 = csub(F[k], s1);
                    F[k] = cadd(F[k], s1);
                    const orc_c32 s3 = cadd(s0, s2);
                    orc_c32 s4 = csub(s0, s2);
                    const float sg = inverse ? -1.0f : 1.0f; /* negative_if_inverse */
                    const orc_c32 rot = {s4.im * sg, -s4.re * sg};
                    s4 = rot;
                    F[k + 2 * m] = csub(F[k], s3);
                    F[k] = cadd(F[k], s3);
                    F[k + m] = cadd(s5, s4);
                    F[k + 3 * m] = csub(s5, s4);
                }
            }
        }
    }
}

/* include/lora_phy/LoRaDetector.hpp:39-74 */
size_t orc_detect(const orc_c32* tw, size_t N, const orc_c32* in, orc_c32* spectrum, float* power,
                  float* power_avg, float* f_index, float* max_mag2)
{
    orc_fft(tw, (int)N, 0, in, spectrum);
    size_t max_index = 0;
    float max_value = 0.0f;
    double total = 0.0;
    for (size_t i = 0; i < N; ++i) { /* :46-58: strict '>' from 0 => lowest index wins ties */
        const float re = spectrum[i].re, im = spectrum[i].im;
        const float mag2 = re * re + im * im;
        total += mag2;
        if (mag2 > max_value) { max_index = i; max_value = mag2; }
    }
    const float power_scale = (float)(20 * log10((double)N)); /* :29, stored in a float member */
    const float noise = sqrtf((float)(total - max_value));    /* :60 */
    const float fundamental = sqrtf(max_value);               /* :61 */
    *power_avg = 20 * log10f(noise) - power_scale;            /* :63 */
    *power = 20 * log10f(fundamental) - power_scale;          /* :64 */
    const orc_c32 l = spectrum[max_index > 0 ? max_index - 1 : N - 1];
    const orc_c32 r = spectrum[max_index < N - 1 ? max_index + 1 : 0];
    const float left = hypotf(l.re, l.im), right = hypotf(r.re, r.im); /* std::abs */
    const double demon = (2.0 * fundamental) - right - left;  /* :69 */
    if (demon == 0.0) *f_index = 0.0f;
    else *f_index = (float)(0.5 * (right - left) / demon);    /* :70-71; right-left is a float op */
    if (max_mag2) *max_mag2 = max_value;
    return max_index;
}
