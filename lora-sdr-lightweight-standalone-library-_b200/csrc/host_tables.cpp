// host_tables.cpp -- see host_tables.h.  Compiled with -ffp-contract=off: the reference is built for
// baseline x86-64 where none of these float expressions can fuse.
#include "host_tables.h"

#include <cmath>

namespace lora_b200 {

std::vector<std::complex<float>> build_twiddles(int n)
{
    std::vector<std::complex<float>> tw(static_cast<size_t>(n));
    // -2*pi/n: for n a power of two the division is exact, so float(-2*pi_double/n) == -2*float(pi)/n
    const float phinc = static_cast<float>(-2.0 * std::acos(-1.0) / n);
    for (int i = 0; i < n; ++i) tw[i] = std::exp(std::complex<float>(0.0f, static_cast<float>(i) * phinc));
    return tw;
}

int host_gen_chirp(std::complex<float>* out, int n, int osr, int nn, float f0, bool down, float ampl,
                   float& phase, float bw_scale)
{
    const double pi = 3.14159265358979323846;
    const float lo = static_cast<float>(-pi * bw_scale / osr);
    const float hi = static_cast<float>(pi * bw_scale / osr);
    const float inc = static_cast<float>((2 * pi * bw_scale) / (n * osr * osr));
    const float span = hi - lo;
    float f = lo + f0;
    float acc = phase;
    for (int i = 0; i < nn; ++i) {
        f += inc;
        if (f > hi) f -= span;
        acc = down ? acc - f : acc + f;
        out[i] = std::complex<float>(ampl * std::cos(acc), ampl * std::sin(acc));
    }
    phase = static_cast<float>(static_cast<double>(acc) - std::floor(static_cast<double>(acc) / (2 * pi)) * 2 * pi);
    return nn;
}

std::vector<std::complex<float>> build_downchirp(int n, float bw_scale)
{
    std::vector<std::complex<float>> d(static_cast<size_t>(n));
    float phase = 0.0f;
    host_gen_chirp(d.data(), n, 1, n, 0.0f, true, 1.0f, phase, bw_scale);
    return d;
}

std::vector<float> build_downchirp_phase(int n, float bw_scale)
{
    // the same recurrence as host_gen_chirp(n, 1, n, 0, down, 1, phase = 0), keeping the float32 phase itself
    const double pi = 3.14159265358979323846;
    const float lo = static_cast<float>(-pi * bw_scale);
    const float hi = static_cast<float>(pi * bw_scale);
    const float inc = static_cast<float>((2 * pi * bw_scale) / n);
    const float span = hi - lo;
    float f = lo + 0.0f;
    float acc = 0.0f;
    std::vector<float> turns(static_cast<size_t>(n));
    for (int i = 0; i < n; ++i) {
        f += inc;
        if (f > hi) f -= span;
        acc = acc - f;
        const double t = static_cast<double>(acc) / (2 * pi);
        turns[i] = static_cast<float>((t - std::nearbyint(t)) * (2 * pi));
    }
    return turns;
}

std::vector<float> build_hann(int n)
{
    std::vector<float> w(static_cast<size_t>(n));
    const float pi_f = static_cast<float>(3.14159265358979323846);
    for (int i = 0; i < n; ++i)
        w[i] = 0.5f - 0.5f * std::cos(2.0f * pi_f * static_cast<float>(i) / (static_cast<float>(n) - 1.0f));
    return w;
}

float detector_power_scale(int n) { return static_cast<float>(20 * std::log10(static_cast<double>(n))); }

void build_whitening(uint8_t* seq_std, uint8_t* seq_p54, int len)
{
    const uint64_t seeds[2][2] = {{0x6572D100E85C2EFFull, 0xE85C2EFFFFFFFFFFull},
                                  {0x05121100F8ECFEEFull, 0xF8ECFEEFEFEFEFEFull}};
    uint8_t* dst[2] = {seq_std, seq_p54};
    for (int v = 0; v < 2; ++v) {
        uint64_t r[2] = {seeds[v][0], seeds[v][1]};
        for (int i = 0; i < len; ++i) {
            uint64_t& x = r[i & 1];
            dst[v][i] = static_cast<uint8_t>(x & 0xff);
            x = (x >> 8) | (((x >> 32) ^ (x >> 24) ^ (x >> 16) ^ x) << 56);
        }
    }
}

} // namespace lora_b200
