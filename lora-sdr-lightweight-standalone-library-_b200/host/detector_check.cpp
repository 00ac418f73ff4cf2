// detector_check.cpp -- the upstream detector test (LoRa-SDR/TestDetector.cpp:9-35) written against this build's
// caller-side headers: kissfft<float>::init / transform and LoRaDetector<float>::feed / detect, which forward to the
// device (include/lora_phy/kissfft.hh, LoRaDetector.hpp).  For N = 2^sf (default 1024, as upstream) every symbol value
// must come back as the detected bin with power above -10 dB; then FFT followed by the inverse FFT must return N times
// the input to float accuracy.  Prints "detector ok" and exits 0, or says what failed.  With --dump=FILE it also
// writes, for a few symbols, the dechirped window, the spectrum and the four detector outputs as raw little-endian
// floats so tests/test_gpu_fft.py can compare the bits with the oracle.
#include <lora_phy/ChirpGenerator.hpp>
#include <lora_phy/LoRaDetector.hpp>
#include <lora_phy/kissfft.hh>

#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

int main(int argc, char** argv)
{
    unsigned sf = 10;
    const char* dump = nullptr;
    unsigned stride = 1;
    for (int i = 1; i < argc; ++i) {
        if (!std::strncmp(argv[i], "--sf=", 5)) sf = unsigned(std::atoi(argv[i] + 5));
        else if (!std::strncmp(argv[i], "--dump=", 7)) dump = argv[i] + 7;
        else if (!std::strncmp(argv[i], "--stride=", 9)) stride = unsigned(std::atoi(argv[i] + 9));
    }
    const size_t N = size_t(1) << sf;
    typedef std::complex<float> cf;
    std::vector<cf> down(N), chirp(N), in(N), out(N), back(N);
    float phase = 0.0f;
    genChirp(down.data(), int(N), 1, int(N), 0.0f, true, 1.0f, phase);

    kissfft_plan<float> plan, plan_inv;
    kissfft<float>::init(plan, int(N), false);
    kissfft<float>::init(plan_inv, int(N), true);
    kissfft<float> fft(plan), ifft(plan_inv);
    LoRaDetector<float> det(N, in.data(), out.data(), fft);
    FILE* f = dump ? std::fopen(dump, "wb") : nullptr;

    for (size_t sym = 0; sym < N; sym += stride ? stride : 1) {
        phase = float(M_PI / 4); // some phase offset, as upstream
        genChirp(chirp.data(), int(N), 1, int(N), float(2 * M_PI * sym) / N, false, 1.0f, phase);
        for (size_t i = 0; i < N; ++i) det.feed(i, down[i] * chirp[i]);
        float power = 0, power_avg = 0, f_index = 0;
        const size_t index = det.detect(power, power_avg, f_index);
        if (index != sym || !(power > -10.0f)) {
            std::fprintf(stderr, "symbol %zu detected as %zu, power %g (no CUDA device?)\n", sym, index, power);
            return 1;
        }
        if (f && (sym % 97 == 0 || sym + 1 == N)) {
            const float rec[4] = {float(index), power, power_avg, f_index};
            std::fwrite(in.data(), sizeof(cf), N, f);
            std::fwrite(out.data(), sizeof(cf), N, f);
            std::fwrite(rec, sizeof(float), 4, f);
        }
    }
    // forward then inverse: N * input
    fft.transform(in.data(), out.data());
    ifft.transform(out.data(), back.data());
    if (fft.last_error() || ifft.last_error()) { std::fprintf(stderr, "transform failed\n"); return 1; }
    double worst = 0.0;
    for (size_t i = 0; i < N; ++i) worst = std::fmax(worst, double(std::abs(back[i] / float(N) - in[i])));
    if (!(worst < 1e-5)) { std::fprintf(stderr, "inverse(forward(x)) / N differs from x by %g\n", worst); return 1; }
    if (f) std::fclose(f);
    std::printf("detector ok: %zu symbols at N = %zu, round-trip error %.2g\n", (N + stride - 1) / stride, N, worst);
    return 0;
}
