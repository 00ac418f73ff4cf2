// host_tables.h -- init-time tables, built on the host the way the reference's init builds them.
#pragma once
#include <complex>
#include <cstddef>
#include <cstdint>
#include <vector>

namespace lora_b200 {

// exp(-2*pi*j*i/N) through std::exp(std::complex<float>) with a float phase increment, as
// include/lora_phy/kissfft.hh:24-29 does (so the table holds the host libm's bits, like the reference's).
std::vector<std::complex<float>> build_twiddles(int n);
// The table demodulate() regenerates every symbol: genChirp(N, 1, N, 0, down, 1, phase=0, bw_scale)
// (src/phy/phy.cpp:202-204 -> include/lora_phy/ChirpGenerator.hpp:24-50).
std::vector<std::complex<float>> build_downchirp(int n, float bw_scale);
// Phase of every entry of that table in radians, reduced to [-pi, pi] in double: the float32 accumulator the
// recurrence holds when it emits the entry.  Used by the fast demod flavour, which applies dechirp and CFO
// rotation as one phase.
std::vector<float> build_downchirp_phase(int n, float bw_scale);
// 0.5 - 0.5*cos(2*pi*i/(N-1)) in float (src/phy/phy.cpp:36-43).
std::vector<float> build_hann(int n);
// float(20*log10(N)), include/lora_phy/LoRaDetector.hpp:29.
float detector_power_scale(int n);
// One chirp with the reference's float32 phase-accumulator recurrence; returns samples written.
int host_gen_chirp(std::complex<float>* out, int n, int osr, int nn, float f0, bool down, float ampl,
                   float& phase, float bw_scale);
// First `len` bytes of the two whitening sequences (index 0: RDD != 1, index 1: RDD == 1), produced by
// stepping the interleaved LFSRs of include/lora_phy/LoRaCodes.hpp:176-189 on a zero buffer.
void build_whitening(uint8_t* seq_std, uint8_t* seq_p54, int len);

} // namespace lora_b200
