// lora_phy/ChirpGenerator.hpp -- genChirp<T> for CALLERS of the `lora_phy` API, B200 build.
//
// The reference's tests and its performance harness build a down-chirp table with genChirp and dechirp on the
// CPU before calling lora_demodulate (tests/e2e_chain_test.cpp:79-93, tests/performance_test.cpp:92-116).  This
// header keeps that caller-side helper available with the reference's signature and arithmetic (reference
// include/lora_phy/ChirpGenerator.hpp:24-50: float sweep variable even for other T, one wrap test per sample,
// phase wrapped once per call in double) so such callers compile unchanged.  It is not used by this library:
// the device modulator emulates the same recurrence in csrc/mod_kernels.cu, the device down-chirp table comes
// from csrc/host_tables.cpp, and a caller that wants the dechirp itself on the GPU passes
// LORA_B200_DEMOD_LEGACY_DECHIRP to lora_b200_demod_batch instead of multiplying on the host.
#pragma once

#include <cmath>
#include <complex>

template <typename Type>
int genChirp(std::complex<Type>* samps, int N, int osr, int NN, Type f0, bool down, const Type ampl, Type& phaseAccum,
             Type bw_scale = Type(1))
{
    const Type lo = -M_PI * bw_scale / osr;                       // sweep limits and slope: computed in double,
    const Type hi = M_PI * bw_scale / osr;                        // stored in Type, as the reference does
    const Type slope = (2 * M_PI * bw_scale) / (N * osr * osr);
    const Type sign = down ? Type(-1) : Type(1);
    float sweep = lo + f0;
    int produced = 0;
    for (; produced < NN; ++produced) {
        sweep += slope;
        if (sweep > hi) sweep -= (hi - lo);
        phaseAccum += sign * sweep;                               // x + (-1*f) == x - f exactly
        samps[produced] = std::polar(ampl, phaseAccum);
    }
    phaseAccum -= floor(phaseAccum / (2 * M_PI)) * 2 * M_PI;
    return produced;
}
