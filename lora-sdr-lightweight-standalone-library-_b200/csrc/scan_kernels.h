// scan_kernels.h -- the light normalisation scan (LoRaDemod.cpp:59-77) that runs beside the demod kernel.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "packed_math.cuh"

namespace lora_b200 {

struct ScanArgs {
    const float2* iq;
    size_t packet_stride, count;
    int n_packets;
    int N;             // samples per symbol (down-chirp table length)
    const c64* down;   // non-null: examine iq * down[n mod N] (fused caller-dechirp, performance_test.cpp:109-114)
    float one;         // run-time 1.0f (packed_math.cuh)
    float4* est;       // out, per packet: z = scale, w = max_amp (x, y zeroed)
};

bool scan_light_supported(const ScanArgs& a);
cudaError_t launch_scan_light(const ScanArgs& a, int sm_count, cudaStream_t st);

} // namespace lora_b200
