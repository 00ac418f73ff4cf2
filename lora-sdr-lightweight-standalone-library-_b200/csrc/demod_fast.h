// demod_fast.h -- argument blocks and launchers of the register-blocked receive kernels (SF7..SF12).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include <vector>

#include "fft_blocked.cuh"

namespace lora_b200 {

// How windows are resolved.
enum {
    PRECISION_FAST = 0,  // fused-math kernel on every window + margin test + bit-faithful recheck of the flagged ones
    PRECISION_EXACT = 1  // bit-faithful kernel on every window
};

// w / d for w < 2^31 with one multiply-high (d = windows per packet, fixed per launch).
struct FastDiv {
    unsigned d, magic, shift;
};
FastDiv make_fastdiv(unsigned d);

struct BlockedDemodArgs {
    const float2* iq;
    size_t packet_stride, count;
    unsigned n_windows;     // n_packets * total (< 2^31; the C ABI splits larger batches)
    int total;              // windows per packet, sync included
    FastDiv div_total;
    int osr;
    int dechirp, use_scale, have_sync;
    int dechirp_absolute;   // fused caller-dechirp: table index = absolute sample position mod N
    const c64* twp;         // per-stage twiddle table (Blk::toff): the values of kissfft.hh:24-29, regrouped
    const c64* down;        // reference down-chirp table (ChirpGenerator.hpp:24-50), stored twice back to back
    const float* down_phase; // phase of down[i] in radians in [-pi, pi], stored twice back to back
    // work split of the non-list kernels (set by the launcher): each packet is cut into chunks_per_packet runs of
    // chunk_len consecutive windows, and a slot takes one run at a time
    unsigned n_packets, chunk_len, chunks_per_packet;
    FastDiv div_cpp;
    const float* window;    // Hann table or nullptr
    const float4* est;      // per packet: x = cfo, y = time_offset, z = scale, w = max_amp
    uint16_t* symbols;
    size_t symbol_stride;
    uint16_t* sync_raw;     // 2 per packet
    float* peaks;           // optional, n_windows
    Tw0 tw0;
    float one;              // run-time 1.0f (packed_math.cuh)
    float margin;           // fast kernel: relative top-1/top-2 margin of |X|^2 below which a window is rechecked
    unsigned* flag_count;   // fast kernel appends, recheck kernel reads
    unsigned* flag_list;
    unsigned flag_cap;
    unsigned long long* stat_rechecked; // running total of rechecked windows (lora_b200_recheck_stats)
    int max_ctas_per_sm;    // 0 = as many as fit; else a cap (room for a concurrent prepass)
};

struct PrepassArgs {
    const float2* iq;
    size_t packet_stride, count;
    int n_packets;
    int do_maxabs;     // legacy modes: LoRaDemod.cpp:59-77
    int scale_from_est; // the scan already ran (scan_kernels.cu): take est[p].z / .w from memory instead
    int symbols;       // windows to estimate over (min(total, 2) inside demodulate; all of them for estimate_offsets)
    int osr;
    int dechirp;       // multiply by the down-chirp first (fused caller-dechirp only)
    int use_scale;     // apply est[p].z
    int legacy_tie;    // LoRaDemod.cpp:101 lowest-index rule on equal power
    const c64* twp;    // per-stage twiddle table
    const c64* down;   // doubled down-chirp table
    const float* window;
    float power_scale; // float(20*log10(N)), LoRaDetector.hpp:29
    float4* est;
    Tw0 tw0;
    float one;
};

// Prepass and fused demod as work items of one persistent launch (demod_fast.cu receive_merged_kernel).
struct MergedArgs {
    PrepassArgs pre;
    BlockedDemodArgs dem;   // est must be pre.est; chunking is set by the launcher
    unsigned* queue;        // n_packets + 1 words, cleared by the launcher: [0] = queue head, [1 + p] = est[p] published
    unsigned* ready;        // = queue + 1
    unsigned n_items, lead, pre_items, dem_items; // set by the launcher
};

cudaError_t launch_prepass_blocked(int sf, const PrepassArgs& a, cudaStream_t st);
// PRECISION_FAST only: queue clear, merged kernel, recheck kernel.
cudaError_t launch_receive_merged(int sf, const MergedArgs& m, int sm_count, cudaStream_t st);
// precision FAST: returns after queueing the fast kernel and the recheck kernel; EXACT: the faithful kernel only.
cudaError_t launch_demod_blocked(int sf, const BlockedDemodArgs& a, int precision, int sm_count, cudaStream_t st);
// Fills the pass-0 twiddle block from the host twiddle table.
void fill_tw0(int sf, const float2* h_tw, Tw0* out);
// The per-stage regrouping of the twiddle table the blocked kernels read (Blk::TWP entries).
std::vector<float2> build_stage_twiddles(int sf, const float2* h_tw);
bool blocked_supported(int sf);

} // namespace lora_b200
