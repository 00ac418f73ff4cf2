// scan_kernels.cu -- the normalisation scan of lora_demodulate (LoRaDemod.cpp:59-77) as a kernel light enough to run
// BESIDE the demodulation kernel of the previous sub-batch.
//
// The legacy receive path reads every packet twice: max over all samples of max(|re|, |im|) first (the scale 1/max
// feeds the offset estimator's float rounding, so it must be final before any window is demodulated), then the
// windows.  The demod kernel is bound by the fp32 pipe, the scan by HBM; run one after the other
// they add up, run side by side they overlap.  The register-blocked prepass kernel cannot share an SM with three
// demod CTAs (96 registers x 256 threads); this one can: 128 threads, 32 registers (4096 of the 65536 the three demod
// CTAs leave free), no LDG in the loop --
//   * samples arrive by cp.async.bulk (TMA, SASS UBLKCP) into a ring of 8 KB shared-memory stages, completion on an
//     mbarrier per stage: the bytes in flight (24-32 KB per SM) do not depend on registers or on the four warps' issue;
//   * every thread looks at its samples with a floating-point FILTER: |x|^2 = re^2 + im^2 bounds both components of the
//     value the reference examines (x itself, or x * down[n mod N] in fused caller-dechirp mode, |down| <= 1 + 2^-23),
//     so only a sample with |x|^2 (1 + 1e-5) >= (running exact maximum)^2 can raise the maximum.  Those few (about one in
//     10^5 on noise) are evaluated with the reference's own rounded products and comparisons; everything else costs one
//     packed multiply, one add and one compare.  `!(q < bound)` sends NaN down the exact path, where the reference's
//     "a NaN never wins" comparison order applies.
// The result is bit-identical to prepass_blocked_kernel's scan (tests/test_gpu_parity.py::test_overlapped_scan_*).
#include "scan_kernels.h"

#include "demod_kernels.h" // g_launches
#include "bulk_copy.cuh"
#include "packed_math.cuh"

namespace lora_b200 {

namespace {

constexpr int kScanThreads = 128;
constexpr int kStageBytes = 8192;                 // 1024 samples
constexpr int kStages = 4;
constexpr int kPerThread = kStageBytes / 16 / kScanThreads; // 16-byte pairs per thread per stage


__global__ void __maxnreg__(32) scan_light_kernel(const ScanArgs a)
{
    __shared__ __align__(128) unsigned char stage[kStages][kStageBytes];
    __shared__ __align__(8) unsigned long long full[kStages];
    __shared__ float redf[kScanThreads / 32];
    const int tid = threadIdx.x;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // This CTA's work: packets blockIdx.x, blockIdx.x + grid, ...; each packet is `chunks` stages of up to 8 KB (the
    // last one shorter; an odd sample count leaves one sample for the tail below).  Items are numbered consecutively
    // across the CTA's packets so the ring never drains at a packet boundary.
    const size_t bytes = (a.count >> 1) << 4;                       // whole 16-byte pairs
    const unsigned chunks = (unsigned)((bytes + kStageBytes - 1) / kStageBytes);
    const unsigned last_bytes = (unsigned)(bytes - (size_t)(chunks - 1) * kStageBytes);
    const unsigned my_packets = blockIdx.x < (unsigned)a.n_packets ? ((unsigned)a.n_packets - blockIdx.x + gridDim.x - 1) / gridDim.x : 0u;
    const unsigned items = my_packets * chunks;
    // producer cursor (thread 0): item number, its packet ordinal and chunk
    unsigned pi = 0, pk = 0, pc = 0;
    auto issue = [&]() {
        const size_t p = blockIdx.x + (size_t)pk * gridDim.x;
        const unsigned n = pc + 1 == chunks ? last_bytes : (unsigned)kStageBytes;
        const unsigned s = pi % kStages;
        mbar_expect_tx(&full[s], n);
        bulk_load(stage[s], reinterpret_cast<const unsigned char*>(a.iq + p * a.packet_stride) + (size_t)pc * kStageBytes, n, &full[s]);
        ++pi;
        if (++pc == chunks) { pc = 0; ++pk; }
    };
    if (tid == 0)
        while (pi < items && pi < (unsigned)kStages) issue();

    float mx = 0.0f;   // exact running maximum of this thread (LoRaDemod.cpp:61-66)
    float b2 = 0.0f;   // filter threshold: mx^2, rounded down a little
    // the reference's comparison for one sample: a NaN real part drops the sample, a NaN imaginary part only itself
    auto take = [&](c64 v, size_t n) {
        if (a.down) v = cmul<true>(v, __ldg(a.down + (n & (size_t)(a.N - 1))), a.one);
        const float2 f = up(v);
        const float r = fabsf(f.x), im = fabsf(f.y);
        const float m = (r < im) ? im : r;
        if (m > mx) { mx = m; b2 = __fmul_rz(__fmul_rz(mx, mx), 0.99998f); }
    };
    unsigned k = 0, c = 0; // consumer cursor: packet ordinal and chunk of item `it`
    for (unsigned it = 0; it < items; ++it) {
        const unsigned s = it % kStages;
        const unsigned n_pairs = (c + 1 == chunks ? last_bytes : (unsigned)kStageBytes) >> 4;
        mbar_wait(&full[s], (it / kStages) & 1u);
        const ulonglong2* sp = reinterpret_cast<const ulonglong2*>(stage[s]);
        bool quiet = true;
#pragma unroll
        for (int j = 0; j < kPerThread; ++j) {
            const unsigned i = tid + j * kScanThreads;
            const ulonglong2 v = i < n_pairs ? sp[i] : make_ulonglong2(0ull, 0ull);
            const float2 q0 = up(mul2(v.x, v.x)), q1 = up(mul2(v.y, v.y));
            quiet = quiet && (__fadd_rn(q0.x, q0.y) < b2) && (__fadd_rn(q1.x, q1.y) < b2);
        }
        if (!quiet) { // rare: some sample of this thread may raise the maximum (or is NaN) -- the reference's arithmetic
            for (int j = 0; j < kPerThread; ++j) {
                const unsigned i = tid + j * kScanThreads;
                if (i < n_pairs) {
                    const ulonglong2 v = sp[i];
                    const size_t n0 = (size_t)c * (kStageBytes / 8) + 2 * (size_t)i;
                    take(v.x, n0);
                    take(v.y, n0 + 1);
                }
            }
        }
        __syncthreads(); // every thread has read stage s: it may be refilled
        if (tid == 0 && pi < items) issue();
        if (++c == chunks) { // packet complete
            const size_t p = blockIdx.x + (size_t)k * gridDim.x;
            if ((a.count & 1) && tid == 0) take(__ldg(reinterpret_cast<const c64*>(a.iq + p * a.packet_stride) + a.count - 1), a.count - 1);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { const float y = __shfl_xor_sync(0xffffffffu, mx, o); if (y > mx) mx = y; }
            if ((tid & 31) == 0) redf[tid >> 5] = mx;
            __syncthreads();
            if (tid == 0) {
#pragma unroll
                for (int w = 1; w < kScanThreads / 32; ++w) if (redf[w] > mx) mx = redf[w];
                // x, y (cfo, time offset) are written by the estimate that follows; z = scale, w = max_amp (:67-76)
                a.est[p] = make_float4(0.0f, 0.0f, (mx > 1.0f) ? __fdiv_rn(1.0f, mx) : 1.0f, mx);
            }
            __syncthreads(); // redf is reused by the next packet
            mx = 0.0f;
            b2 = 0.0f;
            c = 0;
            ++k;
        }
    }
}

} // namespace

bool scan_light_supported(const ScanArgs& a)
{
    // bulk copies need 16-byte aligned sources: every packet base, hence the batch base and the stride
    return a.count >= 2 && (reinterpret_cast<uintptr_t>(a.iq) & 15u) == 0 && (a.packet_stride & 1u) == 0;
}

cudaError_t launch_scan_light(const ScanArgs& a, int sm_count, cudaStream_t st)
{
    if (a.n_packets <= 0) return cudaSuccess;
    int grid = sm_count;
    if (grid > a.n_packets) grid = a.n_packets;
    scan_light_kernel<<<grid, kScanThreads, 0, st>>>(a);
    ++g_launches;
    return cudaGetLastError();
}

} // namespace lora_b200
