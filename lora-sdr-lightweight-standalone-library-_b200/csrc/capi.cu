// capi.cu -- the extern "C" boundary declared in include/lora_phy_b200.h.
//
// Thin by design: argument checks that mirror the reference's guards, table upload at context creation,
// grow-only device scratch (nothing is allocated once a batch size has been seen, the GPU counterpart of
// the reference's zero-allocation rule, tests/no_alloc_test.cpp), and kernel launches.  No arithmetic of
// the receive path happens on the host, and there is no CPU fallback: a missing device is an error.
#include "lora_phy_b200.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "code_kernels.h"
#include "demod_fast.h"
#include "demod_kernels.h"
#include "fft_kernels.h"
#include "host_tables.h"
#include "mod_kernels.h"
#include "scan_kernels.h"
#include "std_kernels.h"
#include "stream_kernels.h"

using namespace lora_b200;

static thread_local std::string g_error;

static int fail(int code, const char* what, cudaError_t e = cudaSuccess)
{
    g_error = what;
    if (e != cudaSuccess) { g_error += ": "; g_error += cudaGetErrorString(e); }
    return code;
}

#define CU(call)                                                          \
    do {                                                                  \
        cudaError_t e_ = (call);                                          \
        if (e_ != cudaSuccess) return fail(LORA_B200_ECUDA, #call, e_);   \
    } while (0)

// A grow-only device (or pinned host) buffer.
struct Buf {
    void* p = nullptr;
    size_t cap = 0;
    bool pinned = false;
    cudaError_t reserve(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) { pinned ? cudaFreeHost(p) : cudaFree(p); p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 4; // head-room so slowly growing batches do not reallocate each time
        cudaError_t e = pinned ? cudaMallocHost(&p, want) : cudaMalloc(&p, want);
        if (e != cudaSuccess) { want = bytes; e = pinned ? cudaMallocHost(&p, want) : cudaMalloc(&p, want); }
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release()
    {
        if (p) { pinned ? cudaFreeHost(p) : cudaFree(p); }
        p = nullptr; cap = 0;
    }
    template <typename T> T* as() const { return static_cast<T*>(p); }
};

struct lora_b200_ctx {
    lora_b200_config cfg{};
    int device = 0, sm_count = 148;
    int n = 0;
    float bw_scale = 1.0f, power_scale = 0.0f;
    std::vector<std::complex<float>> h_tw, h_down;
    std::vector<float> h_window;
    float2* d_tw = nullptr;
    float2* d_down = nullptr;
    float* d_window = nullptr;   // Hann table, or nullptr when window == none
    float* d_down_phase = nullptr; // phase of d_down in radians, 2N entries (fast demod flavour)
    float2* d_twp = nullptr;       // per-stage twiddle table of the register-blocked FFT
    Tw0 tw0{};                   // pass-0 twiddles of the register-blocked FFT (kernel arguments)
    int precision = PRECISION_FAST;
    Buf flags;                   // recheck list: [0] = count, [1..] = window numbers
    unsigned long long stat_windows = 0; // windows demodulated in PRECISION_FAST since the last stats call
    unsigned long long* d_stat_rechecked = nullptr; // device-side running count of rechecked windows
    uint8_t* d_whiten = nullptr; // 2 x WHITEN_LEN bytes
    float2* d_stream_tables = nullptr; // up (N), down (N), fine (128 N): built by the first stream_detect call
    Buf est, sync_raw;           // per-packet scratch of the demod kernels
    Buf sched;                   // merged launch: queue head + one published-flag per packet
    Buf s_iq, s_sym, s_metrics, s_payload, s_results, s_misc, s_raw; // staging for the host-buffer entry points
    Buf h_pin;                   // pinned bounce buffer
    cudaStream_t stream = nullptr;
    // prepass / demod pipeline of large batches: the scan+estimate of sub-batch k+1 runs beside the demod of k
    cudaStream_t s_pre = nullptr, s_demod = nullptr, s_copy = nullptr;
    std::vector<cudaEvent_t> ev_h2d;   // one per ingest chunk of the host-buffer receive calls
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    std::vector<cudaEvent_t> ev_pre;
    int overlap = 2;             // lora_b200_set_overlap: 2 = prepass and demod as items of one persistent launch (measured 1-11 % faster); 1 = scan of sub-batch k+1 beside the demod of sub-batch k (measured slower); 0 = two launches
    std::mutex mu;
    bool timing = false;                                     // lora_b200_set_kernel_timing
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev_pool; // reused event pairs
    size_t ev_used = 0;
};

extern "C" const char* lora_b200_last_error(void) { return g_error.c_str(); }
extern "C" unsigned long long lora_b200_launch_count(void) { return g_launches; }

extern "C" int lora_b200_ctx_create(const lora_b200_config* cfg, lora_b200_ctx** out)
{
    if (!cfg || !out) return fail(LORA_B200_EINVAL, "null argument");
    if (cfg->sf < 2 || cfg->sf > 12) return fail(LORA_B200_EINVAL, "sf must be 2..12 (N <= 4096, kissfft.hh:34)");
    if (cfg->bw_hz != 125000 && cfg->bw_hz != 250000 && cfg->bw_hz != 500000)
        return fail(LORA_B200_EINVAL, "bw_hz must be 125000, 250000 or 500000");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(LORA_B200_ECUDA, "no CUDA device: this library has no CPU path", e);
    int dev = cfg->device;
    if (dev < 0) CU(cudaGetDevice(&dev));
    if (dev >= count) return fail(LORA_B200_EINVAL, "device ordinal out of range");
    CU(cudaSetDevice(dev));

    lora_b200_ctx* c = new lora_b200_ctx();
    c->cfg = *cfg;
    if (c->cfg.osr == 0) c->cfg.osr = 1;
    c->device = dev;
    c->n = 1 << cfg->sf;
    c->bw_scale = static_cast<float>(cfg->bw_hz) / 125000.0f; // phy.hpp:43-49
    c->power_scale = detector_power_scale(c->n);
    cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, dev);
    c->h_tw = build_twiddles(c->n);
    c->h_down = build_downchirp(c->n, c->bw_scale);
    c->h_window = build_hann(c->n);
    std::vector<float> h_phase = build_downchirp_phase(c->n, c->bw_scale);
    h_phase.resize(2 * static_cast<size_t>(c->n)); // stored twice: no index wrap in the kernels
    std::copy(h_phase.begin(), h_phase.begin() + c->n, h_phase.begin() + c->n);
    std::vector<float2> h_twp;
    if (blocked_supported(static_cast<int>(cfg->sf)))
    {
        fill_tw0(static_cast<int>(cfg->sf), reinterpret_cast<const float2*>(c->h_tw.data()), &c->tw0);
        h_twp = build_stage_twiddles(static_cast<int>(cfg->sf), reinterpret_cast<const float2*>(c->h_tw.data()));
    }
    if (h_twp.empty()) h_twp.assign(1, make_float2(1.0f, 0.0f));
    c->h_pin.pinned = true;
    std::vector<uint8_t> wh(2 * WHITEN_LEN);
    build_whitening(wh.data(), wh.data() + WHITEN_LEN, WHITEN_LEN);

    const size_t nb = sizeof(float2) * c->n;
    bool ok = cudaMalloc(&c->d_tw, nb) == cudaSuccess && cudaMalloc(&c->d_down, 2 * nb) == cudaSuccess &&
              cudaMalloc(&c->d_twp, sizeof(float2) * h_twp.size()) == cudaSuccess &&
              cudaMalloc(&c->d_whiten, wh.size()) == cudaSuccess &&
              cudaMalloc(&c->d_down_phase, sizeof(float) * 2 * c->n) == cudaSuccess &&
              cudaMalloc(&c->d_stat_rechecked, sizeof(unsigned long long)) == cudaSuccess &&
              cudaMemset(c->d_stat_rechecked, 0, sizeof(unsigned long long)) == cudaSuccess &&
              cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess;
    if (ok) {
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        ok = cudaStreamCreateWithPriority(&c->s_pre, cudaStreamNonBlocking, hi) == cudaSuccess &&
             cudaStreamCreateWithPriority(&c->s_demod, cudaStreamNonBlocking, lo) == cudaSuccess &&
             cudaStreamCreateWithFlags(&c->s_copy, cudaStreamNonBlocking) == cudaSuccess &&
             cudaEventCreateWithFlags(&c->ev_begin, cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&c->ev_end, cudaEventDisableTiming) == cudaSuccess;
    }
    if (ok && cfg->window == 1) ok = cudaMalloc(&c->d_window, sizeof(float) * c->n) == cudaSuccess;
    ok = ok && cudaMemcpy(c->d_tw, c->h_tw.data(), nb, cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(c->d_down, c->h_down.data(), nb, cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(c->d_down + c->n, c->h_down.data(), nb, cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(c->d_twp, h_twp.data(), sizeof(float2) * h_twp.size(), cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(c->d_whiten, wh.data(), wh.size(), cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(c->d_down_phase, h_phase.data(), sizeof(float) * 2 * c->n, cudaMemcpyHostToDevice) == cudaSuccess;
    if (ok && c->d_window)
        ok = cudaMemcpy(c->d_window, c->h_window.data(), sizeof(float) * c->n, cudaMemcpyHostToDevice) == cudaSuccess;
    if (!ok) {
        const cudaError_t last = cudaGetLastError();
        lora_b200_ctx_destroy(c);
        return fail(LORA_B200_ECUDA, "context setup", last);
    }
    *out = c;
    return LORA_B200_OK;
}

extern "C" void lora_b200_ctx_destroy(lora_b200_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    cudaFree(c->d_tw); cudaFree(c->d_down); cudaFree(c->d_window); cudaFree(c->d_whiten); cudaFree(c->d_down_phase); cudaFree(c->d_twp);
    cudaFree(c->d_stat_rechecked);
    cudaFree(c->d_stream_tables);
    for (Buf* b : {&c->flags, &c->est, &c->sync_raw, &c->sched, &c->s_iq, &c->s_sym, &c->s_metrics, &c->s_payload, &c->s_results,
                   &c->s_misc, &c->s_raw, &c->h_pin})
        b->release();
    for (auto& pr : c->ev_pool) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->s_pre) cudaStreamDestroy(c->s_pre);
    if (c->s_demod) cudaStreamDestroy(c->s_demod);
    if (c->s_copy) cudaStreamDestroy(c->s_copy);
    for (cudaEvent_t e : c->ev_h2d) cudaEventDestroy(e);
    if (c->ev_begin) cudaEventDestroy(c->ev_begin);
    if (c->ev_end) cudaEventDestroy(c->ev_end);
    for (cudaEvent_t e : c->ev_pre) cudaEventDestroy(e);
    delete c;
}

extern "C" int lora_b200_set_kernel_timing(lora_b200_ctx* c, int enable)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    c->timing = enable != 0;
    c->ev_used = 0;
    return LORA_B200_OK;
}

extern "C" int lora_b200_kernel_time(lora_b200_ctx* c, double* total_ms, unsigned long long* launches)
{
    if (!c || !total_ms || !launches) return fail(LORA_B200_EINVAL, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    double sum = 0.0;
    for (size_t i = 0; i < c->ev_used; ++i) {
        CU(cudaEventSynchronize(c->ev_pool[i].second));
        float ms = 0.0f;
        CU(cudaEventElapsedTime(&ms, c->ev_pool[i].first, c->ev_pool[i].second));
        sum += ms;
    }
    *total_ms = sum;
    *launches = c->ev_used;
    c->ev_used = 0;
    return LORA_B200_OK;
}

extern "C" int lora_b200_set_precision(lora_b200_ctx* c, int precision)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    if (precision != LORA_B200_PRECISION_FAST && precision != LORA_B200_PRECISION_EXACT)
        return fail(LORA_B200_EINVAL, "unknown precision");
    std::lock_guard<std::mutex> lk(c->mu);
    c->precision = precision;
    return LORA_B200_OK;
}

extern "C" int lora_b200_set_overlap(lora_b200_ctx* c, int enable)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    c->overlap = enable < 0 ? 0 : (enable > 3 ? 3 : enable);
    return LORA_B200_OK;
}

extern "C" int lora_b200_recheck_stats(lora_b200_ctx* c, unsigned long long* windows, unsigned long long* rechecked)
{
    if (!c || !windows || !rechecked) return fail(LORA_B200_EINVAL, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    CU(cudaDeviceSynchronize());
    unsigned long long r = 0;
    CU(cudaMemcpy(&r, c->d_stat_rechecked, sizeof(r), cudaMemcpyDeviceToHost));
    CU(cudaMemset(c->d_stat_rechecked, 0, sizeof(r)));
    *windows = c->stat_windows;
    *rechecked = r;
    c->stat_windows = 0;
    return LORA_B200_OK;
}

extern "C" int lora_b200_ctx_table(const lora_b200_ctx* c, int table, void* h_out, size_t n)
{
    if (!c || !h_out || n != static_cast<size_t>(c->n)) return fail(LORA_B200_EINVAL, "bad table request");
    switch (table) {
    case 0: std::memcpy(h_out, c->h_tw.data(), sizeof(float2) * n); return LORA_B200_OK;
    case 1: std::memcpy(h_out, c->h_down.data(), sizeof(float2) * n); return LORA_B200_OK;
    case 2: std::memcpy(h_out, c->h_window.data(), sizeof(float) * n); return LORA_B200_OK;
    default: return fail(LORA_B200_EINVAL, "unknown table");
    }
}

// ---- receive path -------------------------------------------------------------------------------------------
static int demod_batch_locked(lora_b200_ctx* c, int mode, const float2* d_iq, size_t n_packets, size_t count,
                              size_t packet_stride, uint16_t* d_symbols, size_t symbol_stride, float* d_peaks,
                              lora_b200_pkt_metrics* d_metrics, cudaStream_t st)
{
    const size_t N = static_cast<size_t>(c->n), osr = c->cfg.osr, step = N * osr;
    const bool legacy = mode != LORA_B200_DEMOD_WORKSPACE;
    if (mode < 0 || mode > 2) return fail(LORA_B200_EINVAL, "unknown demod mode");
    if (!d_iq || !d_symbols) return fail(LORA_B200_EINVAL, "null buffer");              // phy.cpp:181
    if (n_packets == 0) return LORA_B200_OK;
    if (n_packets > 0x7fffffffu) return fail(LORA_B200_EINVAL, "too many packets");
    if (packet_stride < count) return fail(LORA_B200_EINVAL, "packet_stride < samples_per_packet");
    const size_t total = count / step;
    if (!legacy) {
        if (count % step != 0) return fail(LORA_B200_EINVAL, "sample count not a multiple of N*osr"); // :186
        if (total < 2) return fail(LORA_B200_EINVAL, "fewer than two symbols");                        // :188
    }
    if (mode == LORA_B200_DEMOD_LEGACY_DECHIRP && (osr != 1 || count % N != 0))
        return fail(LORA_B200_EINVAL, "fused caller-dechirp needs osr == 1 and whole symbols");
    const bool have_sync = total >= 2;
    const size_t n_out = have_sync ? total - 2 : total;
    if (symbol_stride < n_out) return fail(LORA_B200_EINVAL, "symbol_stride too small");
    if (total > 0xffff) return fail(LORA_B200_EINVAL, "more than 65535 symbols per packet");

    CU(c->est.reserve(sizeof(float4) * n_packets));
    CU(c->sync_raw.reserve(sizeof(uint16_t) * 2 * n_packets));
    float4* est = c->est.as<float4>();
    const int np = static_cast<int>(n_packets);
    const bool dechirp_in = mode == LORA_B200_DEMOD_LEGACY_DECHIRP;

    if (blocked_supported(static_cast<int>(c->cfg.sf))) {
        // SF7..SF12: register-blocked kernels (demod_fast.cu).  Batches are cut so that a launch sees < 2^31 windows.
        const int sf = static_cast<int>(c->cfg.sf);
        size_t chunk_packets = total ? std::max<size_t>(1, 0x7fffffffull / total) : n_packets;
        // Large legacy batches are cut into 8 sub-batches and pipelined over two internal streams: the normalisation scan
        // (HBM-bound; scan_kernels.cu, 128 threads x 32 registers per SM) of sub-batch k+1 runs BESIDE the offset
        // estimate and the demod kernel (fp32-bound, three 256-thread CTAs per SM) of sub-batch k, so the second read
        // of every sample costs bandwidth the demod kernel leaves idle instead of time.
        ScanArgs probe{d_iq, packet_stride, count, np, c->n, nullptr, 1.0f, est};
        const bool pipelined = c->overlap == 1 && legacy && total > 0 && n_packets >= 64 && scan_light_supported(probe) &&
                               n_packets * count * sizeof(float2) >= (size_t(256) << 20);
        // overlap == 2: prepass and fused demod as work items of ONE persistent launch (demod_fast.cu
        // receive_merged_kernel) -- the HBM-bound scan of some packets runs beside the issue-bound transform of others
        // Measured (profiles/r2_ab_merged_schedule.txt, small-batch table): the persistent launch wins from about 0.6 GB
        // of samples per call (0.3 ms of work) and loses below -- its queue needs a few waves of items to overlap anything.
        const bool merged = c->overlap >= 2 && c->precision == PRECISION_FAST && total > 0 && n_packets >= 256 &&
                            (c->overlap == 3 || n_packets * count * sizeof(float2) >= (size_t(600) << 20));
        if (merged) CU(c->sched.reserve(sizeof(unsigned) * (std::min(n_packets, chunk_packets) + 1)));
        if (pipelined) chunk_packets = std::min(chunk_packets, (n_packets + 7) / 8);
        const size_t n_sub = (n_packets + chunk_packets - 1) / chunk_packets;
        size_t max_windows = std::min(n_packets, chunk_packets) * total;
        if (total > 0 && c->precision == PRECISION_FAST) CU(c->flags.reserve(sizeof(unsigned) * (max_windows + 1)));
        cudaStream_t st_pre = st, st_dem = st;
        if (pipelined) {
            while (c->ev_pre.size() < n_sub) {
                cudaEvent_t e = nullptr;
                CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                c->ev_pre.push_back(e);
            }
            st_pre = c->s_pre; st_dem = c->s_demod;
            CU(cudaEventRecord(c->ev_begin, st));
            CU(cudaStreamWaitEvent(st_pre, c->ev_begin, 0));
            CU(cudaStreamWaitEvent(st_dem, c->ev_begin, 0));
        }
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        size_t sub = 0;
        for (size_t p0 = 0; p0 < n_packets; p0 += chunk_packets, ++sub) {
            const size_t pn = std::min(chunk_packets, n_packets - p0);
            PrepassArgs pa{};
            pa.iq = d_iq + p0 * packet_stride; pa.packet_stride = packet_stride; pa.count = count;
            pa.n_packets = static_cast<int>(pn);
            pa.do_maxabs = legacy && count > 0;
            pa.symbols = static_cast<int>(total < 2 ? total : 2);
            pa.osr = static_cast<int>(osr);
            pa.dechirp = dechirp_in;   // demodulate() estimates on RAW samples (phy.cpp:192-193)
            pa.use_scale = legacy;
            pa.legacy_tie = legacy;
            pa.twp = reinterpret_cast<const c64*>(c->d_twp); pa.down = reinterpret_cast<const c64*>(c->d_down);
            pa.window = c->d_window;
            pa.power_scale = c->power_scale;
            pa.est = est + p0;
            pa.tw0 = c->tw0; pa.one = 1.0f;
            if (pipelined) {
                ScanArgs sa{pa.iq, packet_stride, count, pa.n_packets, c->n, dechirp_in ? pa.down : nullptr, 1.0f, est + p0};
                CU(launch_scan_light(sa, c->sm_count, st_pre));
                CU(cudaEventRecord(c->ev_pre[sub], st_pre));
                CU(cudaStreamWaitEvent(st_dem, c->ev_pre[sub], 0));
                pa.do_maxabs = 0;
                pa.scale_from_est = 1;
            }
            if (!merged) CU(launch_prepass_blocked(sf, pa, st_dem));
            if (total == 0) continue;

            BlockedDemodArgs da{};
            da.iq = pa.iq; da.packet_stride = packet_stride; da.count = count;
            da.n_windows = static_cast<unsigned>(pn * total);
            da.n_packets = static_cast<unsigned>(pn);
            da.total = static_cast<int>(total);
            da.div_total = make_fastdiv(static_cast<unsigned>(total));
            da.osr = static_cast<int>(osr);
            da.dechirp = mode != LORA_B200_DEMOD_LEGACY;
            da.dechirp_absolute = dechirp_in;
            da.use_scale = legacy;
            da.have_sync = have_sync;
            da.twp = pa.twp; da.down = pa.down; da.down_phase = c->d_down_phase; da.window = c->d_window;
            da.est = est + p0;
            da.symbols = d_symbols + p0 * symbol_stride; da.symbol_stride = symbol_stride;
            da.sync_raw = c->sync_raw.as<uint16_t>() + 2 * p0;
            da.peaks = d_peaks ? d_peaks + p0 * total : nullptr;
            da.tw0 = c->tw0; da.one = 1.0f;
            da.margin = 1.0f / 8192.0f;
            da.flag_count = c->flags.as<unsigned>();
            da.flag_list = c->flags.as<unsigned>() ? c->flags.as<unsigned>() + 1 : nullptr;
            da.flag_cap = da.n_windows;
            da.stat_rechecked = c->d_stat_rechecked;
            da.max_ctas_per_sm = 0;
            if (c->timing) {
                if (c->ev_used == c->ev_pool.size()) {
                    cudaEvent_t a = nullptr, b = nullptr;
                    CU(cudaEventCreate(&a));
                    CU(cudaEventCreate(&b));
                    c->ev_pool.emplace_back(a, b);
                }
                e0 = c->ev_pool[c->ev_used].first;
                e1 = c->ev_pool[c->ev_used].second;
                ++c->ev_used;
                CU(cudaEventRecord(e0, st_dem));
            }
            if (merged) {
                MergedArgs ma{};
                ma.pre = pa; ma.dem = da;
                ma.queue = c->sched.as<unsigned>(); ma.ready = ma.queue + 1;
                CU(launch_receive_merged(sf, ma, c->sm_count, st_dem));
            } else {
                CU(launch_demod_blocked(sf, da, c->precision, c->sm_count, st_dem));
            }
            if (c->timing) CU(cudaEventRecord(e1, st_dem));
            if (c->precision == PRECISION_FAST) c->stat_windows += da.n_windows;
        }
        if (pipelined) {
            CU(cudaEventRecord(c->ev_end, st_dem));
            CU(cudaStreamWaitEvent(st, c->ev_end, 0));
        }
        if (d_metrics) {
            FinalizeArgs fa{est, c->sync_raw.as<uint16_t>(), d_metrics, np, static_cast<int>(c->cfg.sf),
                            have_sync, legacy, static_cast<int>(n_out)};
            CU(launch_finalize(fa, st));
        }
        return LORA_B200_OK;
    }

    CU(launch_fill_est(est, np, st));
    if (legacy && count > 0) {
        MaxAbsArgs m{d_iq, packet_stride, count, dechirp_in ? c->d_down : nullptr, c->n, est};
        CU(launch_maxabs(m, np, st));
    }
    if (total > 0) {
        EstimateArgs ea{};
        ea.iq = d_iq; ea.packet_stride = packet_stride;
        ea.symbols = static_cast<int>(total < 2 ? total : 2);
        ea.osr = static_cast<int>(osr);
        ea.dechirp = dechirp_in;   // demodulate() estimates on RAW samples (phy.cpp:192-193)
        ea.use_scale = legacy;
        ea.legacy_tie = legacy;
        ea.tw = c->d_tw; ea.down = c->d_down; ea.window = c->d_window;
        ea.power_scale = c->power_scale;
        ea.est = est;
        CU(launch_estimate(static_cast<int>(c->cfg.sf), ea, np, st));

        DemodArgs da{};
        da.iq = d_iq; da.packet_stride = packet_stride; da.count = count;
        da.n_windows = static_cast<long long>(n_packets) * static_cast<long long>(total);
        da.total = static_cast<int>(total);
        da.osr = static_cast<int>(osr);
        da.dechirp = mode != LORA_B200_DEMOD_LEGACY;
        da.dechirp_absolute = dechirp_in;
        da.use_scale = legacy;
        da.have_sync = have_sync;
        da.tw = c->d_tw; da.down = c->d_down; da.window = c->d_window;
        da.est = est;
        da.symbols = d_symbols; da.symbol_stride = symbol_stride;
        da.sync_raw = c->sync_raw.as<uint16_t>();
        da.peaks = d_peaks;
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (c->timing) {
            if (c->ev_used == c->ev_pool.size()) {
                cudaEvent_t a = nullptr, b = nullptr;
                CU(cudaEventCreate(&a));
                CU(cudaEventCreate(&b));
                c->ev_pool.emplace_back(a, b);
            }
            e0 = c->ev_pool[c->ev_used].first;
            e1 = c->ev_pool[c->ev_used].second;
            ++c->ev_used;
            CU(cudaEventRecord(e0, st));
        }
        CU(launch_demod(static_cast<int>(c->cfg.sf), da, c->sm_count, st));
        if (e1) CU(cudaEventRecord(e1, st));
    }
    if (d_metrics) {
        FinalizeArgs fa{est, c->sync_raw.as<uint16_t>(), d_metrics, np, static_cast<int>(c->cfg.sf),
                        have_sync, legacy, static_cast<int>(n_out)};
        CU(launch_finalize(fa, st));
    }
    return LORA_B200_OK;
}

extern "C" int lora_b200_demod_batch(lora_b200_ctx* c, int mode, const lora_b200_c32* d_iq, size_t n_packets,
                                     size_t samples_per_packet, size_t packet_stride, uint16_t* d_symbols,
                                     size_t symbol_stride, float* d_peaks, lora_b200_pkt_metrics* d_metrics,
                                     void* stream)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    return demod_batch_locked(c, mode, reinterpret_cast<const float2*>(d_iq), n_packets, samples_per_packet,
                              packet_stride, d_symbols, symbol_stride, d_peaks, d_metrics,
                              static_cast<cudaStream_t>(stream));
}

// ---- receive path without allocation ----------------------------------------------------------------------
// lora_b200_demod_batch_static: the blocked receive kernels (prepass -> fused demod -> recheck -> finalize) with their
// per-call scratch -- offset estimates, raw sync symbols, the recheck list -- in a fixed file-scope device array
// instead of the context's grow-only buffers.  Nothing is allocated or freed, on the device or on the host, so the
// call is usable where cudaMalloc is not (stream capture, serving loops).  Every GPU that loads this library holds
// the array (kStaticScratchBytes, about 4.3 MB) until the process exits.  All calls on a device SHARE it and it keeps
// the previous call's contents: the one word that must start at zero (the recheck count) is cleared on the stream by
// the launcher (cudaMemsetAsync), everything else is written before it is read.  Calls that overlap on the GPU --
// two streams, two contexts, two host threads -- race on it; the caller orders them.
constexpr size_t kStaticMaxPackets = LORA_B200_STATIC_MAX_PACKETS;
constexpr size_t kStaticMaxWindows = LORA_B200_STATIC_MAX_WINDOWS;
constexpr size_t kStaticEstOff = 0;                                                        // float4[packets]
constexpr size_t kStaticSyncOff = kStaticEstOff + sizeof(float4) * kStaticMaxPackets;     // uint16_t[2 * packets]
constexpr size_t kStaticFlagOff = kStaticSyncOff + sizeof(uint16_t) * 2 * kStaticMaxPackets; // unsigned[1 + windows]
constexpr size_t kStaticScratchBytes = ((kStaticFlagOff + sizeof(unsigned) * (kStaticMaxWindows + 1)) + 255) / 256 * 256;
static_assert(kStaticSyncOff % 256 == 0 && kStaticFlagOff % 256 == 0, "sections stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// In host code `g_static_scratch` is only a handle for the runtime, not a device address: the address is looked up
// once per device, with that device current, and cached.  (The lookup does not allocate; the array came with the module.)
static int static_scratch_base(int device, unsigned char** out)
{
    static std::mutex mu;
    static unsigned char* base[kMaxDevices] = {};
    if (device < 0 || device >= kMaxDevices) return fail(LORA_B200_EINVAL, "device ordinal beyond the static scratch table");
    std::lock_guard<std::mutex> lk(mu);
    if (!base[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) return fail(LORA_B200_ECUDA, "cudaGetSymbolAddress(static scratch)", e);
        base[device] = static_cast<unsigned char*>(p);
    }
    *out = base[device];
    return LORA_B200_OK;
}

extern "C" int lora_b200_demod_batch_static(lora_b200_ctx* c, int mode, const lora_b200_c32* d_iq_, size_t n_packets,
                                            size_t count, size_t packet_stride, uint16_t* d_symbols,
                                            size_t symbol_stride, float* d_peaks, lora_b200_pkt_metrics* d_metrics,
                                            void* stream)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    const float2* d_iq = reinterpret_cast<const float2*>(d_iq_);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int sf = static_cast<int>(c->cfg.sf);
    const size_t N = static_cast<size_t>(c->n), osr = c->cfg.osr, step = N * osr;
    const bool legacy = mode != LORA_B200_DEMOD_WORKSPACE;
    if (mode < 0 || mode > 2) return fail(LORA_B200_EINVAL, "unknown demod mode");
    if (!blocked_supported(sf)) return fail(LORA_B200_EINVAL, "lora_b200_demod_batch_static needs sf 7..12");
    if (!d_iq || !d_symbols) return fail(LORA_B200_EINVAL, "null buffer");
    if (n_packets == 0) return LORA_B200_OK;
    if (n_packets > kStaticMaxPackets)
        return fail(LORA_B200_EINVAL, "lora_b200_demod_batch_static: more than LORA_B200_STATIC_MAX_PACKETS (16384) packets");
    if (packet_stride < count) return fail(LORA_B200_EINVAL, "packet_stride < samples_per_packet");
    const size_t total = count / step;
    if (!legacy) {
        if (count % step != 0) return fail(LORA_B200_EINVAL, "sample count not a multiple of N*osr");
        if (total < 2) return fail(LORA_B200_EINVAL, "fewer than two symbols");
    }
    if (mode == LORA_B200_DEMOD_LEGACY_DECHIRP && (osr != 1 || count % N != 0))
        return fail(LORA_B200_EINVAL, "fused caller-dechirp needs osr == 1 and whole symbols");
    const bool have_sync = total >= 2;
    const size_t n_out = have_sync ? total - 2 : total;
    if (symbol_stride < n_out) return fail(LORA_B200_EINVAL, "symbol_stride too small");
    if (total > 0xffff) return fail(LORA_B200_EINVAL, "more than 65535 symbols per packet");
    if (n_packets * total > kStaticMaxWindows)
        return fail(LORA_B200_EINVAL,
                    "lora_b200_demod_batch_static: more than LORA_B200_STATIC_MAX_WINDOWS (1048576) symbol windows in one call");

    unsigned char* base = nullptr;
    if (int rc = static_scratch_base(c->device, &base)) return rc;
    float4* est = reinterpret_cast<float4*>(base + kStaticEstOff);
    uint16_t* sync_raw = reinterpret_cast<uint16_t*>(base + kStaticSyncOff);
    unsigned* flags = reinterpret_cast<unsigned*>(base + kStaticFlagOff);
    const int np = static_cast<int>(n_packets);
    const bool dechirp_in = mode == LORA_B200_DEMOD_LEGACY_DECHIRP;

    PrepassArgs pa{};
    pa.iq = d_iq; pa.packet_stride = packet_stride; pa.count = count;
    pa.n_packets = np;
    pa.do_maxabs = legacy && count > 0;
    pa.symbols = static_cast<int>(total < 2 ? total : 2);
    pa.osr = static_cast<int>(osr);
    pa.dechirp = dechirp_in;
    pa.use_scale = legacy;
    pa.legacy_tie = legacy;
    pa.twp = reinterpret_cast<const c64*>(c->d_twp); pa.down = reinterpret_cast<const c64*>(c->d_down);
    pa.window = c->d_window;
    pa.power_scale = c->power_scale;
    pa.est = est;
    pa.tw0 = c->tw0; pa.one = 1.0f;
    CU(launch_prepass_blocked(sf, pa, st));
    if (total > 0) {
        BlockedDemodArgs da{};
        da.iq = d_iq; da.packet_stride = packet_stride; da.count = count;
        da.n_windows = static_cast<unsigned>(n_packets * total);
        da.n_packets = static_cast<unsigned>(n_packets);
        da.total = static_cast<int>(total);
        da.div_total = make_fastdiv(static_cast<unsigned>(total));
        da.osr = static_cast<int>(osr);
        da.dechirp = mode != LORA_B200_DEMOD_LEGACY;
        da.dechirp_absolute = dechirp_in;
        da.use_scale = legacy;
        da.have_sync = have_sync;
        da.twp = pa.twp; da.down = pa.down; da.down_phase = c->d_down_phase; da.window = c->d_window;
        da.est = est;
        da.symbols = d_symbols; da.symbol_stride = symbol_stride;
        da.sync_raw = sync_raw;
        da.peaks = d_peaks;
        da.tw0 = c->tw0; da.one = 1.0f;
        da.margin = 1.0f / 8192.0f;
        da.flag_count = flags;      // cleared by launch_demod_blocked on `st` before the fused kernel appends to it
        da.flag_list = flags + 1;
        da.flag_cap = da.n_windows;
        da.stat_rechecked = c->d_stat_rechecked;
        da.max_ctas_per_sm = 0;
        CU(launch_demod_blocked(sf, da, c->precision, c->sm_count, st));
        if (c->precision == PRECISION_FAST) c->stat_windows += da.n_windows;
    }
    if (d_metrics) {
        FinalizeArgs fa{est, sync_raw, d_metrics, np, sf, have_sync, legacy, static_cast<int>(n_out)};
        CU(launch_finalize(fa, st));
    }
    return LORA_B200_OK;
}

static int estimate_batch_locked(lora_b200_ctx* c, const float2* d_iq, size_t n_packets, size_t count,
                                 size_t packet_stride, lora_b200_pkt_metrics* d_metrics, cudaStream_t st)
{
    const size_t step = static_cast<size_t>(c->n) * c->cfg.osr;
    if (!d_iq || !d_metrics) return fail(LORA_B200_EINVAL, "null buffer");
    const size_t symbols = count / step; // phy.cpp:86-87: every whole symbol offered is used
    if (n_packets == 0 || symbols == 0) return LORA_B200_OK; // :83,88 silently return
    if (symbols > 0x7fffffff) return fail(LORA_B200_EINVAL, "too many symbols");
    CU(c->est.reserve(sizeof(float4) * n_packets));
    CU(c->sync_raw.reserve(sizeof(uint16_t) * 2 * n_packets));
    const int np = static_cast<int>(n_packets);
    if (blocked_supported(static_cast<int>(c->cfg.sf))) {
        PrepassArgs pa{};
        pa.iq = d_iq; pa.packet_stride = packet_stride; pa.count = count;
        pa.n_packets = np;
        pa.symbols = static_cast<int>(symbols);
        pa.osr = static_cast<int>(c->cfg.osr);
        pa.twp = reinterpret_cast<const c64*>(c->d_twp); pa.down = reinterpret_cast<const c64*>(c->d_down);
        pa.window = c->d_window;
        pa.power_scale = c->power_scale;
        pa.est = c->est.as<float4>();
        pa.tw0 = c->tw0; pa.one = 1.0f;
        CU(launch_prepass_blocked(static_cast<int>(c->cfg.sf), pa, st));
        FinalizeArgs fb{c->est.as<float4>(), c->sync_raw.as<uint16_t>(), d_metrics, np, static_cast<int>(c->cfg.sf), 0, 0, 0};
        CU(launch_finalize(fb, st));
        return LORA_B200_OK;
    }
    CU(launch_fill_est(c->est.as<float4>(), np, st));
    EstimateArgs ea{};
    ea.iq = d_iq; ea.packet_stride = packet_stride;
    ea.symbols = static_cast<int>(symbols);
    ea.osr = static_cast<int>(c->cfg.osr);
    ea.tw = c->d_tw; ea.down = c->d_down; ea.window = c->d_window;
    ea.power_scale = c->power_scale;
    ea.est = c->est.as<float4>();
    CU(launch_estimate(static_cast<int>(c->cfg.sf), ea, np, st));
    FinalizeArgs fa{c->est.as<float4>(), c->sync_raw.as<uint16_t>(), d_metrics, np, static_cast<int>(c->cfg.sf), 0, 0, 0};
    CU(launch_finalize(fa, st));
    return LORA_B200_OK;
}

extern "C" int lora_b200_estimate_batch(lora_b200_ctx* c, const lora_b200_c32* d_iq, size_t n_packets,
                                        size_t samples_per_packet, size_t packet_stride,
                                        lora_b200_pkt_metrics* d_metrics, void* stream)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    return estimate_batch_locked(c, reinterpret_cast<const float2*>(d_iq), n_packets, samples_per_packet,
                                 packet_stride, d_metrics, static_cast<cudaStream_t>(stream));
}

// ---- integer back end ------------------------------------------------------------------------------------
static int check_chain_cfg(const lora_b200_chain_cfg* g)
{
    if (!g) return fail(LORA_B200_EINVAL, "null chain config");
    const unsigned ppm = g->ppm ? g->ppm : g->sf;
    if (g->sf < 5 || g->sf > 12) return fail(LORA_B200_EINVAL, "chain sf must be 5..12");
    if (ppm > g->sf) return fail(LORA_B200_EINVAL, "failed check: PPM <= SF (LoRaDecoder.cpp:202)");
    if (ppm < 5) return fail(LORA_B200_EINVAL, "ppm must be >= 5 (the header needs 5 codewords)");
    if (g->rdd > 4) return fail(LORA_B200_EINVAL, "unknown coding rate (LoRaDecoder.cpp:143-151)");
    if (g->data_length > 255) return fail(LORA_B200_EINVAL, "data_length > 255");
    return LORA_B200_OK;
}

extern "C" int lora_b200_decode_batch(lora_b200_ctx* c, const uint16_t* d_symbols, size_t n_packets,
                                      size_t symbols_per_packet, size_t symbol_stride, uint8_t* d_bytes,
                                      size_t byte_stride, uint8_t* d_crc_ok, void* stream)
{
    if (!c || !d_symbols || !d_bytes) return fail(LORA_B200_EINVAL, "null argument"); // phy.cpp:244
    if (n_packets == 0) return LORA_B200_OK;
    if (n_packets > 0x7fffffffu || symbol_stride < symbols_per_packet || byte_stride < symbols_per_packet / 2)
        return fail(LORA_B200_EINVAL, "bad batch geometry");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    SimpleDecodeArgs a{d_symbols, symbols_per_packet, symbol_stride, d_bytes, byte_stride, d_crc_ok,
                       static_cast<int>(n_packets)};
    CU(launch_simple_decode(a, static_cast<cudaStream_t>(stream)));
    return LORA_B200_OK;
}

extern "C" int lora_b200_encode_batch(lora_b200_ctx* c, const uint8_t* d_bytes, size_t n_packets,
                                      size_t bytes_per_packet, size_t byte_stride, uint16_t* d_symbols,
                                      size_t symbol_stride, void* stream)
{
    if (!c || !d_bytes || !d_symbols) return fail(LORA_B200_EINVAL, "null argument"); // phy.cpp:58
    if (n_packets == 0) return LORA_B200_OK;
    if (n_packets > 0x7fffffffu || byte_stride < bytes_per_packet || symbol_stride < 2 * bytes_per_packet)
        return fail(LORA_B200_EINVAL, "bad batch geometry");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    SimpleEncodeArgs a{d_bytes, bytes_per_packet, byte_stride, d_symbols, symbol_stride, static_cast<int>(n_packets)};
    CU(launch_simple_encode(a, static_cast<cudaStream_t>(stream)));
    return LORA_B200_OK;
}

static int chain_decode_locked(lora_b200_ctx* c, const lora_b200_chain_cfg* g, const uint16_t* d_symbols,
                               size_t n_packets, size_t nsym, size_t symbol_stride, uint8_t* d_payload,
                               lora_b200_chain_result* d_results, cudaStream_t st)
{
    if (int rc = check_chain_cfg(g)) return rc;
    if (!d_symbols || !d_payload || !d_results) return fail(LORA_B200_EINVAL, "null buffer");
    if (n_packets == 0) return LORA_B200_OK;
    if (n_packets > 0x7fffffffu || symbol_stride < nsym || nsym > 0xffff)
        return fail(LORA_B200_EINVAL, "bad batch geometry");
    ChainDecodeArgs a{*g, d_symbols, nsym, symbol_stride, d_payload, d_results, c->d_whiten, static_cast<int>(n_packets),
                      nullptr, nullptr, 0};
    CU(launch_chain_decode(a, st));
    return LORA_B200_OK;
}

extern "C" int lora_b200_chain_decode_batch(lora_b200_ctx* c, const lora_b200_chain_cfg* g,
                                            const uint16_t* d_symbols, size_t n_packets, size_t nsym,
                                            size_t symbol_stride, uint8_t* d_payload,
                                            lora_b200_chain_result* d_results, void* stream)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    return chain_decode_locked(c, g, d_symbols, n_packets, nsym, symbol_stride, d_payload, d_results,
                               static_cast<cudaStream_t>(stream));
}

extern "C" int lora_b200_chain_encode_batch(lora_b200_ctx* c, const lora_b200_chain_cfg* g,
                                            const uint8_t* d_payload, size_t n_packets, size_t payload_len,
                                            size_t payload_stride, uint16_t* d_symbols, size_t symbol_stride,
                                            size_t* symbols_per_packet, void* stream)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    if (int rc = check_chain_cfg(g)) return rc;
    if (payload_len > 255) return fail(LORA_B200_EINVAL, "payload longer than 255 bytes");
    const size_t nsym = chain_symbol_count(*g, payload_len);
    if (nsym == 0) return fail(LORA_B200_EINVAL, "invalid chain configuration");
    if (symbols_per_packet) *symbols_per_packet = nsym;
    if (!d_symbols) return LORA_B200_OK; // size query
    if (!d_payload && payload_len) return fail(LORA_B200_EINVAL, "null payload");
    if (n_packets == 0) return LORA_B200_OK;
    if (n_packets > 0x7fffffffu || symbol_stride < nsym || payload_stride < payload_len)
        return fail(LORA_B200_EINVAL, "bad batch geometry");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    ChainEncodeArgs a{*g, d_payload, payload_len, payload_stride, d_symbols, symbol_stride, c->d_whiten,
                      static_cast<int>(n_packets)};
    CU(launch_chain_encode(a, static_cast<cudaStream_t>(stream)));
    return LORA_B200_OK;
}

// ---- deployed LoRa coding chain (std_kernels.cu) ---------------------------------------------------------------
static int check_std_cfg(const lora_b200_std_cfg* g)
{
    if (!g) return fail(LORA_B200_EINVAL, "null chain config");
    if (g->sf < 7 || g->sf > 12) return fail(LORA_B200_EINVAL, "sf must be 7..12");
    if (g->cr < 1 || g->cr > 4) return fail(LORA_B200_EINVAL, "cr must be 1..4 (4/5..4/8)");
    if (g->payload_len > 255) return fail(LORA_B200_EINVAL, "payload_len > 255");
    return LORA_B200_OK;
}

extern "C" int lora_b200_std_decode_batch(lora_b200_ctx* c, const lora_b200_std_cfg* g, const uint16_t* d_symbols,
                                          size_t n_packets, size_t nsym, size_t symbol_stride, uint8_t* d_payload,
                                          lora_b200_std_result* d_results, void* stream)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    if (int rc = check_std_cfg(g)) return rc;
    if (!d_symbols || !d_payload || !d_results) return fail(LORA_B200_EINVAL, "null buffer");
    if (n_packets == 0) return LORA_B200_OK;
    if (n_packets > 0x7fffffffu || symbol_stride < nsym || nsym > 0xffff) return fail(LORA_B200_EINVAL, "bad batch geometry");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    StdDecodeArgs a{*g, d_symbols, nsym, symbol_stride, d_payload, d_results, static_cast<int>(n_packets), nullptr, nullptr, 0};
    CU(launch_std_decode(a, static_cast<cudaStream_t>(stream)));
    return LORA_B200_OK;
}

extern "C" int lora_b200_std_encode_batch(lora_b200_ctx* c, const lora_b200_std_cfg* g, const uint8_t* d_payload,
                                          size_t n_packets, size_t payload_len, size_t payload_stride,
                                          uint16_t* d_symbols, size_t symbol_stride, size_t* symbols_per_packet,
                                          void* stream)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    if (int rc = check_std_cfg(g)) return rc;
    if (payload_len > 255) return fail(LORA_B200_EINVAL, "payload longer than 255 bytes");
    const size_t nsym = std_symbol_count(*g, payload_len);
    if (nsym == 0) return fail(LORA_B200_EINVAL, "invalid chain configuration");
    if (symbols_per_packet) *symbols_per_packet = nsym;
    if (!d_symbols) return LORA_B200_OK; // size query
    if (!d_payload && payload_len) return fail(LORA_B200_EINVAL, "null payload");
    if (n_packets == 0) return LORA_B200_OK;
    if (n_packets > 0x7fffffffu || symbol_stride < nsym || payload_stride < payload_len)
        return fail(LORA_B200_EINVAL, "bad batch geometry");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    StdEncodeArgs a{*g, d_payload, payload_len, payload_stride, d_symbols, nsym, symbol_stride, static_cast<int>(n_packets)};
    CU(launch_std_encode(a, static_cast<cudaStream_t>(stream)));
    return LORA_B200_OK;
}

extern "C" int lora_b200_std_stream_decode_batch(lora_b200_ctx* c, const lora_b200_std_cfg* g, const int16_t* d_symbols,
                                                 const lora_b200_stream_packet* d_packets,
                                                 const lora_b200_stream_result* d_results, size_t n_streams,
                                                 uint32_t max_packets, uint32_t max_symbols, uint8_t* d_payload,
                                                 lora_b200_std_result* d_std_results, void* stream)
{
    if (!c || !d_symbols || !d_packets || !d_results || !d_payload || !d_std_results)
        return fail(LORA_B200_EINVAL, "null argument");
    if (int rc = check_std_cfg(g)) return rc;
    if (n_streams == 0 || max_packets == 0) return LORA_B200_OK;
    if (n_streams * static_cast<size_t>(max_packets) > 0x7fffffffu) return fail(LORA_B200_EINVAL, "bad batch geometry");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    StdDecodeArgs a{*g, reinterpret_cast<const uint16_t*>(d_symbols), 0, max_symbols, d_payload, d_std_results,
                    static_cast<int>(n_streams * max_packets), d_packets, d_results, max_packets};
    CU(launch_std_decode(a, static_cast<cudaStream_t>(stream)));
    return LORA_B200_OK;
}

// ---- transmit side --------------------------------------------------------------------------------------
static int modulate_locked(lora_b200_ctx* c, const uint16_t* d_symbols, size_t n_packets, size_t nsym,
                           size_t symbol_stride, float amplitude, uint8_t sync, float2* d_iq,
                           size_t packet_stride, cudaStream_t st)
{
    if (!d_iq || (!d_symbols && nsym)) return fail(LORA_B200_EINVAL, "null buffer");
    if (n_packets == 0) return LORA_B200_OK;
    const size_t step = static_cast<size_t>(c->n) * c->cfg.osr;
    if (n_packets > 0x7fffffffu || packet_stride < (nsym + 2) * step || (nsym && symbol_stride < nsym))
        return fail(LORA_B200_EINVAL, "bad batch geometry");
    ModulateArgs a{};
    a.symbols = d_symbols; a.n_symbols = nsym; a.symbol_stride = symbol_stride;
    a.iq = d_iq; a.packet_stride = packet_stride;
    a.n_packets = static_cast<int>(n_packets);
    a.sf = static_cast<int>(c->cfg.sf); a.osr = static_cast<int>(c->cfg.osr);
    a.bw_scale = c->bw_scale;
    a.amplitude = std::fmax(-1.0f, std::fmin(1.0f, amplitude)); // LoRaMod.cpp:18
    const unsigned shift = c->cfg.sf > 4 ? c->cfg.sf - 4 : 0;   // :20-22
    a.sw0 = static_cast<uint16_t>((sync >> 4) << shift);
    a.sw1 = static_cast<uint16_t>((sync & 0x0f) << shift);
    CU(launch_modulate(a, c->sm_count, st));
    return LORA_B200_OK;
}

extern "C" int lora_b200_modulate_batch(lora_b200_ctx* c, const uint16_t* d_symbols, size_t n_packets,
                                        size_t symbols_per_packet, size_t symbol_stride, float amplitude,
                                        uint8_t sync, lora_b200_c32* d_iq, size_t packet_stride, void* stream)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    return modulate_locked(c, d_symbols, n_packets, symbols_per_packet, symbol_stride, amplitude, sync,
                           reinterpret_cast<float2*>(d_iq), packet_stride, static_cast<cudaStream_t>(stream));
}

extern "C" int lora_b200_awgn_batch(lora_b200_ctx* c, lora_b200_c32* d_iq, size_t n_packets,
                                    size_t samples_per_packet, size_t packet_stride, const float* d_snr_db,
                                    uint64_t seed, void* stream)
{
    if (!c || !d_iq || !d_snr_db) return fail(LORA_B200_EINVAL, "null argument");
    if (n_packets == 0 || samples_per_packet == 0) return LORA_B200_OK;
    if (n_packets > 0x7fffffffu || packet_stride < samples_per_packet) return fail(LORA_B200_EINVAL, "bad batch geometry");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    AwgnArgs a{reinterpret_cast<float2*>(d_iq), samples_per_packet, packet_stride, d_snr_db, seed, static_cast<int>(n_packets)};
    CU(launch_awgn(a, c->sm_count, static_cast<cudaStream_t>(stream)));
    return LORA_B200_OK;
}

// ---- host-buffer entry points -----------------------------------------------------------------------------
extern "C" int lora_b200_demodulate_host(lora_b200_ctx* c, int mode, const lora_b200_c32* h_iq, size_t count,
                                         uint16_t* h_symbols, size_t symbol_cap, size_t* n_symbols,
                                         lora_b200_pkt_metrics* h_metrics)
{
    if (!c || !h_iq || !h_symbols) return fail(LORA_B200_EINVAL, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    const size_t step = static_cast<size_t>(c->n) * c->cfg.osr;
    const size_t total = count / step;
    const size_t n_out = total >= 2 ? total - 2 : total;
    if (mode == LORA_B200_DEMOD_WORKSPACE && n_out > symbol_cap && total >= 2 && count % step == 0)
        return fail(LORA_B200_EINVAL, "symbol buffer too small"); // phy.cpp:190
    if (n_symbols) *n_symbols = 0;
    if (count == 0) return mode == LORA_B200_DEMOD_WORKSPACE ? fail(LORA_B200_EINVAL, "no samples") : LORA_B200_OK;
    CU(c->s_iq.reserve(sizeof(float2) * count));
    CU(c->s_sym.reserve(sizeof(uint16_t) * (n_out + 1)));
    CU(c->s_metrics.reserve(sizeof(lora_b200_pkt_metrics)));
    cudaStream_t st = c->stream;
    CU(cudaMemcpyAsync(c->s_iq.p, h_iq, sizeof(float2) * count, cudaMemcpyHostToDevice, st));
    int rc = demod_batch_locked(c, mode, c->s_iq.as<float2>(), 1, count, count, c->s_sym.as<uint16_t>(), n_out + 1,
                                nullptr, c->s_metrics.as<lora_b200_pkt_metrics>(), st);
    if (rc) return rc;
    lora_b200_pkt_metrics m{};
    CU(cudaMemcpyAsync(&m, c->s_metrics.p, sizeof(m), cudaMemcpyDeviceToHost, st));
    const size_t ncopy = n_out < symbol_cap ? n_out : symbol_cap;
    if (ncopy) CU(cudaMemcpyAsync(h_symbols, c->s_sym.p, sizeof(uint16_t) * ncopy, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (total == 0) { m.cfo = NAN; m.time_offset = NAN; } // 0/0 in LoRaDemod.cpp:125
    if (h_metrics) *h_metrics = m;
    if (n_symbols) *n_symbols = n_out;
    return LORA_B200_OK;
}

extern "C" int lora_b200_estimate_host(lora_b200_ctx* c, const lora_b200_c32* h_iq, size_t count, float* cfo,
                                       float* time_offset)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    if (!h_iq || count == 0) return LORA_B200_OK; // phy.cpp:83: silently returns
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    const size_t step = static_cast<size_t>(c->n) * c->cfg.osr;
    if (count / step == 0) return LORA_B200_OK;
    CU(c->s_iq.reserve(sizeof(float2) * count));
    CU(c->s_metrics.reserve(sizeof(lora_b200_pkt_metrics)));
    cudaStream_t st = c->stream;
    CU(cudaMemcpyAsync(c->s_iq.p, h_iq, sizeof(float2) * count, cudaMemcpyHostToDevice, st));
    int rc = estimate_batch_locked(c, c->s_iq.as<float2>(), 1, count, count, c->s_metrics.as<lora_b200_pkt_metrics>(), st);
    if (rc) return rc;
    lora_b200_pkt_metrics m{};
    CU(cudaMemcpyAsync(&m, c->s_metrics.p, sizeof(m), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (cfo) *cfo = m.cfo;
    if (time_offset) *time_offset = m.time_offset;
    return LORA_B200_OK;
}

extern "C" int lora_b200_modulate_host(lora_b200_ctx* c, const uint16_t* h_symbols, size_t nsym, float amplitude,
                                       uint8_t sync, lora_b200_c32* h_iq, size_t iq_cap, size_t* n_samples)
{
    if (!c || !h_iq || (!h_symbols && nsym)) return fail(LORA_B200_EINVAL, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    const size_t produced = (nsym + 2) * static_cast<size_t>(c->n) * c->cfg.osr;
    if (n_samples) *n_samples = produced;
    if (produced > iq_cap) return fail(LORA_B200_EINVAL, "iq buffer too small"); // phy.cpp:74
    CU(c->s_iq.reserve(sizeof(float2) * produced));
    CU(c->s_sym.reserve(sizeof(uint16_t) * (nsym + 1)));
    cudaStream_t st = c->stream;
    if (nsym) CU(cudaMemcpyAsync(c->s_sym.p, h_symbols, sizeof(uint16_t) * nsym, cudaMemcpyHostToDevice, st));
    int rc = modulate_locked(c, c->s_sym.as<uint16_t>(), 1, nsym, nsym + 1, amplitude, sync, c->s_iq.as<float2>(), produced, st);
    if (rc) return rc;
    CU(cudaMemcpyAsync(h_iq, c->s_iq.p, sizeof(float2) * produced, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return LORA_B200_OK;
}

// ---- host-buffer receive: chunked, double-buffered ingest ----------------------------
// The reference's runner reads the whole capture and then calls once (runners/rx_runner.cpp:74-79).  Here a batch of
// host packets is cut into up to 8 chunks of whole packets (>= 16 MB each): chunk k+1 crosses PCIe on the copy stream
// (one cudaMemcpyAsync per chunk, source best in pinned memory) while the conversion (int16 / int8 ingest), the demod
// kernels and the chain decoder of chunk k run on the compute stream; the records of a chunk travel back as soon as
// it is decoded.  `sample_bytes` = 8 (complex<float>), 4 (int16 I,Q) or 2 (int8 I,Q).
static int receive_host_chunked(lora_b200_ctx* c, int mode, const lora_b200_chain_cfg* g, const void* h_iq, int sample_bytes,
                                float scale, size_t n_packets, size_t count, uint8_t* h_payload,
                                lora_b200_chain_result* h_results, lora_b200_pkt_metrics* h_metrics)
{
    const size_t step = static_cast<size_t>(c->n) * c->cfg.osr;
    if (count / step < 2) return fail(LORA_B200_EINVAL, "fewer than two symbols");
    const size_t nsym = count / step - 2;
    const size_t samples = count * n_packets;
    CU(c->s_iq.reserve(sizeof(float2) * samples));
    if (sample_bytes != 8) CU(c->s_raw.reserve(static_cast<size_t>(sample_bytes) * samples));
    CU(c->s_sym.reserve(sizeof(uint16_t) * nsym * n_packets + 2));
    CU(c->s_metrics.reserve(sizeof(lora_b200_pkt_metrics) * n_packets));
    CU(c->s_payload.reserve(static_cast<size_t>(LORA_B200_PAYLOAD_STRIDE) * n_packets));
    CU(c->s_results.reserve(sizeof(lora_b200_chain_result) * n_packets));
    CU(c->est.reserve(sizeof(float4) * n_packets));          // sized once: the per-chunk calls must not regrow them
    CU(c->sync_raw.reserve(sizeof(uint16_t) * 2 * n_packets));
    if (c->precision == PRECISION_FAST) CU(c->flags.reserve(sizeof(unsigned) * (n_packets * (nsym + 2) + 1)));

    const size_t packet_bytes = count * static_cast<size_t>(sample_bytes);
    size_t n_chunks = (packet_bytes * n_packets) / (size_t(16) << 20);
    n_chunks = std::max<size_t>(1, std::min<size_t>(n_chunks, std::min<size_t>(8, n_packets)));
    const size_t per = (n_packets + n_chunks - 1) / n_chunks;
    n_chunks = (n_packets + per - 1) / per;
    while (c->ev_h2d.size() < n_chunks) {
        cudaEvent_t e = nullptr;
        CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        c->ev_h2d.push_back(e);
    }
    cudaStream_t st = c->stream, cp = c->s_copy;
    // the staging buffers may still be read by work queued earlier on the compute stream
    CU(cudaEventRecord(c->ev_begin, st));
    CU(cudaStreamWaitEvent(cp, c->ev_begin, 0));
    uint8_t* d_stage = sample_bytes == 8 ? reinterpret_cast<uint8_t*>(c->s_iq.p) : reinterpret_cast<uint8_t*>(c->s_raw.p);
    for (size_t k = 0; k < n_chunks; ++k) { // every copy is queued up front; the copy engine runs them back to back
        const size_t p0 = k * per, pn = std::min(per, n_packets - p0);
        CU(cudaMemcpyAsync(d_stage + p0 * packet_bytes, static_cast<const uint8_t*>(h_iq) + p0 * packet_bytes, pn * packet_bytes,
                           cudaMemcpyHostToDevice, cp));
        CU(cudaEventRecord(c->ev_h2d[k], cp));
    }
    for (size_t k = 0; k < n_chunks; ++k) {
        const size_t p0 = k * per, pn = std::min(per, n_packets - p0);
        CU(cudaStreamWaitEvent(st, c->ev_h2d[k], 0));
        float2* d_iq = c->s_iq.as<float2>() + p0 * count;
        if (sample_bytes == 4) CU(launch_convert_i16(c->s_raw.as<short2>() + p0 * count, d_iq, pn * count, scale, c->sm_count, st));
        if (sample_bytes == 2) CU(launch_convert_i8(c->s_raw.as<char2>() + p0 * count, d_iq, pn * count, scale, c->sm_count, st));
        uint16_t* d_sym = c->s_sym.as<uint16_t>() + p0 * nsym;
        lora_b200_pkt_metrics* d_met = c->s_metrics.as<lora_b200_pkt_metrics>() + p0;
        uint8_t* d_pay = c->s_payload.as<uint8_t>() + p0 * LORA_B200_PAYLOAD_STRIDE;
        lora_b200_chain_result* d_res = c->s_results.as<lora_b200_chain_result>() + p0;
        int rc = demod_batch_locked(c, mode, d_iq, pn, count, count, d_sym, nsym, nullptr, d_met, st);
        if (rc) return rc;
        // a dropped packet writes no payload bytes: the caller gets zeros there, not whatever the staging buffer held
        CU(cudaMemsetAsync(d_pay, 0, static_cast<size_t>(LORA_B200_PAYLOAD_STRIDE) * pn, st));
        rc = chain_decode_locked(c, g, d_sym, pn, nsym, nsym, d_pay, d_res, st);
        if (rc) return rc;
        CU(cudaMemcpyAsync(h_payload + p0 * LORA_B200_PAYLOAD_STRIDE, d_pay, static_cast<size_t>(LORA_B200_PAYLOAD_STRIDE) * pn,
                           cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(h_results + p0, d_res, sizeof(lora_b200_chain_result) * pn, cudaMemcpyDeviceToHost, st));
        if (h_metrics) CU(cudaMemcpyAsync(h_metrics + p0, d_met, sizeof(lora_b200_pkt_metrics) * pn, cudaMemcpyDeviceToHost, st));
    }
    CU(cudaStreamSynchronize(st));
    return LORA_B200_OK;
}

extern "C" int lora_b200_receive_host(lora_b200_ctx* c, int mode, const lora_b200_chain_cfg* g,
                                      const lora_b200_c32* h_iq, size_t n_packets, size_t count,
                                      uint8_t* h_payload, lora_b200_chain_result* h_results,
                                      lora_b200_pkt_metrics* h_metrics)
{
    if (!c || !h_iq || !h_payload || !h_results) return fail(LORA_B200_EINVAL, "null argument");
    if (n_packets == 0) return LORA_B200_OK;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    return receive_host_chunked(c, mode, g, h_iq, 8, 1.0f, n_packets, count, h_payload, h_results, h_metrics);
}

extern "C" int lora_b200_receive_host_i16(lora_b200_ctx* c, int mode, const lora_b200_chain_cfg* g,
                                          const int16_t* h_iq, size_t n_packets, size_t count, float scale,
                                          uint8_t* h_payload, lora_b200_chain_result* h_results,
                                          lora_b200_pkt_metrics* h_metrics)
{
    if (!c || !h_iq || !h_payload || !h_results) return fail(LORA_B200_EINVAL, "null argument");
    if (n_packets == 0) return LORA_B200_OK;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    return receive_host_chunked(c, mode, g, h_iq, 4, scale, n_packets, count, h_payload, h_results, h_metrics);
}

extern "C" int lora_b200_receive_host_i8(lora_b200_ctx* c, int mode, const lora_b200_chain_cfg* g,
                                         const int8_t* h_iq, size_t n_packets, size_t count, float scale,
                                         uint8_t* h_payload, lora_b200_chain_result* h_results,
                                         lora_b200_pkt_metrics* h_metrics)
{
    if (!c || !h_iq || !h_payload || !h_results) return fail(LORA_B200_EINVAL, "null argument");
    if (n_packets == 0) return LORA_B200_OK;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    return receive_host_chunked(c, mode, g, h_iq, 2, scale, n_packets, count, h_payload, h_results, h_metrics);
}

// ---- small host-buffer entry points used by include/lora_phy/phy.hpp -------------------------------------------
extern "C" int lora_b200_encode_host(lora_b200_ctx* c, const uint8_t* h_bytes, size_t n_bytes, uint16_t* h_symbols,
                                     size_t symbol_cap, size_t* produced)
{
    if (!c || !h_bytes || !h_symbols) return fail(LORA_B200_EINVAL, "null argument"); // phy.cpp:58
    const size_t n_sym = 2 * n_bytes;
    if (produced) *produced = n_sym;
    if (n_sym > symbol_cap) return fail(LORA_B200_EINVAL, "symbol buffer too small");   // :61
    if (n_bytes == 0) return LORA_B200_OK;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    CU(c->s_misc.reserve(n_bytes));
    CU(c->s_sym.reserve(sizeof(uint16_t) * n_sym));
    cudaStream_t st = c->stream;
    CU(cudaMemcpyAsync(c->s_misc.p, h_bytes, n_bytes, cudaMemcpyHostToDevice, st));
    SimpleEncodeArgs a{c->s_misc.as<uint8_t>(), n_bytes, n_bytes, c->s_sym.as<uint16_t>(), n_sym, 1};
    CU(launch_simple_encode(a, st));
    CU(cudaMemcpyAsync(h_symbols, c->s_sym.p, sizeof(uint16_t) * n_sym, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return LORA_B200_OK;
}

extern "C" int lora_b200_decode_host(lora_b200_ctx* c, const uint16_t* h_symbols, size_t n_symbols, uint8_t* h_bytes,
                                     size_t byte_cap, size_t* produced, int* crc_ok)
{
    if (!c || !h_symbols || !h_bytes) return fail(LORA_B200_EINVAL, "null argument"); // phy.cpp:244
    const size_t n_bytes = n_symbols / 2;
    if (produced) *produced = n_bytes;
    if (crc_ok) *crc_ok = 0;
    if (n_bytes > byte_cap) return fail(LORA_B200_EINVAL, "payload buffer too small");  // :246 (checked before writing)
    if (n_bytes == 0) return LORA_B200_OK;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    CU(c->s_sym.reserve(sizeof(uint16_t) * n_symbols));
    CU(c->s_misc.reserve(n_bytes + 16));
    cudaStream_t st = c->stream;
    CU(cudaMemcpyAsync(c->s_sym.p, h_symbols, sizeof(uint16_t) * n_symbols, cudaMemcpyHostToDevice, st));
    uint8_t* d_bytes = c->s_misc.as<uint8_t>();
    SimpleDecodeArgs a{c->s_sym.as<uint16_t>(), n_symbols, n_symbols, d_bytes, n_bytes, d_bytes + n_bytes, 1};
    CU(launch_simple_decode(a, st));
    uint8_t flag = 0;
    CU(cudaMemcpyAsync(h_bytes, d_bytes, n_bytes, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&flag, d_bytes + n_bytes, 1, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (crc_ok) *crc_ok = flag;
    return LORA_B200_OK;
}

extern "C" int lora_b200_compensate_host(lora_b200_ctx* c, float cfo, float time_offset, lora_b200_c32* h_iq, size_t count)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    if (!h_iq || count == 0) return LORA_B200_OK; // phy.cpp:150 silently returns
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    CU(c->s_iq.reserve(sizeof(float2) * count));
    CU(c->s_payload.reserve(sizeof(float2) * count));
    cudaStream_t st = c->stream;
    CU(cudaMemcpyAsync(c->s_iq.p, h_iq, sizeof(float2) * count, cudaMemcpyHostToDevice, st));
    CompensateArgs a{};
    a.in = c->s_iq.as<float2>();
    a.out = c->s_payload.as<float2>();
    a.count = count;
    // phy.cpp:156: -2*pi*cfo / (N*osr), left to right in float
    volatile float num = -2.0f * 3.14159274101257324f * cfo;
    volatile float den = static_cast<float>(c->n) * static_cast<float>(c->cfg.osr);
    a.rate = num / den;
    const long long off = static_cast<long long>(std::round(time_offset));                    // :163
    a.offset = (off != 0 && static_cast<unsigned long long>(off < 0 ? -off : off) < count) ? off : 0; // :164,169
    CU(launch_compensate(a, st));
    CU(cudaMemcpyAsync(h_iq, a.out, sizeof(float2) * count, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return LORA_B200_OK;
}

// ---- stand-alone FFT and detector (kissfft.hh:100-185, LoRaDetector.hpp:39-74) -----------------------------------
static_assert(sizeof(lora_b200_detect_result) == sizeof(DetectOut), "detector record must match the kernel's");

extern "C" int lora_b200_fft_detect_batch(lora_b200_ctx* c, const lora_b200_c32* d_in, size_t n_windows, int inverse,
                                          lora_b200_c32* d_spectrum, lora_b200_detect_result* d_results, void* stream)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    if (!d_in || (!d_spectrum && !d_results)) return fail(LORA_B200_EINVAL, "null buffer");
    if (inverse && d_results) return fail(LORA_B200_EINVAL, "the detector runs on the forward transform");
    if (n_windows == 0) return LORA_B200_OK;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    FftArgs a{};
    a.in = reinterpret_cast<const float2*>(d_in);
    a.n_windows = n_windows;
    a.inverse = inverse != 0;
    a.tw = c->d_tw;
    a.power_scale = c->power_scale;
    a.out = reinterpret_cast<float2*>(d_spectrum);
    a.det = reinterpret_cast<DetectOut*>(d_results);
    CU(launch_fft_detect(static_cast<int>(c->cfg.sf), a, c->sm_count, static_cast<cudaStream_t>(stream)));
    return LORA_B200_OK;
}

extern "C" int lora_b200_fft_detect_host(lora_b200_ctx* c, const lora_b200_c32* h_in, size_t n_windows, int inverse,
                                         lora_b200_c32* h_spectrum, lora_b200_detect_result* h_results)
{
    if (!c) return fail(LORA_B200_EINVAL, "null context");
    if (!h_in || (!h_spectrum && !h_results)) return fail(LORA_B200_EINVAL, "null buffer");
    if (n_windows == 0) return LORA_B200_OK;
    const size_t bytes = sizeof(float2) * n_windows * static_cast<size_t>(c->n);
    cudaStream_t st;
    float2 *d_in, *d_out;
    DetectOut* d_det;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        CU(cudaSetDevice(c->device));
        CU(c->s_iq.reserve(bytes));
        CU(c->s_payload.reserve(bytes));
        CU(c->s_results.reserve(sizeof(DetectOut) * n_windows));
        st = c->stream;
        d_in = c->s_iq.as<float2>(); d_out = c->s_payload.as<float2>(); d_det = c->s_results.as<DetectOut>();
        CU(cudaMemcpyAsync(d_in, h_in, bytes, cudaMemcpyHostToDevice, st));
    }
    const int rc = lora_b200_fft_detect_batch(c, reinterpret_cast<const lora_b200_c32*>(d_in), n_windows, inverse,
                                              h_spectrum ? reinterpret_cast<lora_b200_c32*>(d_out) : nullptr,
                                              h_results ? reinterpret_cast<lora_b200_detect_result*>(d_det) : nullptr, st);
    if (rc != LORA_B200_OK) return rc;
    if (h_spectrum) CU(cudaMemcpyAsync(h_spectrum, d_out, bytes, cudaMemcpyDeviceToHost, st));
    if (h_results) CU(cudaMemcpyAsync(h_results, d_det, sizeof(DetectOut) * n_windows, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return LORA_B200_OK;
}

// ---- continuous-stream front end and full frames ---------------------------------------------------------------
static_assert(sizeof(lora_b200_stream_packet) == sizeof(StreamPacket) && sizeof(lora_b200_stream_result) == sizeof(StreamResult),
              "stream records must match the kernel's");

extern "C" size_t lora_b200_frame_samples(const lora_b200_ctx* c, size_t n_symbols, unsigned padding)
{
    if (!c) return 0;
    return frame_samples(static_cast<int>(c->cfg.sf), static_cast<int>(c->cfg.osr), n_symbols, padding);
}

extern "C" int lora_b200_frame_modulate_batch(lora_b200_ctx* c, const uint16_t* d_symbols, size_t n_packets,
                                              size_t n_symbols, size_t symbol_stride, float amplitude, uint8_t sync,
                                              unsigned padding, lora_b200_c32* d_iq, size_t packet_stride, void* stream)
{
    if (!c || !d_iq || (!d_symbols && n_symbols)) return fail(LORA_B200_EINVAL, "null argument");
    if (n_packets == 0) return LORA_B200_OK;
    const size_t need = frame_samples(static_cast<int>(c->cfg.sf), static_cast<int>(c->cfg.osr), n_symbols, padding);
    if (n_packets > 0x7fffffffu || packet_stride < need || symbol_stride < n_symbols || n_symbols > 0x7fffff00u)
        return fail(LORA_B200_EINVAL, "bad batch geometry");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    FrameModArgs a{d_symbols, n_symbols, symbol_stride, reinterpret_cast<float2*>(d_iq), packet_stride,
                   static_cast<int>(n_packets), static_cast<int>(c->cfg.sf), static_cast<int>(c->cfg.osr), amplitude, sync, padding};
    CU(launch_frame_modulate(a, c->sm_count, static_cast<cudaStream_t>(stream)));
    return LORA_B200_OK;
}

static int stream_detect_locked(lora_b200_ctx* c, const lora_b200_stream_cfg* cfg, const lora_b200_c32* d_iq,
                                size_t n_streams, size_t samples_per_stream, size_t stream_stride, int16_t* d_symbols,
                                lora_b200_stream_packet* d_packets, lora_b200_stream_result* d_results, cudaStream_t st)
{
    if (!cfg || !d_results) return fail(LORA_B200_EINVAL, "null argument");
    if (n_streams == 0) return LORA_B200_OK;
    if (!d_iq || (!d_symbols && cfg->max_symbols) || (!d_packets && cfg->max_packets)) return fail(LORA_B200_EINVAL, "null argument");
    if (!blocked_supported(static_cast<int>(c->cfg.sf)))
        return fail(LORA_B200_EINVAL, "stream detection needs sf 7..12");
    if (n_streams > 0x7fffffffu || stream_stride < samples_per_stream) return fail(LORA_B200_EINVAL, "bad batch geometry");
    const size_t n = static_cast<size_t>(c->n);
    if (!c->d_stream_tables) {
        std::vector<float2> up, down, fine;
        build_stream_tables(static_cast<int>(c->cfg.sf), up, down, fine);
        std::vector<float2> all(up);
        all.insert(all.end(), down.begin(), down.end());
        all.insert(all.end(), fine.begin(), fine.end());
        float2* tables = nullptr;
        CU(cudaMalloc(&tables, sizeof(float2) * all.size()));
        const cudaError_t up_rc = cudaMemcpy(tables, all.data(), sizeof(float2) * all.size(), cudaMemcpyHostToDevice);
        if (up_rc != cudaSuccess) { // never leave a non-null, half-written table behind
            cudaFree(tables);
            return fail(LORA_B200_ECUDA, "stream table upload", up_rc);
        }
        c->d_stream_tables = tables;
    }
    StreamDetectArgs a{};
    a.iq = reinterpret_cast<const float2*>(d_iq);
    a.stream_stride = stream_stride;
    a.count = samples_per_stream;
    a.osr = c->cfg.osr; // LoRaDemod consumes N samples per symbol: an oversampled stream is decimated on the fly
    a.n_streams = static_cast<int>(n_streams);
    a.twp = reinterpret_cast<const c64*>(c->d_twp);
    a.tw0 = c->tw0;
    a.one = 1.0f;
    a.up = reinterpret_cast<const c64*>(c->d_stream_tables);
    a.down = a.up + n;
    a.fine = a.up + 2 * n;
    a.power_scale = c->power_scale;
    a.thresh = cfg->thresh_db;
    a.mtu = cfg->mtu;
    a.sync = cfg->sync;
    a.frame_header = static_cast<int>(cfg->frame_header);
    a.symbols = d_symbols;
    a.sym_cap = cfg->max_symbols;
    a.packets = reinterpret_cast<StreamPacket*>(d_packets);
    a.pkt_cap = cfg->max_packets;
    a.results = reinterpret_cast<StreamResult*>(d_results);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (c->timing) {
        if (c->ev_used == c->ev_pool.size()) {
            cudaEvent_t ea = nullptr, eb = nullptr;
            CU(cudaEventCreate(&ea));
            CU(cudaEventCreate(&eb));
            c->ev_pool.emplace_back(ea, eb);
        }
        e0 = c->ev_pool[c->ev_used].first;
        e1 = c->ev_pool[c->ev_used].second;
        ++c->ev_used;
        CU(cudaEventRecord(e0, st));
    }
    CU(launch_stream_detect(static_cast<int>(c->cfg.sf), a, c->sm_count, st));
    if (e1) CU(cudaEventRecord(e1, st));
    return LORA_B200_OK;
}

extern "C" int lora_b200_stream_detect_batch(lora_b200_ctx* c, const lora_b200_stream_cfg* cfg, const lora_b200_c32* d_iq,
                                             size_t n_streams, size_t samples_per_stream, size_t stream_stride,
                                             int16_t* d_symbols, lora_b200_stream_packet* d_packets,
                                             lora_b200_stream_result* d_results, void* stream)
{
    if (!c) return fail(LORA_B200_EINVAL, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    return stream_detect_locked(c, cfg, d_iq, n_streams, samples_per_stream, stream_stride, d_symbols, d_packets, d_results,
                                static_cast<cudaStream_t>(stream));
}

// One capture, host to host: copy in, frame synchroniser, deployed coding chain, copy out.
extern "C" int lora_b200_std_receive_stream_host(lora_b200_ctx* c, const lora_b200_std_cfg* g,
                                                 const lora_b200_stream_cfg* cfg, const lora_b200_c32* h_iq, size_t count,
                                                 uint8_t* h_payload, lora_b200_std_result* h_results,
                                                 lora_b200_stream_packet* h_packets, uint32_t* n_packets)
{
    if (!c || !cfg || !h_iq || !h_payload || !h_results || !n_packets) return fail(LORA_B200_EINVAL, "null argument");
    if (int rc = check_std_cfg(g)) return rc;
    if (cfg->max_packets == 0 || cfg->max_symbols == 0) return fail(LORA_B200_EINVAL, "no room for packets");
    *n_packets = 0;
    if (count == 0) return LORA_B200_OK;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const size_t np = cfg->max_packets;
    CU(c->s_iq.reserve(sizeof(float2) * count));
    CU(c->s_sym.reserve(sizeof(int16_t) * cfg->max_symbols));
    CU(c->s_misc.reserve(sizeof(lora_b200_stream_packet) * np + sizeof(lora_b200_stream_result)));
    CU(c->s_payload.reserve(static_cast<size_t>(LORA_B200_PAYLOAD_STRIDE) * np));
    CU(c->s_results.reserve(sizeof(lora_b200_std_result) * np));
    lora_b200_stream_result* d_res = c->s_misc.as<lora_b200_stream_result>();
    lora_b200_stream_packet* d_pk = reinterpret_cast<lora_b200_stream_packet*>(d_res + 1);
    CU(cudaMemcpyAsync(c->s_iq.p, h_iq, sizeof(float2) * count, cudaMemcpyHostToDevice, st));
    lora_b200_stream_cfg sc = *cfg;
    StdDecodeArgs a{*g, c->s_sym.as<uint16_t>(), 0, cfg->max_symbols, c->s_payload.as<uint8_t>(),
                    c->s_results.as<lora_b200_std_result>(), static_cast<int>(np), d_pk, d_res, cfg->max_packets};
    if (sc.mtu == 0) {
        // The synchroniser (LoRaDemod.cpp:286-310) collects data symbols until `mtu` or a squelched window; it does not
        // read the header.  mtu = 0 asks for every frame's own length: implicit-header frames have a known one, explicit
        // headers are read by the synchroniser itself after the eighth data symbol (frame_header mode).
        if (g->implicit_header) {
            sc.mtu = static_cast<uint32_t>(std_symbol_count(*g, g->payload_len));
        } else {
            lora_b200_std_cfg longest = *g;
            longest.cr = 4; longest.has_crc = 1;
            sc.mtu = static_cast<uint32_t>(std_symbol_count(longest, 255));
            sc.frame_header = g->ldro ? 2u : 1u;
        }
    }
    if (int rc = stream_detect_locked(c, &sc, c->s_iq.as<lora_b200_c32>(), 1, count, count, c->s_sym.as<int16_t>(), d_pk,
                                      d_res, st))
        return rc;
    CU(cudaMemsetAsync(c->s_payload.p, 0, static_cast<size_t>(LORA_B200_PAYLOAD_STRIDE) * np, st)); // bytes no frame writes read as zero
    CU(launch_std_decode(a, st));
    lora_b200_stream_result res{};
    CU(cudaMemcpyAsync(&res, d_res, sizeof(res), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(h_payload, c->s_payload.p, static_cast<size_t>(LORA_B200_PAYLOAD_STRIDE) * np, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(h_results, c->s_results.p, sizeof(lora_b200_std_result) * np, cudaMemcpyDeviceToHost, st));
    if (h_packets) CU(cudaMemcpyAsync(h_packets, d_pk, sizeof(lora_b200_stream_packet) * np, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    *n_packets = res.n_packets < cfg->max_packets ? res.n_packets : cfg->max_packets;
    return LORA_B200_OK;
}

extern "C" int lora_b200_stream_decode_batch(lora_b200_ctx* c, const lora_b200_chain_cfg* g, const int16_t* d_symbols,
                                             const lora_b200_stream_packet* d_packets,
                                             const lora_b200_stream_result* d_results, size_t n_streams,
                                             uint32_t max_packets, uint32_t max_symbols, uint8_t* d_payload,
                                             lora_b200_chain_result* d_chain_results, void* stream)
{
    if (!c || !d_symbols || !d_packets || !d_results || !d_payload || !d_chain_results)
        return fail(LORA_B200_EINVAL, "null argument");
    if (int rc = check_chain_cfg(g)) return rc;
    if (n_streams == 0 || max_packets == 0) return LORA_B200_OK;
    if (n_streams * static_cast<size_t>(max_packets) > 0x7fffffffu) return fail(LORA_B200_EINVAL, "bad batch geometry");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    ChainDecodeArgs a{*g, reinterpret_cast<const uint16_t*>(d_symbols), 0, max_symbols, d_payload, d_chain_results,
                      c->d_whiten, static_cast<int>(n_streams * max_packets), d_packets, d_results, max_packets};
    CU(launch_chain_decode(a, static_cast<cudaStream_t>(stream)));
    return LORA_B200_OK;
}
