/* lora_phy_b200.h -- C ABI of the B200 LoRa receive path (liblora_phy_b200.so).
 *
 * This is the drop-in boundary.  Plain pointers and sizes, no C++ or torch types.  The reference has no
 * FFI of its own (it is a C++ static library, CMakeLists.txt:12-21); each entry point below names the
 * reference function it replaces, and include/lora_phy/phy.hpp wraps them back into the reference's C++
 * names so runners/rx_runner.cpp and tx_runner.cpp build unchanged.
 *
 * Conventions
 *   - Every function returns LORA_B200_OK (0) or a negative code; lora_b200_last_error() gives text.
 *   - `d_` pointers are device pointers on the context's GPU, `h_` pointers are host pointers.
 *   - Device entry points are stream-ordered on `stream` (a cudaStream_t passed as void*; NULL = the
 *     legacy default stream) and never synchronise or allocate after the first call at a given size.
 *   - ONE stream in flight per context: a context owns per-call scratch (offset estimates, sync symbols, the recheck
 *     list, staging buffers), so calls on the same context must be issued on one stream, or ordered by the caller
 *     (events) when they use several.  Independent work belongs on independent contexts -- they are cheap
 *     (32-100 KB of tables) and this is how the gateway runs one context per spreading factor.  Host-side entry is
 *     serialised by a mutex per context; that orders enqueueing, not execution.
 *   - IQ is interleaved little-endian float32 (re, im) = std::complex<float>; symbols are uint16_t.
 *   - There is no CPU fallback anywhere: with no usable CUDA device every call fails with
 *     LORA_B200_ECUDA.
 */
#ifndef LORA_PHY_B200_H
#define LORA_PHY_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LORA_B200_OK 0
#define LORA_B200_EINVAL (-1) /* bad argument (the reference returns -1 for these too, phy.cpp:181-190) */
#define LORA_B200_ECUDA (-2)  /* CUDA runtime error or no device */
#define LORA_B200_ENOMEM (-3)

typedef struct lora_b200_ctx lora_b200_ctx;
typedef struct { float re, im; } lora_b200_c32;

/* Which reference entry point's semantics demod follows. */
enum {
    /* lora_phy::demodulate, src/phy/phy.cpp:178-239: offsets estimated on the RAW first two symbols,
     * every window multiplied by the down-chirp table, then by the CFO rotation. */
    LORA_B200_DEMOD_WORKSPACE = 0,
    /* lora_phy::lora_demodulate, src/phy/LoRaDemod.cpp:49-195: input already dechirped by the caller;
     * max-abs normalisation when any |re| or |im| exceeds 1, offsets estimated on the normalised data. */
    LORA_B200_DEMOD_LEGACY = 1,
    /* The pipeline of tests/performance_test.cpp:103-116 and tests/e2e_chain_test.cpp:79-104: caller-side
     * dechirp by the genChirp down table, then lora_demodulate -- with the dechirped copy never written
     * to memory.  Bit-identical to running those two steps on the CPU. */
    LORA_B200_DEMOD_LEGACY_DECHIRP = 2
};

typedef struct {
    unsigned sf;     /* 1..12 (N = 1 << sf <= 4096, include/lora_phy/kissfft.hh:34) */
    unsigned bw_hz;  /* 125000, 250000 or 500000 (phy.hpp:36-40)                      */
    unsigned osr;    /* oversampling ratio, 0 means 1                                 */
    int window;      /* 0 = none, 1 = Hann (phy.hpp:29-32)                            */
    int device;      /* CUDA device ordinal, -1 = current device                      */
} lora_b200_config;

/* Per-packet result of demod, 16 bytes. */
typedef struct {
    float cfo;           /* lora_metrics::cfo (phy.hpp:65-69)                                          */
    float time_offset;   /* lora_metrics::time_offset                                                  */
    float max_amp;       /* legacy modes: max(|re|,|im|) over the packet (LoRaDemod.cpp:61-67), else 0 */
    uint8_t sync_word;   /* high nibbles of the first two windows (phy.cpp:235-237)                    */
    uint8_t normalised;  /* legacy modes: 1 when max_amp > 1 and samples were scaled                   */
    uint16_t n_symbols;  /* symbols written for this packet                                            */
} lora_b200_pkt_metrics;

/* Configuration of the full coding chain; mirrors the setters of the Pothos blocks
 * (LoRa-SDR/LoRaDecoder.cpp:131-189, LoRa-SDR/LoRaEncoder.cpp:99-130). */
typedef struct {
    unsigned sf;
    unsigned ppm;          /* 0 = full symbol set; sf-2 for low-data-rate optimisation */
    unsigned rdd;          /* 0..4 = coding rate 4/4 .. 4/8                             */
    int explicit_header;
    int crc;               /* encoder: append CRC; decoder: drop on CRC mismatch (crcc) */
    int whitening;         /* encoder only                                              */
    int interleaving;      /* decoder only                                              */
    int error_check;       /* decoder only                                              */
    int hdr_out;           /* decoder only                                              */
    unsigned data_length;  /* decoder, implicit header mode                             */
} lora_b200_chain_cfg;

/* Per-packet result of the chain decoder, 16 bytes. */
typedef struct {
    int16_t n_bytes;       /* bytes emitted; -1 dropped, -2 no output (< 8 symbols), -3 reference-undefined */
    int16_t drop_line;     /* 0, or the LoRaDecoder.cpp line of the drop() taken                            */
    uint8_t parity_error;  /* `error` flag   */
    uint8_t uncorrectable; /* `bad` flag     */
    uint8_t hdr_length, hdr_has_crc, hdr_rdd, hdr_checksum;
    uint8_t crc_present, crc_ok;
    uint8_t reserved[4];
} lora_b200_chain_result;

#define LORA_B200_PAYLOAD_STRIDE 264 /* bytes reserved per packet by the chain decoder (<= 255+3+2, padded) */

/* The LoRa coding chain as it is deployed on the air (not in the reference, whose chain is the
 * SX127x-style variant of the Pothos blocks above): symbol = bin - 1, first eight symbols at reduced rate (sf-2 bits,
 * CR 4/8) carrying the explicit header, diagonal interleaver, Hamming 4/5..4/8, whitening LFSR x^8+x^6+x^5+x^4+1,
 * CRC-16/CCITT with the last two payload bytes folded in.  It decodes the standards-style frames of the reference's
 * golden capture (vectors_binary/bw_125k_sf_7_cr_1_...unknown), which rx_runner cannot. */
typedef struct {
    unsigned sf;            /* 7..12                                                                    */
    unsigned cr;            /* 1..4 = 4/5..4/8; an explicit header overrides it on receive               */
    int ldro;               /* low-data-rate optimisation: every block carries sf-2 bits per symbol      */
    int implicit_header;
    int has_crc;            /* encoder; decoder in implicit-header mode                                  */
    unsigned payload_len;   /* decoder, implicit-header mode                                             */
} lora_b200_std_cfg;

/* Per-packet result of the deployed-chain decoder, 16 bytes. */
typedef struct {
    int16_t n_bytes;        /* payload bytes; -1 header invalid (or no such packet), -2 fewer than 8 symbols, -3 truncated */
    uint16_t symbols_used;  /* symbols of the frame actually consumed                                   */
    uint8_t hdr_ok, hdr_length, hdr_cr, has_crc, crc_ok;
    uint8_t corrected;      /* some codeword had one bit corrected                                      */
    uint8_t uncorrectable;  /* some codeword had an error that could only be detected                   */
    uint8_t reserved[5];
} lora_b200_std_result;

/* ---- lifetime ------------------------------------------------------------------------------------------ */
/* Replaces lora_phy::init (phy.cpp:26-49) / lora_demod_init (LoRaDemod.cpp:10-32): builds the twiddle,
 * down-chirp and window tables on the host exactly as the reference does and uploads them. */
int lora_b200_ctx_create(const lora_b200_config* cfg, lora_b200_ctx** out);
void lora_b200_ctx_destroy(lora_b200_ctx* ctx);
const char* lora_b200_last_error(void);
/* Number of kernels this library has launched in this process (for bench.py's gpu_launches). */
unsigned long long lora_b200_launch_count(void);
/* Kernel timing for bench.py's roofline: when enabled, every launch of the fused dechirp+FFT+argmax kernel
 * is bracketed by CUDA events on the stream it is launched on.  lora_b200_kernel_time synchronises those
 * events, returns their summed duration and count since the last call, and resets the counters. */
int lora_b200_set_kernel_timing(lora_b200_ctx* ctx, int enable);
int lora_b200_kernel_time(lora_b200_ctx* ctx, double* total_ms, unsigned long long* launches);
/* How lora_b200_demod_batch resolves windows (SF7..SF12; smaller SF always use the exact kernel).
 *   FAST  (default) every window goes through the FMA-fused kernel; a window whose two largest |X[k]|^2 are
 *         closer than the kernel's error bound (relative margin 2^-13, widened with the CFO phase) is recomputed
 *         by the bit-faithful kernel in a second launch.  Symbol indices are bit-identical to the reference;
 *         d_peaks is within 1e-4 relative of the reference's float32 value (BASELINE.json's tolerance).
 *   EXACT every window goes through the bit-faithful kernel: d_peaks is bit-identical as well. */
#define LORA_B200_PRECISION_FAST 0
#define LORA_B200_PRECISION_EXACT 1
int lora_b200_set_precision(lora_b200_ctx* ctx, int precision);
/* How the normalisation scan + offset estimate ("prepass") and the fused demod of lora_b200_demod_batch are scheduled;
 * results never depend on it.
 *   0  two launches back to back on the caller's stream
 *   1  large legacy batches (>= 256 MB of IQ) cut into 8 sub-batches, the scan of sub-batch k+1 on an internal stream
 *      beside the demod kernel of sub-batch k (measured slower, kept as an option)
 *   2  (default) one persistent launch that takes prepass items and demod items from one queue, so every SM scans some
 *      packets while it transforms others -- for calls of >= 600 MB of samples (and >= 256 packets, SF7..SF12, FAST
 *      precision), where it is measured faster; smaller calls run as 0
 *   3  as 2 without the size threshold (tests) */
int lora_b200_set_overlap(lora_b200_ctx* ctx, int enable);
/* Windows demodulated in FAST precision, and how many of them were rechecked, since the last call
 * (synchronises the device; for reporting only). */
int lora_b200_recheck_stats(lora_b200_ctx* ctx, unsigned long long* windows, unsigned long long* rechecked);
/* Copies of the host-built tables, for tests: table 0 twiddles, 1 down-chirp (n complex), 2 window (n floats). */
int lora_b200_ctx_table(const lora_b200_ctx* ctx, int table, void* h_out, size_t n);

/* ---- receive path, device buffers ------------------------------------------------------------------------- */
/* demodulate / lora_demodulate over a batch of equal-length packets.
 *   d_iq              n_packets packets, packet p starting at d_iq + p*packet_stride
 *   samples_per_packet the `sample_count` each reference call would see
 *   d_symbols         packet p's symbols at d_symbols + p*symbol_stride: total-2 of them (the two sync
 *                     windows are stripped, phy.cpp:226-233), or all when total < 2 in legacy mode
 *   d_peaks           optional, n_packets * total floats: max |X[k]|^2 of every window incl. sync
 *   d_metrics         optional, one record per packet
 * WORKSPACE mode requires samples_per_packet % (N*osr) == 0 and >= 2 symbols, as phy.cpp:186-190 does. */
int lora_b200_demod_batch(lora_b200_ctx* ctx, int mode, const lora_b200_c32* d_iq, size_t n_packets,
                          size_t samples_per_packet, size_t packet_stride, uint16_t* d_symbols,
                          size_t symbol_stride, float* d_peaks, lora_b200_pkt_metrics* d_metrics,
                          void* stream);

/* lora_b200_demod_batch for callers that cannot allocate (code under CUDA stream capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same results bit for bit, SF7..SF12 only.  The per-call scratch of the receive kernels (offset
 * estimates, raw sync symbols, the recheck list) lives in a fixed device array inside the library -- about 4.3 MB that
 * every GPU which loads the library holds until the process exits -- so the call issues no cudaMalloc / cudaFree and
 * no host allocation; it only enqueues kernels and one 4-byte memset on `stream`.  Limits, checked with
 * LORA_B200_EINVAL and a message in lora_b200_last_error(): at most LORA_B200_STATIC_MAX_PACKETS packets and
 * LORA_B200_STATIC_MAX_WINDOWS symbol windows (n_packets * (samples_per_packet / (N*osr))) per call.
 * All calls on one device share the array: calls that overlap on the GPU (different streams, contexts or host
 * threads) race on it -- order them on one stream or with events.  It keeps the last call's contents; the word that
 * must start at zero is cleared on `stream` by the call itself.  Kernel timing (lora_b200_set_kernel_timing) and the
 * overlapped scan (lora_b200_set_overlap) do not apply to this entry point. */
#define LORA_B200_STATIC_MAX_PACKETS 16384
#define LORA_B200_STATIC_MAX_WINDOWS 1048576
int lora_b200_demod_batch_static(lora_b200_ctx* ctx, int mode, const lora_b200_c32* d_iq, size_t n_packets,
                                 size_t samples_per_packet, size_t packet_stride, uint16_t* d_symbols,
                                 size_t symbol_stride, float* d_peaks, lora_b200_pkt_metrics* d_metrics,
                                 void* stream);

/* lora_phy::estimate_offsets (phy.cpp:78-145) over a batch; writes cfo/time_offset of d_metrics. */
int lora_b200_estimate_batch(lora_b200_ctx* ctx, const lora_b200_c32* d_iq, size_t n_packets,
                             size_t samples_per_packet, size_t packet_stride,
                             lora_b200_pkt_metrics* d_metrics, void* stream);

/* lora_phy::decode / lora_decode (phy.cpp:241-256, src/phy/LoRaDecoder.cpp:6-19): Hamming(8,4) pairs to
 * bytes plus the CRC flag.  Packet p: symbols at d_symbols + p*symbol_stride, bytes at d_bytes + p*byte_stride. */
int lora_b200_decode_batch(lora_b200_ctx* ctx, const uint16_t* d_symbols, size_t n_packets,
                           size_t symbols_per_packet, size_t symbol_stride, uint8_t* d_bytes,
                           size_t byte_stride, uint8_t* d_crc_ok, void* stream);

/* LoRaDecoder::work (LoRa-SDR/LoRaDecoder.cpp:196-397): Gray, diagonal deinterleave, dewhiten, Hamming /
 * parity, header, CRC.  Payload of packet p at d_payload + p*LORA_B200_PAYLOAD_STRIDE. */
int lora_b200_chain_decode_batch(lora_b200_ctx* ctx, const lora_b200_chain_cfg* cfg,
                                 const uint16_t* d_symbols, size_t n_packets, size_t symbols_per_packet,
                                 size_t symbol_stride, uint8_t* d_payload,
                                 lora_b200_chain_result* d_results, void* stream);

/* ---- transmit side (used to synthesise inputs, and by tx_runner) ------------------------------------------- */
/* lora_encode (src/phy/LoRaEncoder.cpp:6-18). */
int lora_b200_encode_batch(lora_b200_ctx* ctx, const uint8_t* d_bytes, size_t n_packets,
                           size_t bytes_per_packet, size_t byte_stride, uint16_t* d_symbols,
                           size_t symbol_stride, void* stream);
/* LoRaEncoder::work (LoRa-SDR/LoRaEncoder.cpp:161-233); every packet has payload_len bytes.
 * Returns the symbol count per packet through *symbols_per_packet (may be queried with d_symbols NULL). */
int lora_b200_chain_encode_batch(lora_b200_ctx* ctx, const lora_b200_chain_cfg* cfg,
                                 const uint8_t* d_payload, size_t n_packets, size_t payload_len,
                                 size_t payload_stride, uint16_t* d_symbols, size_t symbol_stride,
                                 size_t* symbols_per_packet, void* stream);
/* lora_modulate (src/phy/LoRaMod.cpp:8-43): [sync0][sync1][symbols...] chirps with the reference's float32
 * phase-accumulator recurrence.  Packet p's (symbols_per_packet+2)*N*osr samples at d_iq + p*packet_stride. */
int lora_b200_modulate_batch(lora_b200_ctx* ctx, const uint16_t* d_symbols, size_t n_packets,
                             size_t symbols_per_packet, size_t symbol_stride, float amplitude,
                             uint8_t sync, lora_b200_c32* d_iq, size_t packet_stride, void* stream);
/* Adds complex white Gaussian noise in place: per-component sigma/sqrt(2) with sigma = 10^(-snr_db/20)
 * (tests/awgn_sweep_gtest.cpp:76-80).  d_snr_db holds one SNR per packet.  Counter-based generator, so the
 * result depends only on (seed, packet, sample). */
int lora_b200_awgn_batch(lora_b200_ctx* ctx, lora_b200_c32* d_iq, size_t n_packets,
                         size_t samples_per_packet, size_t packet_stride, const float* d_snr_db,
                         uint64_t seed, void* stream);

/* ---- host-buffer entry points (what include/lora_phy/phy.hpp calls) ------------------------------------------ */
/* One packet, host pointers; copies in, runs the kernels above, copies out, synchronises. */
int lora_b200_demodulate_host(lora_b200_ctx* ctx, int mode, const lora_b200_c32* h_iq, size_t sample_count,
                              uint16_t* h_symbols, size_t symbol_cap, size_t* n_symbols,
                              lora_b200_pkt_metrics* h_metrics);
int lora_b200_estimate_host(lora_b200_ctx* ctx, const lora_b200_c32* h_iq, size_t sample_count,
                            float* cfo, float* time_offset);
int lora_b200_modulate_host(lora_b200_ctx* ctx, const uint16_t* h_symbols, size_t symbol_count,
                            float amplitude, uint8_t sync, lora_b200_c32* h_iq, size_t iq_cap,
                            size_t* n_samples);
/* lora_phy::encode / decode / compensate_offsets with host buffers (phy.cpp:55-63, 241-256, 147-176).  *produced is
 * always set to the count the call yields (2*n_bytes, n_symbols/2); a too-small output buffer fails before anything
 * is written.  compensate works in place on h_iq with the cfo / time_offset the caller holds in its metrics. */
int lora_b200_encode_host(lora_b200_ctx* ctx, const uint8_t* h_bytes, size_t n_bytes, uint16_t* h_symbols,
                          size_t symbol_cap, size_t* produced);
int lora_b200_decode_host(lora_b200_ctx* ctx, const uint16_t* h_symbols, size_t n_symbols, uint8_t* h_bytes,
                          size_t byte_cap, size_t* produced, int* crc_ok);
int lora_b200_compensate_host(lora_b200_ctx* ctx, float cfo, float time_offset, lora_b200_c32* h_iq, size_t count);
/* ---- stand-alone FFT and detector ------------------------------------------------------------------------------------
 * kissfft<float>::transform (include/lora_phy/kissfft.hh:100-185) and LoRaDetector<float>::detect
 * (include/lora_phy/LoRaDetector.hpp:39-74) for n_windows windows of N = 2^sf samples each, bit for bit the CPU's:
 * the spectrum in natural bin order (inverse != 0: the plan built with inverse = true, kissfft.hh:24-29,176-181), and
 * per window the first-strict-maximum bin of |X|^2, the parabolic fIndex (both bit for bit), and power / powerAvg in
 * dB minus 20*log10(N) (through the device's log10f: within 3e-5 dB of glibc's; they only feed comparisons).  These are what include/lora_phy/kissfft.hh and LoRaDetector.hpp of this build forward to. */
typedef struct {
    uint32_t index;      /* LoRaDetector::detect return value */
    float power;         /* :64 */
    float power_avg;     /* :63 */
    float f_index;       /* :66-71 */
} lora_b200_detect_result;
int lora_b200_fft_detect_batch(lora_b200_ctx* ctx, const lora_b200_c32* d_in, size_t n_windows, int inverse,
                               lora_b200_c32* d_spectrum /* nullable */, lora_b200_detect_result* d_results /* nullable */,
                               void* stream);
int lora_b200_fft_detect_host(lora_b200_ctx* ctx, const lora_b200_c32* h_in, size_t n_windows, int inverse,
                              lora_b200_c32* h_spectrum /* nullable */, lora_b200_detect_result* h_results /* nullable */);

/* Batched host-buffer demod + chain decode: the end-to-end call bench.py times (pinned or pageable host
 * memory in, payloads and results out).  Payload bytes the decoder does not write (dropped packets, bytes past
 * n_bytes) come back as zero. */
int lora_b200_receive_host(lora_b200_ctx* ctx, int mode, const lora_b200_chain_cfg* cfg,
                           const lora_b200_c32* h_iq, size_t n_packets, size_t samples_per_packet,
                           uint8_t* h_payload, lora_b200_chain_result* h_results,
                           lora_b200_pkt_metrics* h_metrics);
/* The same call for SDR sample formats: interleaved int16 I,Q as most front ends
 * deliver them, 4 bytes per sample over PCIe instead of 8.  Samples are converted on the device to
 * float32 * scale (exact for scale = 2^-15) and then take exactly the path of lora_b200_receive_host, so the results
 * equal those of the float call on the converted samples.  The runners' float32 file format
 * (runners/rx_runner.cpp:74-79) is unchanged. */
int lora_b200_receive_host_i16(lora_b200_ctx* ctx, int mode, const lora_b200_chain_cfg* cfg,
                               const int16_t* h_iq, size_t n_packets, size_t samples_per_packet, float scale,
                               uint8_t* h_payload, lora_b200_chain_result* h_results,
                               lora_b200_pkt_metrics* h_metrics);
/* The same for interleaved int8 I/Q (RTL-SDR class front ends): a quarter of the PCIe bytes of the float call. */
int lora_b200_receive_host_i8(lora_b200_ctx* ctx, int mode, const lora_b200_chain_cfg* cfg,
                              const int8_t* h_iq, size_t n_packets, size_t samples_per_packet, float scale,
                              uint8_t* h_payload, lora_b200_chain_result* h_results,
                              lora_b200_pkt_metrics* h_metrics);

/* ---- continuous-stream front end and full frames (upstream Pothos blocks) --------------------------------------
 * LoRaDemod::work (LoRa-SDR/LoRaDemod.cpp:145-327): the frame synchroniser state machine -- 10-chirp preamble with
 * fractional-bin fine tuning, two sync-word chirps, two down-chirps for the frequency error, the quarter chirp, then
 * data symbols until `mtu` symbols or a squelched window -- run over n_streams independent sample streams (gateway
 * channels, or segments of a long capture), one stream per group of N/16 threads.  sf 7..12.  With osr > 1 the
 * stream is decimated on the fly (samples 0, osr, 2*osr, ... -- the indexing of phy.cpp:216); positions and
 * `consumed` are then counted in decimated samples.
 * LoRaMod::work (LoRa-SDR/LoRaMod.cpp:109-238): the frame generator that produces such streams. */
typedef struct lora_b200_stream_cfg {
    uint8_t sync;          /* LoRaDemod::setSync, default 0x12 */
    float thresh_db;       /* LoRaDemod::setThreshold, default -30 */
    uint32_t mtu;          /* LoRaDemod::setMTU, default 256 */
    uint32_t max_packets;  /* capacity of d_packets per stream */
    uint32_t max_symbols;  /* capacity of d_symbols per stream */
    uint32_t frame_header; /* 0 = as LoRaDemod: data symbols until mtu or a squelched window (:286-310).  Not in the
                              reference: 1 = after eight data symbols read the deployed chain's explicit header and stop
                              at the frame length it announces (capped by mtu; an invalid header ends the frame at
                              eight symbols); 2 = the same for frames sent with low-data-rate optimisation */
} lora_b200_stream_cfg;

typedef struct lora_b200_stream_packet {
    uint64_t start;        /* sample index (within the stream) of the first data window */
    uint32_t first_symbol; /* offset of the packet's symbols in the stream's symbol region */
    uint32_t n_symbols;
    int32_t freq_error;    /* the "error" signal: averaged down-chirp bin error (:242-270) */
    uint32_t reserved;
} lora_b200_stream_packet;

typedef struct lora_b200_stream_result {
    uint64_t consumed;     /* samples consumed by the state machine */
    uint32_t n_packets;    /* packets found; only the first max_packets are stored */
    uint32_t n_symbols;    /* symbols stored (<= max_symbols) */
} lora_b200_stream_result;

/* Samples in one generated frame: (10 + 2 + 2 + n_symbols + max(padding, 1)) * N*osr + N*osr/4 -- the block counts a
 * padding symbol before it compares with the setting (LoRa-SDR/LoRaMod.cpp:219-221), so 0 behaves like 1. */
size_t lora_b200_frame_samples(const lora_b200_ctx* ctx, size_t n_symbols, unsigned padding);
int lora_b200_frame_modulate_batch(lora_b200_ctx* ctx, const uint16_t* d_symbols, size_t n_packets, size_t n_symbols,
                                   size_t symbol_stride, float amplitude, uint8_t sync, unsigned padding,
                                   lora_b200_c32* d_iq, size_t packet_stride, void* stream);
/* d_symbols: n_streams x max_symbols int16 (the reference's output packets carry int16 symbols, :297);
 * d_packets: n_streams x max_packets; d_results: n_streams. */
int lora_b200_stream_detect_batch(lora_b200_ctx* ctx, const lora_b200_stream_cfg* cfg, const lora_b200_c32* d_iq,
                                  size_t n_streams, size_t samples_per_stream, size_t stream_stride,
                                  int16_t* d_symbols, lora_b200_stream_packet* d_packets,
                                  lora_b200_stream_result* d_results, void* stream);
/* LoRaDecoder::work (LoRa-SDR/LoRaDecoder.cpp:196-397) on every packet the synchroniser reported -- in the Pothos
 * topology LoRaDemod's output port feeds LoRaDecoder's input port.  Slot p = stream * max_packets + k of d_payload
 * (LORA_B200_PAYLOAD_STRIDE bytes each) and d_chain_results receives packet k of that stream; a slot with no packet
 * gets n_bytes = -1.  d_symbols / d_packets / d_results are the outputs of lora_b200_stream_detect_batch. */
int lora_b200_stream_decode_batch(lora_b200_ctx* ctx, const lora_b200_chain_cfg* cfg, const int16_t* d_symbols,
                                  const lora_b200_stream_packet* d_packets, const lora_b200_stream_result* d_results,
                                  size_t n_streams, uint32_t max_packets, uint32_t max_symbols, uint8_t* d_payload,
                                  lora_b200_chain_result* d_chain_results, void* stream);

/* Deployed-chain counterparts of lora_b200_chain_decode_batch / _encode_batch / lora_b200_stream_decode_batch.
 * d_payload: LORA_B200_PAYLOAD_STRIDE bytes per packet.  The encoder with d_symbols == NULL only reports the symbol
 * count.  In the stream form the symbols are the synchroniser's int16 bins (taken modulo 2^sf). */
int lora_b200_std_decode_batch(lora_b200_ctx* ctx, const lora_b200_std_cfg* cfg, const uint16_t* d_symbols,
                               size_t n_packets, size_t nsym, size_t symbol_stride, uint8_t* d_payload,
                               lora_b200_std_result* d_results, void* stream);
int lora_b200_std_encode_batch(lora_b200_ctx* ctx, const lora_b200_std_cfg* cfg, const uint8_t* d_payload,
                               size_t n_packets, size_t payload_len, size_t payload_stride, uint16_t* d_symbols,
                               size_t symbol_stride, size_t* symbols_per_packet, void* stream);
int lora_b200_std_stream_decode_batch(lora_b200_ctx* ctx, const lora_b200_std_cfg* cfg, const int16_t* d_symbols,
                                      const lora_b200_stream_packet* d_packets, const lora_b200_stream_result* d_results,
                                      size_t n_streams, uint32_t max_packets, uint32_t max_symbols, uint8_t* d_payload,
                                      lora_b200_std_result* d_std_results, void* stream);
/* One capture, host buffers in and out: copy in, frame synchroniser (decimating by the context's osr), deployed chain,
 * copy out.  h_payload: cfg->max_packets x LORA_B200_PAYLOAD_STRIDE bytes; h_results: cfg->max_packets; h_packets
 * (optional): cfg->max_packets records; *n_packets: packets stored.  stream_cfg->mtu = 0 takes every frame's length
 * from its own header (frame_header mode of the synchroniser; implicit-header frames use cfg->payload_len).
 * bin/lora_rx_stream is this call behind a CLI. */
int lora_b200_std_receive_stream_host(lora_b200_ctx* ctx, const lora_b200_std_cfg* cfg,
                                      const lora_b200_stream_cfg* stream_cfg, const lora_b200_c32* h_iq, size_t count,
                                      uint8_t* h_payload, lora_b200_std_result* h_results,
                                      lora_b200_stream_packet* h_packets, uint32_t* n_packets);

#ifdef __cplusplus
}
#endif
#endif
