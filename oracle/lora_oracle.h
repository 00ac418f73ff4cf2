/* lora_oracle.h -- CPU restatement of the reference's LoRa receive path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it, and
 * only as the checker or the timed CPU baseline.  The shipped path is the CUDA library.
 *
 * Parity status: PINNED.  Every function here is checked (tests/test_oracle_vs_ref.py, run in the
 * build container) against oracle/_ref, the reference's own sources compiled unmodified from
 * /root/reference by oracle/Makefile, and against the reference's known-answer tests
 * (tests/roundtrip_test.cpp, whitening_test.cpp, equal_power_bin_test.cpp, gr_lora_sdr_interop.cpp,
 * no_alloc_test.cpp, e2e_chain_test.cpp) and the rx_runner output on the committed golden capture.
 * The four Pothos blocks (LoRa-SDR/LoRa{Encoder,Decoder,Mod,Demod}.cpp) cannot be compiled here
 * (Pothos is absent); their orchestration is restated line by line and pinned through the
 * reference primitives it calls plus the round-trip invariants of LoRa-SDR/TestLoopback.cpp.
 *
 * All citations are file:line under /root/reference.
 */
#ifndef LORA_ORACLE_H
#define LORA_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float re, im; } orc_c32;

/* ---- code primitives: include/lora_phy/LoRaCodes.hpp ------------------------------------------- */
unsigned orc_round_up(unsigned num, unsigned factor);                       /* :24-27   */
uint8_t  orc_checksum8(const uint8_t* p, size_t len);                       /* :32-41   */
uint8_t  orc_header_checksum(const uint8_t* h);                             /* :43-67   */
uint16_t orc_crc16sx(uint16_t crc, uint16_t poly);                          /* :69-79   */
uint8_t  orc_xsum8(uint8_t t);                                              /* :81-86   */
uint16_t orc_data_checksum(const uint8_t* data, int length);                /* :92-105  */
void     orc_whiten_sx1232(uint8_t* buf, uint16_t n);                       /* :111-137 */
void     orc_whiten_table(uint8_t* buf, uint16_t n, int bit_ofs, int rdd);  /* :147-167 */
void     orc_whiten_lfsr(uint8_t* buf, uint16_t n, int bit_ofs, size_t rdd);/* :176-189 */
uint16_t orc_bin_to_gray(uint16_t v);                                       /* :201-207 */
uint16_t orc_gray_to_bin(uint16_t v);                                       /* :212-222 */
uint8_t  orc_enc_h84(uint8_t x);                                            /* :229-243 */
uint8_t  orc_dec_h84(uint8_t b, int* error, int* bad);                      /* :251-281 */
uint8_t  orc_enc_h74(uint8_t x);                                            /* :287-299 */
uint8_t  orc_dec_h74(uint8_t b, int* error);                                /* :306-334 */
uint8_t  orc_chk_p54(uint8_t b, int* error);                                /* :340-345 */
uint8_t  orc_enc_p54(uint8_t b);                                            /* :347-351 */
uint8_t  orc_chk_p64(uint8_t b, int* error);                                /* :357-365 */
uint8_t  orc_enc_p64(uint8_t b);                                            /* :367-371 */
void     orc_interleave(const uint8_t* cw, size_t ncw, uint16_t* sym, size_t ppm, size_t rdd);   /* :376-393 */
void     orc_deinterleave(const uint16_t* sym, size_t nsym, uint8_t* cw, size_t ppm, size_t rdd); /* :396-412 */

/* ---- chirp, FFT, detector ----------------------------------------------------------------------- */
/* include/lora_phy/ChirpGenerator.hpp:24-50 (float instantiation) */
int  orc_gen_chirp(orc_c32* out, int N, int osr, int NN, float f0, int down, float ampl,
                   float* phase_accum, float bw_scale);
/* include/lora_phy/kissfft.hh:24-29 (twiddles), :80-97 (radix list) */
void orc_fft_twiddles(orc_c32* tw, int nfft, int inverse);
int  orc_fft_factor(int nfft, int* radix, int* remainder);
/* include/lora_phy/kissfft.hh:100-185, radix 2 and 4 only (nfft = 2^k) */
void orc_fft(const orc_c32* tw, int nfft, int inverse, const orc_c32* src, orc_c32* dst);
/* include/lora_phy/LoRaDetector.hpp:39-74; spectrum[] receives the N bins */
size_t orc_detect(const orc_c32* tw, size_t N, const orc_c32* in, orc_c32* spectrum,
                  float* power, float* power_avg, float* f_index, float* max_mag2);

/* ---- lightweight PHY: src/phy/ ------------------------------------------------------------------- */
typedef struct {
    unsigned sf;          /* spreading factor                        */
    unsigned bw_hz;       /* 125000 / 250000 / 500000                */
    unsigned osr;         /* oversampling ratio (0 -> 1)             */
    int      hann;        /* window_type::window_hann when non-zero  */
    int      has_window;  /* ws->window != nullptr (new API only)    */
    uint8_t  sync_word;
} orc_params;

typedef struct {
    int   crc_ok;
    float cfo;
    float time_offset;
    uint8_t sync_word;    /* ws->sync_word after demodulate (phy.cpp:235-237) */
} orc_metrics;

/* src/phy/LoRaEncoder.cpp:6-18, src/phy/LoRaDecoder.cpp:6-19 */
size_t orc_lora_encode(const uint8_t* bytes, size_t n, uint16_t* symbols);
size_t orc_lora_decode(const uint16_t* symbols, size_t n, uint8_t* bytes);
/* src/phy/phy.cpp:241-256 (decode + crc flag) */
long   orc_decode(const uint16_t* symbols, size_t n, uint8_t* payload, size_t cap, int* crc_ok);
/* src/phy/LoRaMod.cpp:8-43 */
size_t orc_lora_modulate(const uint16_t* symbols, size_t n, orc_c32* out, unsigned sf, unsigned osr,
                         unsigned bw_hz, float amplitude, uint8_t sync);
/* src/phy/phy.cpp:78-145 */
void   orc_estimate_offsets(const orc_params* p, const orc_c32* samples, size_t count,
                            float* cfo, float* time_offset);
/* src/phy/phy.cpp:147-176 */
void   orc_compensate_offsets(const orc_params* p, float cfo, float time_offset,
                              orc_c32* samples, size_t count);
/* src/phy/phy.cpp:178-239.  peaks (optional) receives max |X|^2 of every window incl. the two sync ones */
long   orc_demodulate(const orc_params* p, const orc_c32* iq, size_t count, uint16_t* symbols,
                      size_t cap, orc_metrics* m, float* peaks);
/* src/phy/LoRaDemod.cpp:49-195.  scratch may be NULL (then returns 0 when normalisation is needed) */
size_t orc_lora_demodulate(unsigned sf, int hann, const orc_c32* samples, size_t count,
                           uint16_t* out_symbols, unsigned osr, uint8_t* out_sync,
                           orc_c32* scratch, size_t scratch_len, orc_metrics* m, float* peaks);

/* ---- full LoRa coding chain (Pothos blocks restated) ------------------------------------------------ */
typedef struct {
    unsigned sf;
    unsigned ppm;          /* 0 = full symbol set (ppm == sf)          */
    unsigned rdd;          /* 0..4 = CR 4/4 .. 4/8                     */
    int explicit_header;
    int crc;               /* encoder: append CRC; decoder: crcc check  */
    int whitening;         /* encoder only (decoder never reads it)     */
    int interleaving;      /* decoder only                              */
    int error_check;       /* decoder only                              */
    int hdr_out;           /* decoder only: emit header bytes           */
    unsigned data_length;  /* decoder, implicit mode                    */
} orc_chain_cfg;

/* LoRa-SDR/LoRaEncoder.cpp:132-233.  Returns symbol count (or the needed count when symbols==NULL). */
size_t orc_chain_encode(const orc_chain_cfg* c, const uint8_t* payload, size_t len,
                        uint16_t* symbols, size_t cap);
/* What the decoder saw on the way; filled as far as the block got before returning. */
typedef struct {
    int drop_line;     /* 0 = emitted; else the LoRaDecoder.cpp line of the `return this->drop()` taken */
    int parity_error;  /* `error` flag: some codeword had a non-zero syndrome / parity failure        */
    int uncorrectable; /* `bad` flag: some Hamming(8,4) word had a 2-bit syndrome                     */
    int hdr_length;    /* explicit mode: bytes[0]                                                      */
    int hdr_has_crc;   /* explicit mode: bytes[1] & 1                                                  */
    int hdr_rdd;       /* explicit mode: (bytes[1] >> 1) & 7                                           */
    int hdr_checksum;  /* explicit mode: bytes[2] after XOR with headerChecksum (0 = header valid)      */
    int crc_present;   /* a CRC was computed                                                           */
    int crc_ok;        /* computed CRC == packet CRC                                                   */
} orc_chain_info;

/* LoRa-SDR/LoRaDecoder.cpp:196-397.  Returns bytes emitted; -1 when the block drops the packet
 * (info->drop_line says where); -2 when it returns without output (fewer than 8 symbols, :208);
 * -3 where the reference itself is undefined (explicit header, no CRC bit, hdr off and length < 2:
 * `dataLength -= 5` wraps, :381-384).  With interleaving off the Gray'd symbols are written to out as
 * little-endian uint16 and 2*numSymbols is returned (:264-270). */
long   orc_chain_decode(const orc_chain_cfg* c, const uint16_t* symbols, size_t nsym,
                        uint8_t* out, size_t cap, orc_chain_info* info);

/* ---- the LoRa coding chain as deployed on the air (std_chain.c) ---------------------
 * Not in the reference: pinned by the 16 CRC-checked frames of the reference's golden capture, see std_chain.c. */
typedef struct {
    unsigned sf;            /* 7..12 */
    unsigned cr;            /* 1..4 = 4/5 .. 4/8 (explicit header: decoder takes it from the header) */
    int ldro;               /* low data rate optimisation: every block at sf-2 bits per symbol */
    int implicit_header;
    int has_crc;            /* encoder; decoder in implicit-header mode */
    unsigned payload_len;   /* decoder, implicit-header mode */
} orc_std_cfg;
typedef struct {
    int n_bytes;            /* payload bytes; -1 header invalid, -2 fewer than 8 symbols / bad cfg, -3 frame truncated */
    int hdr_ok, hdr_length, hdr_cr, has_crc, crc_ok;
    int corrected, uncorrectable; /* some codeword had one bit corrected / an error that could not be */
    int symbols_used;
} orc_std_info;
uint8_t  orc_std_header_checksum(unsigned n0, unsigned n1, unsigned n2);
void     orc_std_whitening(uint8_t* seq, size_t n);
uint16_t orc_std_crc(const uint8_t* payload, size_t len);
size_t   orc_std_symbol_count(const orc_std_cfg* c, size_t len);
size_t   orc_std_frame_symbols(unsigned sf, int ldro, const uint16_t* sym);
size_t   orc_std_encode(const orc_std_cfg* c, const uint8_t* payload, size_t len, uint16_t* out, size_t cap);
long     orc_std_decode(const orc_std_cfg* c, const uint16_t* symbols, size_t nsym, uint8_t* out, size_t cap,
                        orc_std_info* info);

#ifdef __cplusplus
}
#endif
/* ---- stream.c: upstream Pothos blocks restated (LoRa-SDR/LoRaMod.cpp:109-238, LoRaDemod.cpp:98-327) ---------- */
size_t orc_frame_modulate(unsigned sf, unsigned ovs, const uint16_t* symbols, size_t n_sym, unsigned char sync,
                          unsigned padding, float ampl, orc_c32* out);
void   orc_stream_tables(unsigned sf, orc_c32* up, orc_c32* down, orc_c32* fine);
size_t orc_stream_demod(unsigned sf, unsigned char sync, float thresh, size_t mtu, const orc_c32* samples,
                        size_t count, int16_t* out_symbols, size_t sym_cap, uint32_t* pkt_len, uint64_t* pkt_start,
                        int32_t* pkt_err, size_t pkt_cap, uint64_t* consumed);
/* The same state machine with one addition that is NOT in the reference: frame_header = 1 (2 with LDRO) reads the
 * deployed chain's explicit header after eight data symbols and ends the frame at the length it announces (capped by
 * mtu; an invalid header ends it at eight symbols).  frame_header = 0 is orc_stream_demod. */
size_t orc_stream_demod_hdr(unsigned sf, unsigned char sync, float thresh, size_t mtu, int frame_header,
                            const orc_c32* samples, size_t count, int16_t* out_symbols, size_t sym_cap, uint32_t* pkt_len,
                            uint64_t* pkt_start, int32_t* pkt_err, size_t pkt_cap, uint64_t* consumed);

#endif
