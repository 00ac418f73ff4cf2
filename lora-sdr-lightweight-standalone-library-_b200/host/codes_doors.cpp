// codes_doors.cpp -- extern "C" doors into this build's include/lora_phy/LoRaCodes.hpp, for the tests only.
//
// Same door names and signatures as oracle/ref_shim.cpp gives the reference's header (ref_X) and oracle/codes.c gives
// the restatement (orc_X), with prefix hdr_: tests/test_codes_header.py calls all three with identical arguments over
// each function's whole input range.  Nothing in the product links this file.
#include <lora_phy/LoRaCodes.hpp>

extern "C" {
unsigned hdr_round_up(unsigned a, unsigned b) { return roundUp(a, b); }
uint8_t hdr_checksum8(const uint8_t* p, size_t n) { return checksum8(p, n); }
uint8_t hdr_header_checksum(const uint8_t* h) { return headerChecksum(h); }
uint16_t hdr_crc16sx(uint16_t crc, uint16_t poly) { return crc16sx(crc, poly); }
uint8_t hdr_xsum8(uint8_t t) { return xsum8(t); }
uint16_t hdr_data_checksum(const uint8_t* d, int n) { return sx1272DataChecksum(d, n); }
void hdr_whiten_sx1232(uint8_t* b, uint16_t n) { SX1232RadioComputeWhitening(b, n); }
void hdr_whiten_table(uint8_t* b, uint16_t n, int ofs, int rdd) { Sx1272ComputeWhitening(b, n, ofs, rdd); }
void hdr_whiten_lfsr(uint8_t* b, uint16_t n, int ofs, size_t rdd) { Sx1272ComputeWhiteningLfsr(b, n, ofs, rdd); }
uint16_t hdr_bin_to_gray(uint16_t v) { return binaryToGray16(v); }
uint16_t hdr_gray_to_bin(uint16_t v) { return grayToBinary16(v); }
uint8_t hdr_enc_h84(uint8_t x) { return encodeHamming84sx(x); }
uint8_t hdr_dec_h84(uint8_t b, int* error, int* bad)
{
    bool e = *error != 0, d = *bad != 0;
    const uint8_t r = decodeHamming84sx(b, e, d);
    *error = e; *bad = d;
    return r;
}
uint8_t hdr_enc_h74(uint8_t x) { return encodeHamming74sx(x); }
uint8_t hdr_dec_h74(uint8_t b, int* error) { bool e = *error != 0; const uint8_t r = decodeHamming74sx(b, e); *error = e; return r; }
uint8_t hdr_chk_p54(uint8_t b, int* error) { bool e = *error != 0; const uint8_t r = checkParity54(b, e); *error = e; return r; }
uint8_t hdr_enc_p54(uint8_t b) { return encodeParity54(b); }
uint8_t hdr_chk_p64(uint8_t b, int* error) { bool e = *error != 0; const uint8_t r = checkParity64(b, e); *error = e; return r; }
uint8_t hdr_enc_p64(uint8_t b) { return encodeParity64(b); }
void hdr_interleave(const uint8_t* cw, size_t ncw, uint16_t* sym, size_t ppm, size_t rdd) { diagonalInterleaveSx(cw, ncw, sym, ppm, rdd); }
void hdr_deinterleave(const uint16_t* sym, size_t nsym, uint8_t* cw, size_t ppm, size_t rdd) { diagonalDeterleaveSx(sym, nsym, cw, ppm, rdd); }
void hdr_deinterleave2(const uint16_t* sym, size_t nsym, uint8_t* cw, size_t ppm, size_t rdd) { diagonalDeterleaveSx2(sym, nsym, cw, ppm, rdd); }
}
