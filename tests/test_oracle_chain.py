"""Invariants of the full coding chain (oracle/chain.c), restating the Pothos tests the reference cannot
run here: LoRa-SDR/TestLoopback.cpp:14-62 (encoder -> decoder identity, SF 7..12 x CR 4/4..4/8),
LoRa-SDR/TestCodesSx.cpp (single-bit correction / detection), and the known answers of
tests/whitening_test.cpp:30-49.  CPU only.
"""
import ctypes as C

import numpy as np
import pytest

from _oracle import ORC, chain_cfg, u8p


def test_whitening_and_crc_known_answers():
    plain = np.frombuffer(bytes.fromhex("deadbeef700d"), np.uint8).copy()
    buf = plain.copy()
    f = ORC.fn("whiten_lfsr", None, [u8p, C.c_uint16, C.c_int, C.c_size_t])
    f(buf, len(buf), 0, 4)
    assert bytes(buf) == bytes.fromhex("215290102cf2")
    f(buf, len(buf), 0, 4)
    assert bytes(buf) == bytes(plain)
    crc = ORC.fn("data_checksum", C.c_uint16, [u8p, C.c_int])(plain, 4)
    assert crc == 0x0D70


def test_table_and_lfsr_whitening_agree():
    # the TX side whitens with the 510-bit table, the RX side with the LFSRs; they must be one sequence
    ft = ORC.fn("whiten_table", None, [u8p, C.c_uint16, C.c_int, C.c_int])
    fl = ORC.fn("whiten_lfsr", None, [u8p, C.c_uint16, C.c_int, C.c_size_t])
    for rdd in range(5):
        for ofs in range(0, 17):
            a, b = np.zeros(600, np.uint8), np.zeros(600, np.uint8)
            ft(a, 600, ofs, rdd)
            fl(b, 600, ofs, rdd)
            assert bytes(a) == bytes(b), (rdd, ofs)


@pytest.mark.parametrize("sf", range(7, 13))
@pytest.mark.parametrize("rdd", range(5))
def test_encoder_decoder_identity(sf, rdd):
    r = np.random.default_rng(sf * 10 + rdd)
    for ppm in (0, sf - 2):  # full symbol set, and the reduced set LDRO uses
        for n in (1, 2, 5, 16, 17, 50, 255):
            for crc in (True, False):
                payload = bytes(r.integers(0, 256, n, dtype=np.uint8))
                cfg = chain_cfg(sf, rdd, ppm=ppm, crc=crc)
                sym = ORC.chain_encode(cfg, payload)
                ppm_eff = ppm or sf
                ncw = -(-((n + 2 * crc) * 2 + 5) // ppm_eff) * ppm_eff
                assert len(sym) == 8 + (ncw // ppm_eff - 1) * (4 + rdd)
                assert int(sym.max()) < (1 << sf)
                ret, out, info = ORC.chain_decode(cfg, sym)
                if not crc and n < 2:
                    assert ret == -3  # the reference's own size_t wrap (LoRaDecoder.cpp:381-384)
                    continue
                want = payload if crc else payload[: n - 2]  # `dataLength -= 5` even without a CRC
                assert ret == len(want) and bytes(out) == want, (sf, rdd, ppm, n, crc)
                assert info.drop_line == 0 and info.hdr_length == n and info.hdr_rdd == rdd
                assert info.hdr_checksum == 0 and info.parity_error == 0
                assert info.crc_present == int(crc) and (info.crc_ok == 1 or not crc)
                # header output on: 3 header bytes + payload + (zeroed) crc
                cfg_h = chain_cfg(sf, rdd, ppm=ppm, crc=crc, hdr_out=True)
                ret, out, _ = ORC.chain_decode(cfg_h, sym)
                assert ret == n + (5 if crc else 3) and bytes(out[3:3 + n]) == payload
                if crc:
                    assert bytes(out[3 + n:]) == b"\0\0"


@pytest.mark.parametrize("rdd,correctable", [(3, True), (4, True), (1, False), (2, False)])
def test_single_symbol_bit_error(rdd, correctable):
    """One flipped symbol bit lands in one codeword: 4/7 and 4/8 correct it, 4/5 and 4/6 only detect."""
    sf = 9
    payload = bytes(range(40))
    cfg = chain_cfg(sf, rdd, crc=True)
    sym = ORC.chain_encode(cfg, payload)
    flips = 0
    for pos in range(8, len(sym)):
        for b in range(sf):
            bad = sym.copy()
            # flip a bit of the Gray-domain value: go through the same +-1 neighbour the channel would
            bad[pos] ^= 1 << b
            ret, out, info = ORC.chain_decode(cfg, bad)
            flips += 1
            if correctable:
                assert ret == len(payload) and bytes(out) == payload, (pos, b)
                assert info.crc_ok == 1
            else:
                assert info.parity_error == 1 or bytes(out) == payload
    assert flips > 100
    # with errorCheck on, a detected-but-uncorrected error drops the packet (:363)
    if not correctable:
        bad = sym.copy(); bad[10] ^= 1
        cfg_e = chain_cfg(sf, rdd, crc=True, error_check=True)
        ret, _, info = ORC.chain_decode(cfg_e, bad)
        assert ret == -1 and info.drop_line in (342, 363)


def test_decoder_drop_paths_and_short_input():
    cfg = chain_cfg(8, 2, crc=True)
    payload = b"hello lora"
    sym = ORC.chain_encode(cfg, payload)
    assert ORC.chain_decode(cfg, sym[:7])[0] == -2          # fewer than 8 symbols: no output (:208)
    bad = sym.copy(); bad[9] ^= 0x3f   # row 1 of the first payload block carries DATA bit 1 (row >= 4 would be parity only)
    ret, out, info = ORC.chain_decode(cfg, bad)
    assert ret == -1 and info.drop_line == 371 and info.crc_ok == 0
    cfg_nc = chain_cfg(8, 2, crc=False)                      # crcc off: same packet is emitted anyway
    ret, out, info = ORC.chain_decode(cfg_nc, bad)
    assert ret == len(payload) and info.crc_present == 1 and info.crc_ok == 0
    ret, _, info = ORC.chain_decode(cfg, sym[:9])            # header says more bytes than were received
    assert ret == -1 and info.drop_line == 313


def test_implicit_header_mode():
    for sf, rdd in ((7, 1), (10, 4), (12, 3)):
        payload = bytes(range(1, 21))
        for crc in (True, False):
            cfg = chain_cfg(sf, rdd, explicit=False, crc=crc, data_length=len(payload))
            sym = ORC.chain_encode(cfg, payload)
            ret, out, info = ORC.chain_decode(cfg, sym)
            want = payload + (b"\0\0" if crc else b"")
            assert ret == len(want) and bytes(out) == want, (sf, rdd, crc)


def test_interleaving_off_returns_gray_symbols():
    cfg = chain_cfg(9, 4, interleaving=False)
    sym = np.array([0, 1, 2, 3, 100, 511, 256, 255], np.uint16)
    ret, out, _ = ORC.chain_decode(cfg, sym)
    got = np.frombuffer(bytes(out), np.uint16)
    assert ret == 16 and list(got) == [int(v) ^ (int(v) >> 1) for v in sym]
