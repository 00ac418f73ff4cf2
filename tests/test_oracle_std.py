"""CPU: the deployed LoRa coding chain (oracle/std_chain.c).

The reference has no such chain, so there is no reference output to compare with.  What pins the restatement is the
reference's own golden capture: through the restated Pothos frame synchroniser and this chain, every frame decodes to
the text its transmitter sent, with a valid header checksum and a valid payload CRC."""
import os

import numpy as np
import pytest

from _oracle import ORC, std_cfg

HERE = os.path.dirname(os.path.abspath(__file__))


def capture():
    return np.load(os.path.join(HERE, "golden", "capture_sf7.npz"))["iq"]


def test_golden_capture_decodes_to_its_sixteen_frames():
    """vectors_binary/bw_125k_sf_7_cr_1_ldro_false_crc_true_implheader_false.unknown: 2x oversampled, a frame every 65.25
    symbols.  Decimated from its first sample the dechirped tones fall on whole bins; rx_runner prints noise
    for this file, the synchroniser + deployed chain print the sixteen messages."""
    x = np.ascontiguousarray(capture()[0::2])
    frames, start, _, _ = ORC.stream_demod(7, x, mtu=40, pkt_cap=32)
    assert len(frames) == 16 and [int(s) for s in start] == [1568 + 8352 * k for k in range(16)]
    for k, sym in enumerate(frames):
        ret, data, info = ORC.std_decode(std_cfg(7), sym.astype(np.uint16) & 127)
        text = f"hello world: {k}".encode()
        assert ret == len(text) and bytes(data) == text
        assert info.hdr_ok == 1 and info.hdr_cr == 1 and info.has_crc == 1 and info.crc_ok == 1
        assert info.corrected == 0 and info.uncorrectable == 0 and info.symbols_used == 33
        # the transmitter's side of the same frame
        assert bytes(ORC.std_encode(std_cfg(7, cr=1, crc=True), np.frombuffer(text, np.uint8))) == \
            bytes((sym[:33].astype(np.uint16) & 127))


def test_header_checksum_and_crc_known_answers():
    # header nibbles of the capture's frames: length 14 / 15, CR 4/5 + CRC -> checksum 3 / 5
    assert ORC.std_header_checksum(0, 14, 3) == 3 and ORC.std_header_checksum(0, 15, 3) == 5
    # payload CRCs carried by frames 0 and 10 of the capture
    assert ORC.std_crc(np.frombuffer(b"hello world: 0", np.uint8)) == 54289
    assert ORC.std_crc(np.frombuffer(b"hello world: 10", np.uint8)) == 39881


@pytest.mark.parametrize("sf", range(7, 13))
def test_round_trip_all_modes(sf):
    rng = np.random.default_rng(sf)
    for cr in (1, 2, 3, 4):
        for ldro in (False, True):
            for implicit in (False, True):
                for crc in (False, True):
                    for n in (1, 2, 3, 16, 17, 64, 255):
                        pay = rng.integers(0, 256, n, dtype=np.uint8)
                        enc = std_cfg(sf, cr, ldro, implicit, crc)
                        sym = ORC.std_encode(enc, pay)
                        assert len(sym) == ORC.std_symbol_count(enc, n) and sym.max() < (1 << sf)
                        # the decoder of an explicit-header frame learns cr, crc and the length from the header
                        dec = std_cfg(sf, cr if implicit else 4, ldro, implicit, crc if implicit else False, n if implicit else 0)
                        ret, data, info = ORC.std_decode(dec, sym)
                        assert ret == n and bytes(data) == bytes(pay), (sf, cr, ldro, implicit, crc, n)
                        assert info.hdr_ok == 1 and info.has_crc == int(crc) and info.crc_ok == int(crc)
                        assert info.symbols_used == len(sym) and not info.corrected and not info.uncorrectable


@pytest.mark.parametrize("cr", [1, 2, 3, 4])
def test_symbol_errors(cr):
    """A symbol off by one bin flips one bit of its Gray code, i.e. one bit in one codeword of the block: 4/7 and
    4/8 correct it, 4/5 and 4/6 notice it.  Reduced-rate symbols shrug off +-1 entirely."""
    sf = 9
    rng = np.random.default_rng(cr)
    pay = rng.integers(0, 256, 40, dtype=np.uint8)
    cfg = std_cfg(sf, cr, crc=True)
    sym = ORC.std_encode(cfg, pay)
    hit = sym.copy()
    hit[3] = (hit[3] + 1) % (1 << sf)                    # header block: reduced rate
    ret, data, info = ORC.std_decode(cfg, hit)
    assert ret == 40 and bytes(data) == bytes(pay) and info.crc_ok == 1 and not info.corrected
    hit = sym.copy()
    hit[8 + 2] ^= 1 if (hit[8 + 2] - 1) % 2 == 0 else 0  # keep the change to a single Gray bit
    hit[8 + 2] = (((hit[8 + 2] - 1) ^ 1) + 1) % (1 << sf)
    ret, data, info = ORC.std_decode(cfg, hit)
    if cr >= 3:
        assert ret == 40 and bytes(data) == bytes(pay) and info.crc_ok == 1 and info.corrected == 1
    else:
        assert info.uncorrectable == 1 and info.crc_ok == 0


def test_rejections():
    cfg = std_cfg(8)
    sym = ORC.std_encode(cfg, np.arange(20, dtype=np.uint8))
    assert ORC.std_decode(cfg, sym[:7])[0] == -2                       # no header block
    assert ORC.std_decode(cfg, sym[:-1])[0] == -3                      # frame cut short
    rejected = 0
    for k in range(1, 40):                                              # two header symbols moved to other reduced-rate bins
        bad = sym.copy()
        bad[0] = (bad[0] + 4 * k) % 256
        bad[5] = (bad[5] + 4 * (k + 3)) % 256
        ret, _, info = ORC.std_decode(cfg, bad)
        rejected += ret == -1
        assert ret == -1 or info.corrected or info.uncorrectable       # never silently accepted
    assert rejected > 0


def test_synchroniser_reads_the_frame_length_from_the_header():
    """frame_header mode (not in the reference): no per-capture mtu is needed, and frames of different lengths that follow
    each other closely are each cut at their own last symbol."""
    x = np.ascontiguousarray(capture()[0::2])
    frames, start, _, _ = ORC.stream_demod(7, x, mtu=600, pkt_cap=32, frame_header=1)
    assert len(frames) == 16 and all(len(f) == 33 for f in frames)
    assert [int(s) for s in start] == [1568 + 8352 * k for k in range(16)]
    # a train of unequal frames, one symbol of padding between them
    sf, cfg = 8, std_cfg(8, cr=2)
    rng = np.random.default_rng(3)
    pays = [rng.integers(0, 256, n, dtype=np.uint8) for n in (3, 60, 1, 17, 200)]
    syms = [ORC.std_encode(cfg, p) for p in pays]
    x = np.concatenate([ORC.frame_modulate(sf, s, padding=1, ampl=0.5) for s in syms] + [np.zeros(3 << sf, np.complex64)])
    frames, _, _, _ = ORC.stream_demod(sf, x, mtu=1000, pkt_cap=16, frame_header=1)
    assert [len(f) for f in frames] == [len(s) for s in syms]
    for f, p in zip(frames, pays):
        ret, data, info = ORC.std_decode(std_cfg(sf), f.astype(np.uint16) & 255)
        assert ret == len(p) and bytes(data) == bytes(p) and info.crc_ok == 1
    # the reference's behaviour (frame_header = 0) swallows the following frames into the first packet
    frames0, _, _, _ = ORC.stream_demod(sf, x, mtu=1000, pkt_cap=16)
    assert len(frames0) < len(syms)


@pytest.mark.parametrize("sf", [7, 8, 10, 12])
def test_c_oracle_agrees_with_the_independent_python_statement(sf):
    """tests/_std_python.py is the script that first decoded the capture, written before std_chain.c: frames encoded by
    the C oracle must decode in it, for every coding rate, with and without LDRO and CRC."""
    from _std_python import decode
    rng = np.random.default_rng(100 + sf)
    for cr in (1, 2, 3, 4):
        for ldro in (False, True):
            for crc in (False, True):
                for n in (1, 2, 5, 31, 64):
                    pay = rng.integers(0, 256, n, dtype=np.uint8)
                    sym = ORC.std_encode(std_cfg(sf, cr, ldro, False, crc), pay)
                    ln, has_crc, cr_rx, chk, data, _, ok = decode(sym, sf, ldro)
                    assert (ln, has_crc, cr_rx) == (n, int(crc), cr) and data == bytes(pay), (sf, cr, ldro, crc, n)
                    assert chk == ORC.std_header_checksum(n >> 4, n & 15, cr << 1 | int(crc))
                    assert ok is (True if crc else None)
    x = np.ascontiguousarray(capture()[0::2])
    frames, _, _, _ = ORC.stream_demod(7, x, mtu=33, pkt_cap=32)
    assert [decode(f, 7)[4] for f in frames] == [f"hello world: {k}".encode() for k in range(16)]
    assert all(decode(f, 7)[6] for f in frames)
