"""include/lora_phy/LoRaCodes.hpp of this build (its own bodies behind the reference's names and signatures) against
the reference's header (`REF`, unmodified, where oracle/_ref was built) and the oracle restatement (`ORC`, itself pinned
to REF by tests/test_oracle_vs_ref.py): every function over its whole input range.  CPU only -- these helpers are
caller-side code (runners/lora_phy_vector_dump.cpp:105,153-160), not the device path.
"""
import ctypes as C
import os

import numpy as np
import pytest

from _oracle import ORC, REF, Lib, ROOT, u8p, u16p

_path = os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-_b200", "build", "libcodes_hdr.so")
HDR = Lib(_path, "hdr") if os.path.exists(_path) else None
pytestmark = pytest.mark.skipif(HDR is None, reason="build/libcodes_hdr.so not built (make -C the package)")
OTHERS = [ORC] + ([REF] if REF is not None else [])


def bits(a):
    return np.ascontiguousarray(a).view(np.uint8).tobytes()


@pytest.mark.parametrize("other", OTHERS, ids=lambda l: l.prefix)
def test_scalar_functions_exhaustive(other):
    for name, res in [("bin_to_gray", C.c_uint16), ("gray_to_bin", C.c_uint16)]:
        fa, fb = HDR.fn(name, res, [C.c_uint16]), other.fn(name, res, [C.c_uint16])
        assert all(fa(v) == fb(v) for v in range(65536)), name
    for name in ("enc_h84", "enc_h74", "enc_p54", "enc_p64", "xsum8"):
        fa, fb = HDR.fn(name, C.c_uint8, [C.c_uint8]), other.fn(name, C.c_uint8, [C.c_uint8])
        assert all(fa(v) == fb(v) for v in range(256)), name
    ip = C.POINTER(C.c_int)
    for name in ("dec_h74", "chk_p54", "chk_p64"):
        fa, fb = HDR.fn(name, C.c_uint8, [C.c_uint8, ip]), other.fn(name, C.c_uint8, [C.c_uint8, ip])
        for v in range(256):
            for start in (0, 1):        # the flag is only ever set, never cleared
                ea, eb = C.c_int(start), C.c_int(start)
                assert fa(v, C.byref(ea)) == fb(v, C.byref(eb)) and ea.value == eb.value, (name, v)
    fa = HDR.fn("dec_h84", C.c_uint8, [C.c_uint8, ip, ip])
    fb = other.fn("dec_h84", C.c_uint8, [C.c_uint8, ip, ip])
    for v in range(256):
        for se, sb in ((0, 0), (1, 0), (0, 1), (1, 1)):
            ea, ba, eb, bb = C.c_int(se), C.c_int(sb), C.c_int(se), C.c_int(sb)
            assert fa(v, C.byref(ea), C.byref(ba)) == fb(v, C.byref(eb), C.byref(bb))
            assert (ea.value, ba.value) == (eb.value, bb.value), v
    fa, fb = HDR.fn("crc16sx", C.c_uint16, [C.c_uint16, C.c_uint16]), other.fn("crc16sx", C.c_uint16, [C.c_uint16, C.c_uint16])
    assert all(fa(v, p) == fb(v, p) for v in range(0, 65536, 3) for p in (0x1021, 0x8005))
    fa, fb = HDR.fn("round_up", C.c_uint, [C.c_uint, C.c_uint]), other.fn("round_up", C.c_uint, [C.c_uint, C.c_uint])
    assert all(fa(a, b) == fb(a, b) for a in range(0, 700) for b in range(1, 14))
    for h0 in range(256):
        for h1 in range(256):
            h = np.array([h0, h1, 0], np.uint8)
            assert HDR.fn("header_checksum", C.c_uint8, [u8p])(h) == other.fn("header_checksum", C.c_uint8, [u8p])(h)


@pytest.mark.parametrize("other", OTHERS, ids=lambda l: l.prefix)
def test_buffer_functions(other):
    r = np.random.default_rng(1)
    for n in (0, 1, 2, 3, 5, 16, 255, 300, 509, 510, 511, 1200):
        data = r.integers(0, 256, max(n, 1), dtype=np.uint8)
        assert HDR.fn("data_checksum", C.c_uint16, [u8p, C.c_int])(data, n) == other.fn("data_checksum", C.c_uint16, [u8p, C.c_int])(data, n)
        assert HDR.fn("checksum8", C.c_uint8, [u8p, C.c_size_t])(data, n) == other.fn("checksum8", C.c_uint8, [u8p, C.c_size_t])(data, n)
        a, b = data.copy(), data.copy()
        HDR.fn("whiten_sx1232", None, [u8p, C.c_uint16])(a, n)
        other.fn("whiten_sx1232", None, [u8p, C.c_uint16])(b, n)
        assert bits(a) == bits(b)
        for rdd in range(5):
            for ofs in (0, 1, 2, 5, 7, 12, 16, 100, 255, 509, 510, 511, 777):
                for name, t in (("whiten_table", C.c_int), ("whiten_lfsr", C.c_size_t)):
                    a, b = data.copy(), data.copy()
                    HDR.fn(name, None, [u8p, C.c_uint16, C.c_int, t])(a, n, ofs, rdd)
                    other.fn(name, None, [u8p, C.c_uint16, C.c_int, t])(b, n, ofs, rdd)
                    assert bits(a) == bits(b), (name, n, rdd, ofs)
    # the known answers of tests/whitening_test.cpp:30-49
    buf = np.frombuffer(bytes.fromhex("deadbeef700d"), np.uint8).copy()
    HDR.fn("whiten_lfsr", None, [u8p, C.c_uint16, C.c_int, C.c_size_t])(buf, 6, 0, 4)
    assert bytes(buf) == bytes.fromhex("215290102cf2")
    assert HDR.fn("data_checksum", C.c_uint16, [u8p, C.c_int])(np.frombuffer(bytes.fromhex("deadbeef"), np.uint8).copy(), 4) == 0x0D70


@pytest.mark.parametrize("other", OTHERS, ids=lambda l: l.prefix)
def test_interleavers(other):
    r = np.random.default_rng(2)
    args_i = [u8p, C.c_size_t, u16p, C.c_size_t, C.c_size_t]
    args_d = [u16p, C.c_size_t, u8p, C.c_size_t, C.c_size_t]
    for ppm in range(4, 13):
        for rdd in range(5):
            for blocks in (1, 3):
                cw = r.integers(0, 256, ppm * blocks, dtype=np.uint8)      # stray high bits must be ignored alike
                sa = np.full((4 + rdd) * blocks, 0xffff, np.uint16)        # the lightweight interleaver ASSIGNS its output
                sb = sa.copy()
                HDR.fn("interleave", None, args_i)(cw, len(cw), sa, ppm, rdd)
                other.fn("interleave", None, args_i)(cw, len(cw), sb, ppm, rdd)
                assert bits(sa) == bits(sb), (ppm, rdd)
                sym = r.integers(0, 1 << 16, (4 + rdd) * blocks + 16, dtype=np.uint16)
                ca, cb = np.zeros(ppm * blocks, np.uint8), np.zeros(ppm * blocks, np.uint8)
                HDR.fn("deinterleave", None, args_d)(sym, (4 + rdd) * blocks, ca, ppm, rdd)
                other.fn("deinterleave", None, args_d)(sym, (4 + rdd) * blocks, cb, ppm, rdd)
                assert bits(ca) == bits(cb), (ppm, rdd)
                if other.has("deinterleave2"):                              # the reference's unused second form, quirks kept
                    ca, cb = np.zeros(ppm * blocks, np.uint8), np.zeros(ppm * blocks, np.uint8)
                    HDR.fn("deinterleave2", None, args_d)(sym, (4 + rdd) * blocks, ca, ppm, rdd)
                    other.fn("deinterleave2", None, args_d)(sym, (4 + rdd) * blocks, cb, ppm, rdd)
                    assert bits(ca) == bits(cb), (ppm, rdd)
