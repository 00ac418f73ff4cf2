"""oracle/chain.c and oracle/stream.c against the reference's OWN code for SURVEY.md section 8 rows a18-a21.

`POTHOS` is oracle/_ref/liblora_pothos_ref.so: LoRa-SDR/LoRaDecoder.cpp, LoRaEncoder.cpp, LoRaDemod.cpp and
LoRaMod.cpp compiled UNMODIFIED from /root/reference against the header-only stand-in oracle/pothos_stub (ports
backed by plain memory, a scheduler loop in oracle/pothos_doors.cpp).  Every comparison here is bit for bit:
symbols, bytes, drop decisions, modulated IQ, packet boundaries, frequency errors and consumed-sample counts.
The loops follow LoRa-SDR/TestLoopback.cpp:14-133 (SF 7..12 x CR 4/4..4/8, then the noisy SF10 loopback).  CPU only.
"""
import numpy as np
import pytest

from _oracle import ORC, POTHOS, chain_cfg

pytestmark = pytest.mark.skipif(POTHOS is None, reason="oracle/_ref/liblora_pothos_ref.so was never built")


def same_decode(cfg, sym, ctx):
    ro, bo, io = ORC.chain_decode(cfg, sym)
    rp, bp, _ = POTHOS.chain_decode(cfg, sym)
    if ro == -3:       # the reference itself is undefined there (`dataLength -= 5` wraps, LoRaDecoder.cpp:381-384)
        return ro
    assert ro == rp, (ctx, ro, rp, io.drop_line)
    assert bytes(bo) == bytes(bp), ctx
    return ro


@pytest.mark.parametrize("sf", range(7, 13))
@pytest.mark.parametrize("rdd", range(5))
def test_encoder_and_decoder_equal_the_blocks(sf, rdd):
    """TestLoopback.cpp:14-62 grid, plus LDRO symbol sets, implicit header, CRC and header-output switches."""
    r = np.random.default_rng(1000 + sf * 10 + rdd)
    for ppm in (0, sf - 2, sf - 1):
        for n in (1, 2, 3, 5, 16, 17, 50, 255):
            payload = bytes(r.integers(0, 256, n, dtype=np.uint8))
            for explicit in (True, False):
                for crc in (True, False):
                    for whitening in (True, False):
                        cfg = chain_cfg(sf, rdd, ppm=ppm, explicit=explicit, crc=crc, whitening=whitening,
                                        data_length=n)
                        so = ORC.chain_encode(cfg, payload)
                        sp = POTHOS.chain_encode(cfg, payload)
                        assert len(so) == len(sp) and (so == sp).all(), (sf, rdd, ppm, n, explicit, crc, whitening)
                        if not whitening:
                            continue  # the decoder never reads its whitening flag; the encoder switch is pinned above
                        for hdr_out in (False, True):
                            for error_check in (False, True):
                                cfg_d = chain_cfg(sf, rdd, ppm=ppm, explicit=explicit, crc=crc, hdr_out=hdr_out,
                                                  error_check=error_check, data_length=n)
                                ret = same_decode(cfg_d, so, (sf, rdd, ppm, n, explicit, crc, hdr_out, error_check))
                                assert ret != -1 or (not crc and explicit)


@pytest.mark.parametrize("sf,rdd,ppm", [(7, 1, 0), (7, 4, 0), (8, 2, 0), (9, 3, 0), (10, 0, 0), (11, 1, 9), (12, 4, 10),
                                        (12, 2, 0), (9, 4, 7)])
def test_decoder_on_corrupted_and_arbitrary_symbols(sf, rdd, ppm):
    """Drop decisions and corrected bytes on damaged packets: bit flips, +-1 bin errors, truncation, random symbols."""
    r = np.random.default_rng(sf * 100 + rdd)
    cfgs = [chain_cfg(sf, rdd, ppm=ppm, crc=crc, error_check=ec, hdr_out=h, explicit=ex, data_length=20)
            for crc in (True, False) for ec in (True, False) for h in (False, True) for ex in (True, False)]
    payload = bytes(r.integers(0, 256, 20, dtype=np.uint8))
    outcomes = set()
    for cfg in cfgs:
        enc = chain_cfg(sf, rdd, ppm=ppm, crc=True, explicit=bool(cfg.explicit_header))
        sym = ORC.chain_encode(enc, payload)
        for trial in range(60):
            bad = sym.copy()
            kind = trial % 6
            if kind == 0:
                bad[r.integers(0, len(bad))] ^= np.uint16(1 << int(r.integers(0, sf)))
            elif kind == 1:
                k = r.integers(0, len(bad)); bad[k] = (int(bad[k]) + int(r.choice([-1, 1]))) % (1 << sf)
            elif kind == 2:
                for k in r.integers(0, len(bad), 3):
                    bad[k] ^= np.uint16(r.integers(1, 1 << sf))
            elif kind == 3:
                bad = bad[: int(r.integers(0, len(bad) + 1))]
            elif kind == 4:
                bad = r.integers(0, 1 << sf, int(r.integers(8, 60)), dtype=np.uint16)
            else:
                bad = np.concatenate([bad, r.integers(0, 1 << sf, int(r.integers(1, 9)), dtype=np.uint16)])
            outcomes.add(min(same_decode(cfg, bad, (sf, rdd, ppm, trial)), 0))
    assert {0, -1} <= outcomes


def test_decoder_interleaving_off_and_short_input():
    for sf in (7, 9, 12):
        cfg = chain_cfg(sf, 4, interleaving=False)
        sym = np.random.default_rng(sf).integers(0, 1 << sf, 16, dtype=np.uint16)
        assert same_decode(cfg, sym, sf) == 32
        assert same_decode(chain_cfg(sf, 1), sym[:7], sf) == -2


@pytest.mark.parametrize("sf", [7, 8, 9, 10, 11, 12])
def test_frame_generator_equals_the_block(sf):
    r = np.random.default_rng(sf)
    for n_sym, sync, padding, ampl, ovs in ((20, 0x12, 1, 0.3, 1), (1, 0x34, 3, 1.0, 1), (33, 0xFF, 2, 0.5, 2),
                                             (8, 0x00, 1, 0.3, 4)):
        if sf >= 11 and ovs > 2:
            continue
        sym = r.integers(0, 1 << sf, n_sym, dtype=np.uint16)
        a = ORC.frame_modulate(sf, sym, sync=sync, padding=padding, ampl=ampl, ovs=ovs)
        b = POTHOS.frame_modulate(sf, sym, sync=sync, padding=padding, ampl=ampl, ovs=ovs)
        assert len(a) == len(b)
        assert a.tobytes() == b.tobytes(), (sf, n_sym, sync, padding, ampl, ovs)


def channel(frame, n, lead, tail, cfo_bins, snr_db, rng, ampl=0.3):
    x = np.concatenate([np.zeros(lead, np.complex64), frame, np.zeros(tail, np.complex64)])
    k = np.arange(len(x))
    x = x * np.exp(2j * np.pi * cfo_bins * k / n)
    sigma = ampl * 10 ** (-snr_db / 20) / np.sqrt(2)
    x = x + (rng.standard_normal(len(x)) + 1j * rng.standard_normal(len(x))) * sigma
    return x.astype(np.complex64)


def same_stream(sf, x, **kw):
    po, so, eo, co = ORC.stream_demod(sf, x, **kw)
    pp, sp, ep, cp = POTHOS.stream_demod(sf, x, **kw)
    assert co == cp, ("consumed", co, cp)
    assert len(po) == len(pp), ("packets", len(po), len(pp))
    for a, b in zip(po, pp):
        assert len(a) == len(b) and (a == b).all()
    assert (so == sp).all() and (eo == ep).all()
    return po


@pytest.mark.parametrize("sf", [7, 8, 9, 10])
@pytest.mark.parametrize("cfo_bins,lead_frac,snr_db", [(0.0, 0.0, 80.0), (0.3, 0.29, 20.0), (-2.4, 0.52, 10.0),
                                                       (5.2, 0.09, 0.0), (-0.49, 0.77, -6.0), (17.3, 0.5, 3.0)])
def test_stream_synchroniser_equals_the_block(sf, cfo_bins, lead_frac, snr_db):
    """TestLoopback.cpp:64-133 in spirit: generator -> noisy channel -> synchroniser, at several offsets."""
    n = 1 << sf
    rng = np.random.default_rng(sf * 100 + int(snr_db) + 7)
    sym = rng.integers(0, n, 20, dtype=np.uint16)
    fr = ORC.frame_modulate(sf, sym, padding=4)
    x = channel(fr, n, 3 * n + int(lead_frac * n), 4 * n, cfo_bins, snr_db, rng)
    same_stream(sf, x, mtu=len(sym))
    same_stream(sf, x, mtu=64, thresh=-10.0)
    same_stream(sf, x, mtu=5, thresh=3.0, sync=0x34)


def test_stream_multi_frame_never_squelched_and_degenerate():
    sf, n = 8, 256
    rng = np.random.default_rng(5)
    parts = []
    for k in range(5):
        s = rng.integers(0, n, int(rng.integers(5, 30)), dtype=np.uint16)
        parts.append(channel(ORC.frame_modulate(sf, s, padding=int(rng.integers(1, 4))), n, int(rng.integers(0, 3 * n)), 0,
                             float(rng.uniform(-8, 8)), float(rng.uniform(0, 25)), rng))
    x = np.concatenate(parts + [np.zeros(3 * n, np.complex64)])
    pk = same_stream(sf, x, mtu=64, thresh=-10.0)
    assert len(pk) >= 3
    same_stream(sf, x, mtu=300, thresh=-200.0)       # never squelched: the fine-tune index walks through its wraps
    same_stream(sf, x, mtu=16, thresh=-200.0)
    noise = ((rng.standard_normal(40 * n) + 1j * rng.standard_normal(40 * n)) * 0.1).astype(np.complex64)
    same_stream(sf, noise, thresh=10.0)
    same_stream(sf, noise, thresh=-200.0, mtu=7)
    same_stream(sf, noise[: 2 * n - 1])
    same_stream(sf, noise[:0])
    same_stream(sf, np.zeros(20 * n, np.complex64))  # all-zero windows: snr is NaN, never "squelched"


def test_stream_sf11_sf12_once():
    for sf in (11, 12):
        n = 1 << sf
        rng = np.random.default_rng(sf)
        sym = rng.integers(0, n, 6, dtype=np.uint16)
        x = channel(ORC.frame_modulate(sf, sym, padding=2), n, n + 1234, 3 * n, 1.7, 12.0, rng)
        pk = same_stream(sf, x, mtu=len(sym))
        assert len(pk) == 1 and (pk[0].astype(np.int64) == sym.astype(np.int64)).all()
