"""The restated Pothos stream blocks (oracle/stream.c): frame generator -> channel -> frame synchroniser.

The reference holds no expected output for LoRa-SDR/LoRaMod.cpp / LoRaDemod.cpp (they need the Pothos runtime, which
is not here), so the restatement is pinned by the property the reference's own TestLoopback states: the symbols that
go into the generator come out of the synchroniser, with carrier offset, arbitrary start position and noise.
"""
import numpy as np
import pytest

from _oracle import ORC


def channel(frame, n, lead, tail, cfo_bins, snr_db, rng, ampl=0.3):
    x = np.concatenate([np.zeros(lead, np.complex64), frame, np.zeros(tail, np.complex64)])
    k = np.arange(len(x))
    x = x * np.exp(2j * np.pi * cfo_bins * k / n)
    sigma = ampl * 10 ** (-snr_db / 20) / np.sqrt(2)
    x = x + (rng.standard_normal(len(x)) + 1j * rng.standard_normal(len(x))) * sigma
    return x.astype(np.complex64)


def test_frame_layout():
    sf, n = 7, 128
    sym = np.array([5, 100, 77], np.uint16)
    fr = ORC.frame_modulate(sf, sym, padding=2, ampl=0.5)
    assert len(fr) == (10 + 2 + 2 + 3 + 2) * n + n // 4           # LoRaMod.cpp:138-226
    body = fr[: (14 + 3) * n + n // 4]
    np.testing.assert_allclose(np.abs(body), 0.5, rtol=1e-6)       # std::polar(ampl, phase)
    assert not fr[len(body):].any()                                # :224-226 zero padding
    # the ten preamble chirps are one chirp repeated (phase carried, wrapped to 2*pi each time)
    d = np.angle(fr[n:2 * n] * np.conj(fr[:n]))
    np.testing.assert_allclose(d, d[0], atol=2e-3)


def test_stream_tables():
    up, down, fine = ORC.stream_tables(8)
    np.testing.assert_array_equal(up, np.conj(down))               # LoRaDemod.cpp:104-105
    np.testing.assert_allclose(np.abs(fine), 1.0, rtol=1e-6)
    # fine[i] = exp(j*2*pi*(i+1)/(N*128)) up to the float step's rounding
    i = np.arange(len(fine))
    ph = np.angle(fine * np.exp(-2j * np.pi * (i + 1) / len(fine)))
    assert np.max(np.abs(ph)) < 2e-2


@pytest.mark.parametrize("sf", [7, 8, 9, 10])
@pytest.mark.parametrize("cfo_bins,lead_frac,snr_db", [(0.0, 0.0, 80.0), (0.3, 0.29, 20.0), (-2.4, 0.52, 10.0), (5.2, 0.09, 0.0)])
def test_loopback(sf, cfo_bins, lead_frac, snr_db):
    n = 1 << sf
    rng = np.random.default_rng(sf * 100 + int(snr_db))
    sym = rng.integers(0, n, 20, dtype=np.uint16)
    fr = ORC.frame_modulate(sf, sym, padding=4)
    x = channel(fr, n, 3 * n + int(lead_frac * n), 4 * n, cfo_bins, snr_db, rng)
    pk, start, err, consumed = ORC.stream_demod(sf, x, mtu=len(sym))
    assert len(pk) == 1 and len(pk[0]) == len(sym)
    np.testing.assert_array_equal(pk[0].astype(np.int64), sym.astype(np.int64))
    assert consumed <= len(x) and consumed + 2 * n > len(x)       # :148 stops when fewer than 2N samples are left
    # the data windows start one frame header after the (delayed) frame start; a carrier offset of c bins moves the
    # alignment by about c samples (a chirp cannot tell the two apart) and the down-chirps report about twice that
    want = 3 * n + int(lead_frac * n) + 14 * n + n // 4
    assert abs(int(start[0]) - want) <= abs(cfo_bins) + 3
    assert abs(int(err[0]) - 2 * cfo_bins) <= 4


def test_two_frames_and_squelch_termination():
    sf, n = 8, 256
    rng = np.random.default_rng(5)
    a = rng.integers(0, n, 12, dtype=np.uint16)
    b = rng.integers(0, n, 7, dtype=np.uint16)
    x = np.concatenate([channel(ORC.frame_modulate(sf, a, padding=3), n, 2 * n + 17, 0, 0.2, 25.0, rng),
                        channel(ORC.frame_modulate(sf, b, padding=3), n, 91, 3 * n, -0.4, 25.0, rng)])
    # mtu larger than either packet: each ends at its first squelched window (LoRaDemod.cpp:301), which is delivered too
    pk, start, err, consumed = ORC.stream_demod(sf, x, mtu=64, thresh=-10.0)
    assert len(pk) == 2
    assert len(pk[0]) == len(a) + 1 and len(pk[1]) == len(b) + 1
    np.testing.assert_array_equal(pk[0][: len(a)].astype(np.int64), a.astype(np.int64))
    np.testing.assert_array_equal(pk[1][: len(b)].astype(np.int64), b.astype(np.int64))


def test_noise_only_and_short_streams():
    sf, n = 7, 128
    rng = np.random.default_rng(9)
    x = ((rng.standard_normal(40 * n) + 1j * rng.standard_normal(40 * n)) * 0.1).astype(np.complex64)
    pk, _, _, consumed = ORC.stream_demod(sf, x, thresh=10.0)
    assert pk == [] and consumed == 39 * n                          # every window squelched: N per call
    pk, _, _, consumed = ORC.stream_demod(sf, x[: 2 * n - 1])
    assert pk == [] and consumed == 0                               # :148 needs 2N samples
    pk, _, _, consumed = ORC.stream_demod(sf, x[:0])
    assert pk == [] and consumed == 0
