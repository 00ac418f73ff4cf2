"""The stand-alone FFT and detector entry points (SURVEY.md section 8 rows a3-a5 as direct calls):
lora_b200_fft_detect_host behind kissfft<float>::transform and LoRaDetector<float>::detect of include/lora_phy.
Spectrum, index, power, powerAvg and fIndex bit for bit against the oracle (itself pinned to the unmodified
reference by tests/test_oracle_vs_ref.py)."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-_b200")


def bits(a):
    return np.ascontiguousarray(a).view(np.uint8).tobytes()


def windows(n, seed):
    r = np.random.default_rng(seed)
    noise = lambda s: ((r.standard_normal(n) + 1j * r.standard_normal(n)) * s).astype(np.complex64)
    tone = np.exp(2j * np.pi * 37.3 * np.arange(n) / n).astype(np.complex64)
    two = np.zeros(n, np.complex64); two[::2] = 1          # equal-power bins (tests/equal_power_bin_test.cpp:33-55)
    big = noise(1e-3); big[n // 3] = 1e4                    # |X|^2 terms spanning far more than 2^18: the order of the double sum matters
    imp = np.zeros(n, np.complex64); imp[0] = 1
    return [noise(1.0), noise(1e-3), noise(300.0), np.zeros(n, np.complex64), tone, (tone + noise(0.5)).astype(np.complex64),
            two, big, imp]


@pytest.mark.parametrize("sf", range(2, 13))
def test_fft_and_detector_bit_identical(sf):
    import lora_phy_b200 as L
    from _oracle import ORC
    n = 1 << sf
    ctx = L.Context(sf)
    xs = windows(n, sf)
    batch = np.stack(xs)
    fwd, inv = ctx.fft(batch), ctx.fft(batch, inverse=True)
    res, spec = ctx.detect(batch, want_spectrum=True)
    for k, x in enumerate(xs):
        assert bits(fwd[k]) == bits(ORC.fft(x)), (sf, k)
        assert bits(inv[k]) == bits(ORC.fft(x, inverse=True)), (sf, k, "inverse")
        io, po, pao, fo, so = ORC.detect(x)
        assert int(res[k]["index"]) == io, (sf, k)
        assert bits(spec[k]) == bits(so)
        # fIndex comes from exact hypot and double arithmetic: bit for bit.  power / powerAvg are 20*log10f(.) minus
        # 20*log10(N): the device's and glibc's log10f can differ in the last bit of the first term (|dB| up to ~150,
        # ulp 1.5e-5), which the subtraction carries over: within 3e-5 dB (they only ever feed comparisons).
        got = res[k]["f_index"]
        assert bits(np.float32(got)) == bits(np.float32(fo)) or (np.isnan(got) and np.isnan(fo)), (sf, k, "f_index", got, fo)
        for name, want in (("power", po), ("power_avg", pao)):
            got = np.float32(res[k][name])
            want = np.float32(want)
            same = bits(got) == bits(want) or (np.isnan(got) and np.isnan(want))
            close = np.isfinite(got) and np.isfinite(want) and abs(float(got) - float(want)) <= 3e-5
            assert same or close, (sf, k, name, got, want)


def test_equal_power_tie_rule_n4():
    """tests/equal_power_bin_test.cpp:33-55: input (1,0,1,0) at N = 4 gives two equal bins; the lower index wins."""
    import lora_phy_b200 as L
    res = L.Context(2).detect(np.array([[1, 0, 1, 0]], np.complex64))
    assert int(res[0]["index"]) == 0


@pytest.mark.parametrize("sf", [7, 10, 12])
def test_detector_header_program(tmp_path, sf):
    """LoRa-SDR/TestDetector.cpp:9-35 through this build's kissfft.hh / LoRaDetector.hpp (bin/detector_check): every
    symbol value detected, and what the header calls return carries the oracle's bits."""
    from _oracle import ORC
    exe = os.path.join(PKG, "bin", "detector_check")
    assert os.path.exists(exe), "bin/detector_check was not built"
    dump = tmp_path / "d.bin"
    stride = 1 if sf <= 10 else 13
    r = subprocess.run([exe, f"--sf={sf}", f"--dump={dump}", f"--stride={stride}"], capture_output=True, text=True)
    assert r.returncode == 0 and "detector ok" in r.stdout, (r.stdout, r.stderr)
    n = 1 << sf
    raw = np.fromfile(dump, np.float32)
    rec = 4 * n + 4
    assert len(raw) % rec == 0 and len(raw) >= rec
    for k in range(len(raw) // rec):
        blk = raw[k * rec:(k + 1) * rec]
        x, spec, out = blk[:2 * n].view(np.complex64), blk[2 * n:4 * n].view(np.complex64), blk[4 * n:]
        io, po, pao, fo, so = ORC.detect(x)
        assert int(out[0]) == io and bits(spec) == bits(so)
        assert bits(out[3:4]) == bits(np.array([fo], np.float32))
        np.testing.assert_allclose(out[1:3], np.array([po, pao], np.float32), rtol=0, atol=3e-5)
