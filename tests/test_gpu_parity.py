"""GPU parity: the CUDA path, always called through the C ABI, against (a) the committed golden fixtures
produced by the unmodified reference and (b) the CPU oracle on the same seeded inputs.

Bar: bit-exact for symbols, sync words, payload bytes, CRC flags -- and, because the kernels reproduce the
reference's float operations one for one, also bit-exact for cfo / time_offset / peak |X|^2 / modulated IQ
(the north-star tolerance for those is 1e-4 relative; the assertions below are stricter on purpose and
say so where a looser bound would be acceptable).
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def env():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import lora_phy_b200 as L
    from _oracle import ORC, Params, chain_cfg
    return {"torch": torch, "L": L, "ORC": ORC, "Params": Params, "orc_chain_cfg": chain_cfg}


def bits(a):
    return np.ascontiguousarray(a).view(np.uint8).tobytes()


def unfused_dechirp(iq, down):
    x = iq.reshape(-1, len(down))
    re = (x.real * down.real[None]).astype(np.float32) - (x.imag * down.imag[None]).astype(np.float32)
    im = (x.real * down.imag[None]).astype(np.float32) + (x.imag * down.real[None]).astype(np.float32)
    return (re + 1j * im).astype(np.complex64).reshape(-1)


@pytest.mark.parametrize("sf", [2, 5, 7, 8, 9, 10, 11, 12])
def test_host_built_tables_match_reference_init(env, sf):
    ctx = env["L"].Context(sf, window=True)
    n = 1 << sf
    assert bits(ctx.table(0)) == bits(env["ORC"].twiddles(n))
    assert bits(ctx.table(1)) == bits(env["ORC"].gen_chirp(n, 1, n, 0.0, True, 1.0, 0.0, 1.0)[0])
    if sf >= 7:
        g = np.load(os.path.join(G, "awgn_packets.npz"))
        assert bits(ctx.table(0)) == bits(g[f"sf{sf}_twiddles"]) and bits(ctx.table(1)) == bits(g[f"sf{sf}_downchirp"])


def test_config1_golden_capture_rx_runner_path(env):
    """BASELINE config 1: the committed SF7 capture through demodulate() + decode()."""
    L, torch = env["L"], env["torch"]
    g = np.load(os.path.join(G, "capture_sf7.npz"))
    ctx = L.Context(7)
    ret, sym, m = ctx.demodulate(g["iq"])
    assert ret == 2086 and bits(sym) == bits(g["symbols"])
    assert bits(m["cfo"]) == bits(g["cfo"]) and bits(m["time_offset"]) == bits(g["time_offset"])
    assert int(m["sync_word"]) == int(g["sync_word"])
    d_sym = torch.from_numpy(sym.astype(np.int16)).cuda()[None, :]
    out, ok = ctx.decode_batch(d_sym)
    assert bits(out.cpu().numpy()[0]) == bits(g["payload"]) and bool(ok.item()) == bool(g["crc_ok"])
    # guards of phy.cpp:181-190
    assert ctx.demodulate(g["iq"][:-1])[0] == -1 and ctx.demodulate(g["iq"][:128])[0] == -1


def test_interop_golden_osr2(env):
    g = np.load(os.path.join(G, "interop_sf7_osr2.npz"))
    ctx = env["L"].Context(7, osr=2)
    ret, sym, sync, m = ctx.lora_demodulate(g["iq"])
    assert bits(sym) == bits(g["symbols"]) and sync == 0x29 == int(g["sync"])
    assert bits(m["cfo"]) == bits(g["cfo"]) and bits(m["time_offset"]) == bits(g["time_offset"])


@pytest.mark.parametrize("sf", range(7, 13))
def test_awgn_goldens_all_three_modes(env, sf):
    g = np.load(os.path.join(G, "awgn_packets.npz"))
    ctx = env["L"].Context(sf)
    down = g[f"sf{sf}_downchirp"]
    for k in range(3):
        t = f"sf{sf}_{k}"
        iq = g[f"{t}_iq"]
        ret, sym, m = ctx.demodulate(iq)
        assert bits(sym) == bits(g[f"{t}_demodulate"]), (t, "demodulate")
        assert bits(np.array([m["cfo"], m["time_offset"]], np.float32)) == bits(g[f"{t}_demodulate_m"])
        assert int(m["sync_word"]) == int(g[f"{t}_demodulate_sync"])
        for fused in (True, False):
            x = iq if fused else unfused_dechirp(iq, down)
            ret, sym, sync, m = ctx.lora_demodulate(x, dechirp=fused)
            assert bits(sym) == bits(g[f"{t}_legacy"]) and sync == int(g[f"{t}_legacy_sync"]), (t, fused)
            assert bits(np.array([m["cfo"], m["time_offset"]], np.float32)) == bits(g[f"{t}_legacy_m"])
    assert bits(ctx.lora_modulate(np.arange(3, dtype=np.uint16) * 37 % (1 << sf))) == bits(g[f"sf{sf}_modulated"])


@pytest.mark.parametrize("sf,bw,osr,window", [(7, 125000, 1, False), (8, 125000, 1, True), (9, 250000, 1, False),
                                               (10, 125000, 1, False), (11, 125000, 1, True), (12, 500000, 1, False),
                                               (7, 125000, 2, False), (6, 125000, 1, False), (4, 125000, 1, False)])
def test_batch_vs_oracle_seeded(env, sf, bw, osr, window):
    """Synthetic AWGN packets made on the device, every packet re-run through the CPU oracle."""
    L, torch, ORC, Params = env["L"], env["torch"], env["ORC"], env["Params"]
    ctx = L.Context(sf, bw_hz=bw, osr=osr, window=window)
    n_pk, nsym = 48, (9 if sf <= 10 else 5)
    rng = np.random.default_rng(sf * 31 + osr)
    tx = rng.integers(0, 1 << sf, (n_pk, nsym), dtype=np.uint16)
    d_tx = torch.from_numpy(tx.astype(np.int16)).cuda()
    iq = ctx.modulate_batch(d_tx)
    snr = torch.from_numpy(rng.choice([-20.0, -10.0, 0.0, 10.0], n_pk).astype(np.float32)).cuda()
    ctx.awgn_batch(iq, snr, seed=sf)
    h_iq = iq.cpu().numpy()
    # the modulator itself: bit-exact IQ before noise (north-star bound would be 1e-4 relative)
    clean = ctx.modulate_batch(d_tx).cpu().numpy()
    for p in (0, n_pk - 1):
        assert bits(clean[p]) == bits(ORC.lora_modulate(tx[p], sf, osr, bw))
    for mode, precision in [(m, pr) for m in (L.DEMOD_WORKSPACE, L.DEMOD_LEGACY, L.DEMOD_LEGACY_DECHIRP)
                            for pr in (L.PRECISION_EXACT, L.PRECISION_FAST)]:
        if mode == L.DEMOD_LEGACY_DECHIRP and osr != 1:
            continue
        ctx.set_precision(precision)
        sym, met, peaks = ctx.demod_batch(iq, mode=mode, want_peaks=True)
        sym = sym.cpu().numpy().view(np.uint16)
        met = met.cpu().numpy().view(L.PKT_METRICS).reshape(-1)
        peaks = peaks.cpu().numpy()
        for p in range(n_pk):
            if mode == L.DEMOD_WORKSPACE:
                ret, so, mo, pk = ORC.demodulate(Params(sf, bw, osr, int(window), int(window), 0x12), h_iq[p], want_peaks=True)
                sync_o = mo.sync_word
            else:
                x = h_iq[p]
                if mode == L.DEMOD_LEGACY_DECHIRP:
                    x = unfused_dechirp(x, ORC.gen_chirp(1 << sf, 1, 1 << sf, 0.0, True, 1.0, 0.0, bw / 125000.0)[0])
                ret, so, sync_o, mo, pk = ORC.lora_demodulate(sf, x, osr=osr, hann=window, want_peaks=True)
            assert bits(sym[p]) == bits(so), (sf, mode, p)
            assert int(met[p]["sync_word"]) == sync_o
            assert bits(met[p]["cfo"]) == bits(np.float32(mo.cfo)), (sf, mode, p)
            assert bits(met[p]["time_offset"]) == bits(np.float32(mo.time_offset))
            if precision == L.PRECISION_EXACT or sf < 7:
                assert bits(peaks[p]) == bits(pk), (sf, mode, p)
            else:  # BASELINE.json: FFT peak magnitudes within 1e-4 relative in fp32 (|X|, so 2e-4 on |X|^2)
                np.testing.assert_allclose(np.sqrt(peaks[p]), np.sqrt(pk), rtol=1e-4, atol=0)


def test_legacy_edge_cases(env):
    L, ORC = env["L"], env["ORC"]
    ctx = L.Context(7)
    rng = np.random.default_rng(5)
    y = ((rng.standard_normal(640) + 1j * rng.standard_normal(640)) * 3).astype(np.complex64)
    assert ctx.lora_demodulate(y, scratch=False)[0] == 0                 # LoRaDemod.cpp:69-71
    for count in (128, 192, 256, 389):                                    # < 2 symbols, ragged tails
        ret, sym, sync, m = ctx.lora_demodulate(y[:count])
        ro, so, sy, mo = ORC.lora_demodulate(7, y[:count])
        assert ret == ro and bits(sym) == bits(so) and sync == sy
        assert bits(m["cfo"]) == bits(np.float32(mo.cfo)) and bits(m["time_offset"]) == bits(np.float32(mo.time_offset))
    z = np.zeros(512, np.complex64)                                      # all-zero windows -> index 0
    ret, sym, sync, m = ctx.lora_demodulate(z)
    assert list(sym) == [0, 0] and sync == 0
    c2 = L.Context(2)                                                     # tests/equal_power_bin_test.cpp
    assert list(c2.lora_demodulate(np.array([1, 0, 1, 0], np.complex64))[1]) == [0]
    co, to = ctx.estimate_offsets(y[:512])
    po = env["Params"](7, 125000, 1, 0, 0, 0x12)
    eo = ORC.estimate_offsets(po, y[:512])
    assert bits(co) == bits(eo[0]) and bits(to) == bits(eo[1])


@pytest.mark.parametrize("sf", [7, 9, 12])
def test_normalisation_scan_nan_inf(env, sf):
    """LoRaDemod.cpp:59-77 compares, it does not call fmax: a NaN real part drops the whole sample, a NaN imaginary
    part only itself.  The kernel's fast scan assumes finite data and must fall back to the exact order otherwise."""
    L, ORC = env["L"], env["ORC"]
    ctx = L.Context(sf)
    n = 1 << sf
    rng = np.random.default_rng(sf)
    base = ((rng.standard_normal(6 * n + 37) + 1j * rng.standard_normal(6 * n + 37)) * 2).astype(np.complex64)

    def expect(y):
        mx = np.float32(0)
        for v in y:  # the reference's loop, literally
            a, b = np.abs(np.float32(v.real)), np.abs(np.float32(v.imag))
            m = b if a < b else a
            if m > mx:
                mx = m
        return mx

    cases = []
    y = base.copy(); y[3 * n + 5] = complex(np.nan, 50.0); cases.append(y)       # dropped: 50 must not win
    y = base.copy(); y[4 * n + 9] = complex(40.0, np.nan); cases.append(y)       # 40 wins
    y = base.copy(); y[2 * n + 1] = complex(np.nan, 50.0); y[5 * n + 2] = complex(-60.0, 1.0); cases.append(y)
    y = base.copy(); y[6 * n + 30] = complex(np.nan, 70.0); cases.append(y)      # in the ragged tail
    y = base.copy(); y[3 * n] = complex(1.0, -np.inf); cases.append(y)           # Inf wins, scale = 0
    for y in cases + [base]:
        ret, sym, sync, m = ctx.lora_demodulate(y)
        want = expect(y)
        assert bits(np.float32(m["max_amp"])) == bits(np.float32(want)), (sf, m["max_amp"], want)
        ro, so, sy, mo = ORC.lora_demodulate(sf, y)
        assert ret == ro


def test_time_offset_shifts_both_directions(env):
    L, ORC, Params = env["L"], env["ORC"], env["Params"]
    ctx = L.Context(7, osr=2)
    rng = np.random.default_rng(9)
    y = ORC.lora_modulate(rng.integers(0, 128, 10, dtype=np.uint16), 7, osr=2)
    y = (y + (rng.standard_normal(len(y)) + 1j * rng.standard_normal(len(y))) * 0.3).astype(np.complex64)
    seen = set()
    for shift in (0, -17, 9, 40, -77):
        z = np.roll(y, shift)
        ret, sym, m = ctx.demodulate(z)
        ro, so, mo = ORC.demodulate(Params(7, 125000, 2, 0, 0, 0x12), z)
        assert ret == ro and bits(sym) == bits(so) and bits(m["time_offset"]) == bits(np.float32(mo.time_offset))
        seen.add(int(np.sign(round(float(mo.time_offset)))))
    assert len(seen) >= 2  # the shifted-window branches of phy.cpp:205-212 were really exercised


@pytest.mark.parametrize("sf", [7, 9, 12])
def test_code_chain_vs_oracle(env, sf):
    L, torch, ORC, ocfg = env["L"], env["torch"], env["ORC"], env["orc_chain_cfg"]
    ctx = L.Context(sf)
    rng = np.random.default_rng(sf)
    for rdd in range(5):
        for ppm in (0, sf - 2):
            for plen, crc, explicit in ((16, True, True), (1, True, True), (255, True, True), (33, False, True),
                                        (20, True, False), (20, False, False)):
                n = 24
                pay = rng.integers(0, 256, (n, plen), dtype=np.uint8)
                kw = dict(ppm=ppm, explicit=explicit, crc=crc, data_length=plen)
                cfg, oc = L.chain_cfg(sf, rdd, **kw), ocfg(sf, rdd, **kw)
                sym = ctx.chain_encode_batch(cfg, torch.from_numpy(pay).cuda())
                h_sym = sym.cpu().numpy().view(np.uint16)
                for p in range(n):
                    assert bits(h_sym[p]) == bits(ORC.chain_encode(oc, bytes(pay[p]))), (sf, rdd, ppm, plen)
                # corrupt a third of the packets so drops / corrections / CRC failures all occur
                bad = h_sym.copy()
                for p in range(0, n, 3):
                    for _ in range(1 + p % 4):
                        bad[p, rng.integers(0, bad.shape[1])] ^= np.uint16(1 << rng.integers(0, sf))
                for ec in (False, True):
                    cfg2, oc2 = L.chain_cfg(sf, rdd, error_check=ec, **kw), ocfg(sf, rdd, error_check=ec, **kw)
                    out, res = ctx.chain_decode_batch(cfg2, torch.from_numpy(bad.view(np.int16)).cuda())
                    out = out.cpu().numpy()
                    res = res.cpu().numpy().view(L.CHAIN_RESULT).reshape(-1)
                    for p in range(n):
                        ro, bo, info = ORC.chain_decode(oc2, bad[p])
                        assert int(res[p]["n_bytes"]) == ro, (sf, rdd, ppm, plen, crc, explicit, ec, p)
                        assert int(res[p]["drop_line"]) == info.drop_line
                        if ro >= 0:
                            assert bits(out[p, :ro]) == bits(bo)
                        for f in ("parity_error", "uncorrectable", "hdr_length", "hdr_has_crc", "hdr_rdd",
                                  "hdr_checksum", "crc_present", "crc_ok"):
                            assert int(res[p][f]) == getattr(info, f), (f, sf, rdd, p)
    # short input: fewer than 8 symbols -> no output (:208)
    short = torch.zeros((2, 7), dtype=torch.int16, device="cuda")
    _, res = ctx.chain_decode_batch(L.chain_cfg(sf, 1), short)
    assert list(res.cpu().numpy().view(L.CHAIN_RESULT).reshape(-1)["n_bytes"]) == [-2, -2]


def test_simple_encode_decode_kat(env):
    L, torch = env["L"], env["torch"]
    ctx = L.Context(7)
    data = torch.tensor([[0xDE, 0xAD, 0xBE, 0xEF]], dtype=torch.uint8, device="cuda")
    sym = ctx.encode_batch(data)
    assert list(sym.cpu().numpy().view(np.uint16)[0]) == [0x8D, 0x2E, 0x9A, 0x8D, 0x4B, 0x2E, 0x2E, 0xFF]
    out, ok = ctx.decode_batch(sym)
    assert bytes(out.cpu().numpy()[0]) == b"\xde\xad\xbe\xef"


@pytest.mark.parametrize("sf", [7, 10])
def test_full_size_round_trip_property(env, sf):
    """BASELINE config 2 size (65,536 packets at SF7): encode -> modulate -> noise -> demod -> decode on the
    device.  Size-independent properties instead of a per-packet oracle run: every packet at >= 0 dB decodes
    to the bytes sent with CRC ok, and a checksum of all symbols equals that of a second, independent run."""
    L, torch = env["L"], env["torch"]
    n = 65536 if sf == 7 else 4096
    ctx = L.Context(sf)
    cfg = L.chain_cfg(sf, 1)
    gen = torch.Generator(device="cuda").manual_seed(sf)
    pay = torch.randint(0, 256, (n, 16), dtype=torch.uint8, device="cuda", generator=gen)
    sym = ctx.chain_encode_batch(cfg, pay)
    iq = ctx.modulate_batch(sym)
    snr = torch.full((n,), 0.0, dtype=torch.float32, device="cuda")
    snr[::2] = 10.0
    ctx.awgn_batch(iq, snr, seed=77)
    rx, met, _ = ctx.demod_batch(iq, mode=L.DEMOD_LEGACY_DECHIRP)
    out, res = ctx.chain_decode_batch(cfg, rx.contiguous())
    res_h = res.cpu().numpy().view(L.CHAIN_RESULT).reshape(-1)
    assert (res_h["n_bytes"] == 16).all() and (res_h["crc_ok"] == 1).all()
    assert torch.equal(out[:, :16], pay)
    assert torch.equal(rx, sym)
    rx2, _, _ = ctx.demod_batch(iq, mode=L.DEMOD_LEGACY_DECHIRP)
    assert int(rx.to(torch.int64).sum()) == int(rx2.to(torch.int64).sum())


@pytest.mark.parametrize("sf", range(7, 13))
def test_fast_precision_filter(env, sf):
    """PRECISION_FAST = fused kernel + bit-faithful recheck of near-tie windows.  On mostly-noise input (where
    near ties are common) the symbols must equal PRECISION_EXACT's on every window, the peaks must agree far
    inside the 1e-4 budget, and the recheck rate must be the small fraction the margin predicts."""
    L, torch = env["L"], env["torch"]
    ctx = L.Context(sf)
    n_pk, nsym = (2048 >> max(0, sf - 9)), 14
    gen = torch.Generator(device="cuda").manual_seed(100 + sf)
    tx = torch.randint(0, 1 << sf, (n_pk, nsym), dtype=torch.int16, device="cuda", generator=gen)
    iq = ctx.modulate_batch(tx)
    snr = torch.tensor([-30.0, -20.0, -12.0, 0.0], device="cuda")[torch.arange(n_pk, device="cuda") % 4]
    ctx.awgn_batch(iq, snr.contiguous(), seed=sf)
    for mode in (L.DEMOD_LEGACY_DECHIRP, L.DEMOD_WORKSPACE):
        ctx.set_precision(L.PRECISION_EXACT)
        s0, m0, p0 = ctx.demod_batch(iq, mode=mode, want_peaks=True)
        ctx.set_precision(L.PRECISION_FAST)
        ctx.recheck_stats()
        s1, m1, p1 = ctx.demod_batch(iq, mode=mode, want_peaks=True)
        windows, rechecked = ctx.recheck_stats()
        assert torch.equal(s0, s1), (sf, mode, int((s0 != s1).sum()))
        assert torch.equal(m0, m1)
        rel = ((p1 - p0).abs() / p0).max().item()
        assert rel < 2e-5, (sf, mode, rel)
        assert windows == n_pk * (nsym + 2) and 0 < rechecked < 0.02 * windows, (sf, mode, windows, rechecked)


def test_compensate_encode_decode_host_entry_points(env):
    """compensate_offsets (phy.cpp:147-176) against the oracle in both shift directions and with a shift that is
    too large to apply; encode / decode with host buffers against the oracle's Hamming(8,4) pairs and CRC flag."""
    L, ORC, Params = env["L"], env["ORC"], env["Params"]
    rng = np.random.default_rng(11)
    for sf, osr in ((7, 1), (9, 2)):
        ctx = L.Context(sf, osr=osr)
        y = (rng.standard_normal(1000) + 1j * rng.standard_normal(1000)).astype(np.complex64)
        for cfo, to in ((0.0, 0.0), (0.31, 5.4), (-0.07, -17.5), (0.5, 2000.0), (0.002, -999.4)):
            got = ctx.compensate_offsets(cfo, to, y)
            want = ORC.compensate_offsets(Params(sf, 125000, osr, 0, 0, 0x12), cfo, to, y)
            assert bits(got) == bits(want), (sf, osr, cfo, to)
    ctx = L.Context(7)
    data = rng.integers(0, 256, 37, dtype=np.uint8)
    sym = ctx.encode(data)
    assert bits(sym) == bits(ORC.lora_encode(data))
    sym[5] ^= 0x4  # one correctable bit error
    out, ok = ctx.decode(sym)
    o_out, o_ok = ORC.decode(sym)
    assert bits(out) == bits(o_out) and ok == bool(o_ok)


@pytest.mark.parametrize("sf", [7, 9, 10, 11, 12])
def test_repeatable_under_load(env, sf):
    """The kernels reuse one shared-memory exchange region window after window behind hand-placed barriers (the
    end-of-window barrier is gone: the peak search is the rendezvous).  compute-sanitizer is not available on the
    GPU pool, so a missing barrier has to show up as what it would cause: results that change from run to run, or
    differ between a loaded GPU (all CTAs resident, every slot busy) and the oracle."""
    torch, L, ORC = env["torch"], env["L"], env["ORC"]
    ctx = L.Context(sf)
    n_pk, nsym = (8192 >> max(0, sf - 8)), 10
    gen = torch.Generator(device="cuda").manual_seed(7 * sf)
    tx = torch.randint(0, 1 << sf, (n_pk, nsym), dtype=torch.int16, device="cuda", generator=gen)
    iq = ctx.modulate_batch(tx)
    snr = torch.full((n_pk,), -14.0, dtype=torch.float32, device="cuda")   # near-ties are common at this SNR
    ctx.awgn_batch(iq, snr, seed=sf)
    down = ORC.gen_chirp(1 << sf, 1, 1 << sf, 0.0, True, 1.0, 0.0, 1.0)[0]
    for precision in (L.PRECISION_FAST, L.PRECISION_EXACT):
        ctx.set_precision(precision)
        first = None
        for rep in range(12):
            rx, met, pk = ctx.demod_batch(iq, mode=L.DEMOD_LEGACY_DECHIRP, want_peaks=True)
            cur = (rx.clone(), met.clone(), pk.clone())
            if first is None:
                first = cur
            else:
                for a, b in zip(first, cur):
                    assert torch.equal(a.view(torch.uint8), b.view(torch.uint8)), (sf, precision, rep)
        rx_h = first[0].cpu().numpy().view(np.uint16)
        h_iq = iq[:: max(1, n_pk // 24)].cpu().numpy()
        for k, p in enumerate(range(0, n_pk, max(1, n_pk // 24))):
            _, so, _, _ = ORC.lora_demodulate(sf, unfused_dechirp(h_iq[k], down))
            assert bits(rx_h[p]) == bits(so), (sf, precision, p)


def test_int16_ingest_equals_float_path(env):
    """int16 I/Q in, converted on the device (exactly, scale 2^-15), then the float path."""
    torch, L = env["torch"], env["L"]
    for sf in (7, 10):
        ctx = L.Context(sf)
        cfg = L.chain_cfg(sf, 1)
        gen = torch.Generator(device="cuda").manual_seed(sf)
        pay = torch.randint(0, 256, (33, 16), dtype=torch.uint8, device="cuda", generator=gen)
        iq = ctx.modulate_batch(ctx.chain_encode_batch(cfg, pay), amplitude=0.25)
        ctx.awgn_batch(iq, torch.full((33,), 20.0, device="cuda"), seed=sf)   # sigma is relative to amplitude 1
        h = iq.cpu().numpy()
        q = np.stack([h.real, h.imag], axis=-1)
        q = np.clip(np.round(q * 8192.0), -32768, 32767).astype(np.int16)           # a 14-bit ADC's view of the samples
        as_float = (q.astype(np.float32) * np.float32(1.0 / 32768.0))
        as_float = (as_float[..., 0] + 1j * as_float[..., 1]).astype(np.complex64)
        p1, r1, m1 = ctx.receive_host_i16(cfg, q)
        p2, r2, m2 = ctx.receive_host(cfg, as_float)
        assert bits(p1) == bits(p2) and bits(r1) == bits(r2) and bits(m1) == bits(m2)
        assert (r1["crc_ok"] == 1).all() and bytes(p1[:, :16]) == bytes(pay.cpu().numpy())


def test_int8_ingest_equals_float_path(env):
    """lora_b200_receive_host_i8: int8 I/Q (a quarter of the PCIe bytes) converted exactly on the device; results equal
    the float call on the converted samples, ragged (non multiple of eight) sample counts included."""
    torch, L = env["torch"], env["L"]
    for sf, amp in ((7, 0.5), (9, 0.4)):
        ctx = L.Context(sf)
        cfg = L.chain_cfg(sf, 1)
        gen = torch.Generator(device="cuda").manual_seed(40 + sf)
        pay = torch.randint(0, 256, (21, 16), dtype=torch.uint8, device="cuda", generator=gen)
        iq = ctx.modulate_batch(ctx.chain_encode_batch(cfg, pay), amplitude=amp)
        ctx.awgn_batch(iq, torch.full((21,), 25.0, device="cuda"), seed=sf)
        h = iq.cpu().numpy()
        q = np.clip(np.round(np.stack([h.real, h.imag], axis=-1) * 96.0), -128, 127).astype(np.int8)
        as_float = q.astype(np.float32) * np.float32(1.0 / 128.0)
        as_float = (as_float[..., 0] + 1j * as_float[..., 1]).astype(np.complex64)
        p1, r1, m1 = ctx.receive_host_i8(cfg, q)
        p2, r2, m2 = ctx.receive_host(cfg, as_float)
        assert bits(p1) == bits(p2) and bits(r1) == bits(r2) and bits(m1) == bits(m2)
        assert (r1["crc_ok"] == 1).all() and bytes(p1[:, :16]) == bytes(pay.cpu().numpy())
    # a lone ragged buffer: the tail path of the converter
    ctx = L.Context(7)
    rng = np.random.default_rng(1)
    q = rng.integers(-128, 128, (1, 3 * 128, 2), dtype=np.int8)
    f = (q[..., 0].astype(np.float32) + 1j * q[..., 1].astype(np.float32)).astype(np.complex64) * np.float32(1 / 128.0)
    a = ctx.receive_host_i8(L.chain_cfg(7, 1), q[:, :381], mode=L.DEMOD_LEGACY)   # this entry point takes ragged tails
    b = ctx.receive_host(L.chain_cfg(7, 1), np.ascontiguousarray(f[:, :381]), mode=L.DEMOD_LEGACY)
    assert all(bits(x) == bits(y) for x, y in zip(a, b))


def test_stream_pipelined_sub_batches_give_the_same_results(env):
    """lora_b200_set_overlap: scan+estimate of sub-batch k+1 on a second stream beside the demod of sub-batch k (only
    taken for batches of 256 MB and more).  Off by default because it is no faster; it must still be exact."""
    torch, L = env["torch"], env["L"]
    sf = 9
    ctx = L.Context(sf)
    gen = torch.Generator(device="cuda").manual_seed(99)
    tx = torch.randint(0, 1 << sf, (4608, 28), dtype=torch.int16, device="cuda", generator=gen)
    iq = ctx.modulate_batch(tx)
    assert iq.numel() * 8 >= 256 << 20
    ctx.awgn_batch(iq, torch.full((4608,), -12.0, device="cuda"), seed=5)
    out = []
    for on in (False, True):
        ctx.set_overlap(on)
        rx, met, pk = ctx.demod_batch(iq, mode=L.DEMOD_LEGACY_DECHIRP, want_peaks=True)
        torch.cuda.synchronize()
        out.append((rx.clone(), met.clone(), pk.clone()))
    for a, b in zip(*out):
        assert torch.equal(a.view(torch.uint8), b.view(torch.uint8))


@pytest.mark.gpu
@pytest.mark.parametrize("sf,mode_name", [(7, "dechirp"), (9, "dechirp"), (9, "legacy"), (12, "dechirp")])
def test_overlapped_scan_equals_two_pass_path(env, sf, mode_name):
    """Large legacy batches run the normalisation scan of sub-batch k+1 (csrc/scan_kernels.cu: bulk-copy staged, a
    floating-point filter in front of the reference's comparisons) beside the demod kernel of sub-batch k.  Symbols,
    sync words, cfo, time offset and max_amp must equal the single-stream path bit for bit, including packets that are
    all zero, carry NaN / Inf (LoRaDemod.cpp:59-77 compares, it does not call fmax), have their largest component far
    above the rest, or -- plain legacy mode -- end in a ragged tail with an odd sample count."""
    torch, L = env["torch"], env["L"]
    n = 1 << sf
    mode = L.DEMOD_LEGACY_DECHIRP if mode_name == "dechirp" else L.DEMOD_LEGACY
    nsym = 24
    count = (nsym + 2) * n + (0 if mode_name == "dechirp" else n // 2 + 1)       # odd, ragged in legacy mode
    packets = max(64, -(-(300 << 20) // (count * 8)))                             # >= 256 MB so the call pipelines
    ctx = L.Context(sf)
    g = torch.Generator(device="cuda").manual_seed(sf)
    sym = torch.randint(0, n, (packets, nsym), dtype=torch.int16, device="cuda", generator=g)
    iq = torch.zeros((packets, count + (count & 1)), dtype=torch.complex64, device="cuda")[:, :count]
    iq[:, :(nsym + 2) * n] = ctx.modulate_batch(sym)
    snr = (torch.rand(packets, device="cuda", generator=g) * 30.0 - 20.0).float().contiguous()
    ctx.awgn_batch(iq, snr, seed=sf)
    iq[1] = 0                                                                    # all-zero packet
    iq[2, 3 * n + 5] = complex(float("nan"), 50.0)                               # NaN real part drops the sample
    iq[3, 4 * n + 9] = complex(40.0, float("nan"))                               # NaN imaginary part only itself
    iq[4, 7 * n] = complex(1.0, float("-inf"))
    iq[5] *= 1e-3; iq[5, count - 1] = complex(0.0, -77.0)                         # the very last sample decides
    iq[6] *= 0.01                                                                 # never normalised (max < 1)
    iq[7] *= 1e20; iq[8] *= 1e-25                                                 # |x|^2 overflows / underflows in float
    assert iq.stride(0) % 2 == 0 and iq.data_ptr() % 16 == 0
    out = {}
    for overlap in (0, 1):
        ctx.set_overlap(overlap)
        s, m, pk = ctx.demod_batch(iq, mode=mode, want_peaks=True)
        torch.cuda.synchronize()
        out[overlap] = (s.cpu().numpy().copy(), m.cpu().numpy().copy(), pk.cpu().numpy().copy())
    ctx.set_overlap(1)
    assert bits(out[0][1]) == bits(out[1][1]), "metrics (cfo, time offset, max_amp, sync) differ"
    assert bits(out[0][0]) == bits(out[1][0]), "symbols differ"
    a, b = out[0][2], out[1][2]
    assert ((a == b) | (np.isnan(a) & np.isnan(b))).all(), "peaks differ"
    met = out[1][1].view(L.PKT_METRICS).reshape(-1)
    # The equality above is the test.  The spot checks below pin what the planted samples mean in legacy mode, where
    # the scan sees the samples themselves; in fused-dechirp mode it sees iq * down[n], where a NaN in either component
    # poisons both, so the planted (40, NaN) never counts and the maxima are whatever the noise left.
    assert met["max_amp"][1] == 0 and np.isinf(met["max_amp"][4])
    if mode_name == "legacy":
        assert met["max_amp"][5] == 77.0 and met["max_amp"][3] >= 40.0
        assert met["normalised"][6] == 0 and met["normalised"][0] == 1


@pytest.mark.gpu
@pytest.mark.parametrize("sf", [7, 8, 9, 10, 11, 12])
@pytest.mark.parametrize("mode_name", ["dechirp", "legacy", "workspace"])
def test_merged_launch_equals_two_launch_path(env, sf, mode_name):
    """lora_b200_set_overlap(3) (mode 2 without its size threshold): prepass items and demod items of one persistent
    launch (csrc/demod_fast.cu receive_merged_kernel), the estimate handed from one CTA to another through a release/acquire flag per packet.
    Symbols, sync words, cfo, time offset, max_amp and peaks must equal the two-launch path bit for bit -- with a packet
    count that fills neither the last group of 32 nor the last work item, and with all-zero, NaN, Inf, huge and tiny
    packets among them.  Run twice: the second call reuses the queue and the flags of the first."""
    torch, L = env["torch"], env["L"]
    n = 1 << sf
    mode = {"dechirp": L.DEMOD_LEGACY_DECHIRP, "legacy": L.DEMOD_LEGACY, "workspace": L.DEMOD_WORKSPACE}[mode_name]
    nsym = 21
    count = (nsym + 2) * n + (n // 2 + 1 if mode_name == "legacy" else 0)
    packets = {7: 2099, 8: 1301, 9: 1001, 10: 777, 11: 515, 12: 301}[sf]
    ctx = L.Context(sf)
    g = torch.Generator(device="cuda").manual_seed(1000 + sf)
    sym = torch.randint(0, n, (packets, nsym), dtype=torch.int16, device="cuda", generator=g)
    iq = torch.zeros((packets, count + (count & 1)), dtype=torch.complex64, device="cuda")[:, :count]
    iq[:, :(nsym + 2) * n] = ctx.modulate_batch(sym)
    snr = (torch.rand(packets, device="cuda", generator=g) * 30.0 - 20.0).float().contiguous()
    ctx.awgn_batch(iq, snr, seed=sf)
    iq[1] = 0
    iq[2, 3 * n + 5] = complex(float("nan"), 50.0)
    iq[3, 4 * n + 9] = complex(40.0, float("nan"))
    iq[4, 7 * n] = complex(1.0, float("-inf"))
    iq[5] *= 1e-3
    iq[7] *= 1e20; iq[8] *= 1e-25
    iq[packets - 1] *= 3.0
    out = {}
    total = count // n

    def guarded(rows, cols, dtype):
        """A [rows, cols] view in the middle of a buffer whose margins hold a pattern (compute-sanitizer is not
        available on the GPU pool: a write outside the view shows up as a damaged margin)."""
        pad = 4096
        flat = torch.full((rows * cols + 2 * pad,), 23130 if dtype != torch.uint8 else 0x5A, dtype=dtype, device="cuda")
        return flat, flat[pad:pad + rows * cols].view(rows, cols), pad

    for overlap in (0, 3, 3):
        ctx.set_overlap(overlap)
        ctx.recheck_stats()
        fs, s_buf, pad = guarded(packets, total - 2, torch.int16)
        fm, m_buf, _ = guarded(packets, 16, torch.uint8)
        fp, p_buf, _ = guarded(packets, total, torch.float32)
        s_buf.zero_(); m_buf.zero_(); p_buf.zero_()
        s, m, pk = ctx.demod_batch(iq, mode=mode, want_peaks=True, symbols=s_buf, metrics=m_buf, peaks=p_buf)
        torch.cuda.synchronize()
        for f, fill in ((fs, 23130), (fm, 0x5A), (fp, 23130.0)):
            assert bool((f[:pad] == fill).all()) and bool((f[-pad:] == fill).all()), "write outside an output buffer"
        got = (s.cpu().numpy().copy(), m.cpu().numpy().copy(), pk.cpu().numpy().copy(), ctx.recheck_stats())
        if overlap == 0:
            out[0] = got
            continue
        assert bits(out[0][1]) == bits(got[1]), "metrics differ"
        assert bits(out[0][0]) == bits(got[0]), "symbols differ"
        a, b = out[0][2], got[2]
        assert ((a == b) | (np.isnan(a) & np.isnan(b))).all(), "peaks differ"
        assert out[0][3] == got[3], "a different set of windows was rechecked"
    ctx.set_overlap(2)
    if mode_name == "dechirp":   # the mode the bench runs; the others see input their estimator was not written for
        clean = (snr >= 0.0).cpu().numpy()
        clean[:10] = False
        assert (out[0][0][clean] == sym.cpu().numpy()[clean]).mean() > 0.9


@pytest.mark.gpu
def test_chunked_host_receive_equals_device_path(env):
    """SURVEY 8(f) rank 2, second half: lora_b200_receive_host* cut the batch into chunks whose PCIe copy overlaps the
    kernels of the chunk before (csrc/capi.cu receive_host_chunked).  With enough bytes for several chunks (and a packet
    count that does not divide evenly) payloads, results and metrics must equal the device-resident calls bit for bit,
    for float32, int16 and int8 samples, from pinned and from pageable host memory."""
    torch, L = env["torch"], env["L"]
    sf, packets = 9, 5003                                        # 5003 x 35 x 512 x 8 B = 717 MB: eight ragged chunks
    ctx = L.Context(sf)
    cfg = L.chain_cfg(sf, 1)
    gen = torch.Generator(device="cuda").manual_seed(77)
    pay = torch.randint(0, 256, (packets, 16), dtype=torch.uint8, device="cuda", generator=gen)
    iq = ctx.modulate_batch(ctx.chain_encode_batch(cfg, pay), amplitude=0.25)
    ctx.awgn_batch(iq, (torch.rand(packets, device="cuda", generator=gen) * 25.0 - 5.0).float().contiguous(), seed=3)
    q16 = torch.view_as_real(iq).mul(8192.0).round_().clamp_(-32768, 32767).to(torch.int16)
    q8 = torch.view_as_real(iq).mul(96.0).round_().clamp_(-128, 127).to(torch.int8)
    for name, host, dev in (("f32", iq, iq),
                            ("i16", q16, torch.view_as_complex(q16.float().mul(1.0 / 32768.0).contiguous())),
                            ("i8", q8, torch.view_as_complex(q8.float().mul(1.0 / 128.0).contiguous()))):
        rx, met, _ = ctx.demod_batch(dev.contiguous(), mode=L.DEMOD_LEGACY_DECHIRP)
        out, res = ctx.chain_decode_batch(cfg, rx.contiguous())
        want = (out.cpu().numpy(), res.cpu().numpy(), met.cpu().numpy())
        for pinned in (True, False):
            h = torch.empty(host.shape, dtype=host.dtype, pin_memory=pinned)
            h.copy_(host)
            hn = h.numpy()
            if name == "f32":
                p, r, m = ctx.receive_host(cfg, hn)
            elif name == "i16":
                p, r, m = ctx.receive_host_i16(cfg, hn)
            else:
                p, r, m = ctx.receive_host_i8(cfg, hn)
            assert bits(p) == bits(want[0]) and bits(r) == bits(want[1]) and bits(m) == bits(want[2]), (name, pinned)
    assert (want[1].view(L.CHAIN_RESULT)["crc_ok"] == 1).mean() > 0.5


@pytest.mark.gpu
@pytest.mark.parametrize("sf", [9, 10, 11, 12])
def test_all_tie_windows_are_rechecked_once_and_resolved_exactly(env, sf):
    """Every window a lone impulse: a flat spectrum, i.e. (nearly) every bin ties for the maximum.  The fused flavour
    cannot trust any of them; each window must enter the recheck list exactly once (one elected reporter per window,
    so the list cannot overflow) and come back with the reference's lowest-index winner."""
    torch, L, ORC = env["torch"], env["L"], env["ORC"]
    n = 1 << sf
    packets, total = 96, 12
    x = np.zeros((packets, total * n), np.complex64)
    rng = np.random.default_rng(sf)
    for p in range(packets):
        for s in range(total):
            x[p, s * n + (0 if p % 3 else int(rng.integers(0, n)))] = complex(1.0, 0.0) if p % 2 else complex(0.0, -1.0)
    ctx = L.Context(sf)
    ctx.recheck_stats()
    sym, met, _ = ctx.demod_batch(torch.from_numpy(x).cuda(), mode=L.DEMOD_LEGACY)
    torch.cuda.synchronize()
    windows, rechecked = ctx.recheck_stats()
    assert windows == packets * total and rechecked <= windows
    assert rechecked >= windows // 2, "flat spectra must not be trusted by the fused flavour"
    sym = sym.cpu().numpy()
    for p in range(0, packets, 7):
        ret, so, sync, mo = ORC.lora_demodulate(sf, x[p])
        assert bits(sym[p]) == bits(so), (sf, p)
