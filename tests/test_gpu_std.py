"""GPU: the deployed LoRa coding chain (csrc/std_kernels.cu) through the C ABI, against
oracle/std_chain.c on the same inputs, and the reference's golden capture end to end on the device:
capture -> frame synchroniser (decimating by two on the fly) -> deployed chain -> sixteen CRC-valid payloads.

Bar: bit-exact (integer work).  The oracle itself is pinned by that capture, see tests/test_oracle_std.py."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def env():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import lora_phy_b200 as L
    from _oracle import ORC, std_cfg
    return {"torch": torch, "L": L, "ORC": ORC, "orc_cfg": std_cfg}


def bits(a):
    return np.ascontiguousarray(a).view(np.uint8).tobytes()


def test_golden_capture_end_to_end_on_the_device(env):
    """The reference's rx_runner prints 1043 bytes of noise for this file (reproduced bit for bit by
    test_config1_golden_capture_rx_runner_path).  The stream front end plus the deployed chain recover what was sent."""
    torch, L = env["torch"], env["L"]
    iq = np.load(os.path.join(HERE, "golden", "capture_sf7.npz"))["iq"]
    ctx = L.Context(7, osr=2)                                         # the capture is 2x oversampled: samples 0, 2, 4, ...
    d = torch.from_numpy(np.ascontiguousarray(iq)[None]).cuda()
    sym, pk, res = ctx.stream_detect_batch(d, mtu=40, max_packets=32, max_symbols=2048)
    payload, r = ctx.std_stream_decode_batch(L.std_cfg(7), sym, pk, res)
    pkr, rr = L.stream_records(pk, res)
    assert int(rr[0]["n_packets"]) == 16
    assert [int(p["start"]) for p in pkr[0, :16]] == [1568 + 8352 * k for k in range(16)]
    r = r.cpu().numpy().reshape(-1, 16).view(L.STD_RESULT).reshape(-1)
    payload = payload.cpu().numpy()[0]
    for k in range(16):
        text = f"hello world: {k}".encode()
        assert int(r[k]["n_bytes"]) == len(text) and bytes(payload[k, :len(text)]) == text
        assert r[k]["hdr_ok"] == 1 and r[k]["hdr_cr"] == 1 and r[k]["has_crc"] == 1 and r[k]["crc_ok"] == 1
        assert r[k]["corrected"] == 0 and r[k]["uncorrectable"] == 0 and int(r[k]["symbols_used"]) == 33
    assert (r[16:]["n_bytes"] == -1).all()                             # empty packet slots
    # and the oracle agrees symbol for symbol on the decimated capture
    frames, _, _, _ = env["ORC"].stream_demod(7, np.ascontiguousarray(iq[0::2]), mtu=40, pkt_cap=32)
    s = sym.cpu().numpy()[0]
    for k, f in enumerate(frames):
        a = int(pkr[0, k]["first_symbol"])
        assert bits(s[a:a + len(f)]) == bits(f)


@pytest.mark.parametrize("sf", [7, 9, 12])
def test_encode_decode_vs_oracle(env, sf):
    torch, L, ORC, orc_cfg = env["torch"], env["L"], env["ORC"], env["orc_cfg"]
    ctx = L.Context(sf)
    rng = np.random.default_rng(sf)
    for cr in (1, 2, 3, 4):
        for ldro in (False, True):
            for implicit in (False, True):
                for crc in (False, True):
                    for n in (1, 2, 16, 33, 255):
                        pay = rng.integers(0, 256, (6, n), dtype=np.uint8)
                        g = L.std_cfg(sf, cr, ldro, implicit, crc, n)
                        o = orc_cfg(sf, cr, ldro, implicit, crc, n)
                        sym = ctx.std_encode_batch(g, torch.from_numpy(pay).cuda())
                        h = sym.cpu().numpy().view(np.uint16).copy()
                        for p in range(6):
                            assert bits(h[p]) == bits(ORC.std_encode(o, pay[p])), (sf, cr, ldro, implicit, crc, n)
                        # corrupt: packet 1 one data symbol by one Gray bit, packet 2 a header symbol by +-1 bin,
                        # packet 3 two symbols badly, packet 4 cut short (symbols zeroed at the end do not matter: the
                        # kernel sees the same zeros as the oracle)
                        nsym = h.shape[1]
                        if nsym > 9:
                            h[1, 9] = (((int(h[1, 9]) - 1) ^ 2) + 1) % (1 << sf)
                            h[3, 8] = (int(h[3, 8]) + 37) % (1 << sf)
                        h[2, 2] = (int(h[2, 2]) + 1) % (1 << sf)
                        h[3, nsym - 1] = (int(h[3, nsym - 1]) + 91) % (1 << sf)
                        h[4, nsym - 3:] = 0
                        out, res = ctx.std_decode_batch(g, torch.from_numpy(h.view(np.int16)).cuda())
                        res = res.cpu().numpy().view(L.STD_RESULT).reshape(-1)
                        out = out.cpu().numpy()
                        for p in range(6):
                            ret, data, info = ORC.std_decode(o, h[p])
                            assert int(res[p]["n_bytes"]) == ret, (sf, cr, ldro, implicit, crc, n, p)
                            for f in ("hdr_ok", "hdr_length", "hdr_cr", "has_crc", "crc_ok", "corrected", "uncorrectable"):
                                assert int(res[p][f]) == (getattr(info, f) & 0xFF), (f, sf, cr, ldro, implicit, crc, n, p)
                            if ret >= 0:
                                assert bits(out[p, :ret]) == bits(data) and int(res[p]["symbols_used"]) == info.symbols_used
                        assert bits(out[0, :n]) == bits(pay[0]) and bits(out[5, :n]) == bits(pay[5])


def test_full_chain_loopback_through_the_air(env):
    """Deployed-chain encoder -> frame generator -> AWGN -> synchroniser -> deployed-chain decoder, on the device.
    The frame generator is the Pothos one (sync word as two chirps, 2.25 down-chirps), which is what the synchroniser
    locks onto; the symbols in between are the deployed chain's."""
    torch, L = env["torch"], env["L"]
    for sf, cr, ldro in ((7, 1, False), (9, 4, False), (10, 2, True)):
        ctx = L.Context(sf)
        g = L.std_cfg(sf, cr, ldro)
        gen = torch.Generator(device="cuda").manual_seed(sf)
        pay = torch.randint(0, 256, (24, 19), dtype=torch.uint8, device="cuda", generator=gen)
        sym = ctx.std_encode_batch(g, pay)
        iq = ctx.frame_modulate_batch(sym, padding=2)
        ctx.awgn_batch(iq, torch.full((24,), 12.0, device="cuda"), seed=sf)
        s, pk, res = ctx.stream_detect_batch(iq, mtu=sym.shape[1], max_packets=2, max_symbols=sym.shape[1] + 8)
        out, r = ctx.std_stream_decode_batch(g, s, pk, res)
        r = r.cpu().numpy().reshape(24, 2, 16).view(L.STD_RESULT).reshape(24, 2)
        out = out.cpu().numpy()
        ok = 0
        for p in range(24):
            if r[p, 0]["n_bytes"] == 19 and r[p, 0]["crc_ok"] == 1:
                assert bits(out[p, 0, :19]) == bits(pay[p].cpu().numpy())
                ok += 1
        assert ok >= 20, (sf, ok)


def test_host_call_and_cli_on_the_golden_capture(env, tmp_path):
    """lora_b200_std_receive_stream_host with mtu = 0 (frame length taken from the headers) and the CLI over it."""
    import subprocess
    L = env["L"]
    iq = np.load(os.path.join(HERE, "golden", "capture_sf7.npz"))["iq"]
    frames, res, pk = L.Context(7, osr=2).std_receive_stream_host(L.std_cfg(7), iq)
    assert frames == [f"hello world: {k}".encode() for k in range(16)]
    assert (res["crc_ok"] == 1).all() and (res["symbols_used"] == 33).all()
    assert [int(s) for s in pk["start"]] == [1568 + 8352 * k for k in range(16)]
    exe = os.path.join(os.path.dirname(HERE), "lora-sdr-lightweight-standalone-library-_b200", "bin", "lora_rx_stream")
    path = tmp_path / "capture.bin"
    np.ascontiguousarray(iq).tofile(path)
    out = subprocess.run([exe, f"--in={path}", "--sf=7", "--osr=2"], capture_output=True, text=True, check=True).stdout
    lines = out.strip().split("\n")
    assert lines[0] == "Frames: 16"
    for k in range(16):
        text = f"hello world: {k}"
        assert lines[1 + k] == (f"Frame {k} @{1568 + 8352 * k}: len {len(text)} CR 4/5 CRC ok Payload: "
                                f"{text.encode().hex()}  \"{text}\"")


def test_frame_header_mode_equals_oracle(env):
    """lora_b200_stream_cfg.frame_header: the synchroniser reads the frame length from the explicit header after eight
    data symbols.  Same packets, symbols and consumed-sample counts as the oracle's state machine, on a train of unequal
    frames one padding symbol apart (with noise) and on the golden capture."""
    torch, L, ORC, orc_cfg = env["torch"], env["L"], env["ORC"], env["orc_cfg"]
    rng = np.random.default_rng(11)
    for sf, cr, ldro in ((8, 2, False), (9, 4, True), (7, 1, False)):
        pays = [rng.integers(0, 256, n, dtype=np.uint8) for n in (3, 60, 1, 17, 200)]
        syms = [ORC.std_encode(orc_cfg(sf, cr, ldro), p) for p in pays]
        rows = []
        for r in range(3):
            order = rng.permutation(len(syms))
            x = np.concatenate([ORC.frame_modulate(sf, syms[k], padding=1, ampl=0.5) for k in order] +
                               [np.zeros(3 << sf, np.complex64)])
            x = x + ((rng.standard_normal(len(x)) + 1j * rng.standard_normal(len(x))) * 0.02).astype(np.complex64)
            rows.append((order, x.astype(np.complex64)))
        n = max(len(x) for _, x in rows)
        batch = np.zeros((3, n), np.complex64)
        for r, (_, x) in enumerate(rows):
            batch[r, :len(x)] = x
        ctx = L.Context(sf)
        mode = 2 if ldro else 1
        sym, pk, res = ctx.stream_detect_batch(torch.from_numpy(batch).cuda(), mtu=1000, max_packets=8, max_symbols=2048,
                                               frame_header=mode)
        out, sr = ctx.std_stream_decode_batch(L.std_cfg(sf, cr, ldro), sym, pk, res)
        pkr, rr = L.stream_records(pk, res)
        sym, out = sym.cpu().numpy(), out.cpu().numpy()
        sr = sr.cpu().numpy().reshape(3, 8, 16).view(L.STD_RESULT).reshape(3, 8)
        for r, (order, _) in enumerate(rows):
            want, start, err, consumed = ORC.stream_demod(sf, batch[r], mtu=1000, pkt_cap=8, frame_header=mode)
            assert int(rr[r]["n_packets"]) == len(want) == len(order) and int(rr[r]["consumed"]) == consumed
            for k, w in enumerate(want):
                a = int(pkr[r, k]["first_symbol"])
                assert int(pkr[r, k]["n_symbols"]) == len(w) == len(syms[order[k]]) and int(pkr[r, k]["start"]) == int(start[k])
                assert bits(sym[r, a:a + len(w)]) == bits(w)
                p = pays[order[k]]
                assert int(sr[r, k]["n_bytes"]) == len(p) and sr[r, k]["crc_ok"] == 1 and bits(out[r, k, :len(p)]) == bits(p)
    iq = np.load(os.path.join(HERE, "golden", "capture_sf7.npz"))["iq"]
    sym, pk, res = L.Context(7, osr=2).stream_detect_batch(torch.from_numpy(np.ascontiguousarray(iq)[None]).cuda(), mtu=600,
                                                           max_packets=32, max_symbols=2048, frame_header=1)
    pkr, rr = L.stream_records(pk, res)
    assert int(rr[0]["n_packets"]) == 16 and (pkr[0, :16]["n_symbols"] == 33).all()
