"""The CUDA path against the reference's OWN compiled code, with no restatement in between.

Everywhere else the `-m gpu` tests compare the device with `ORC` (oracle/*.c, the C restatement), and the `not gpu`
tests pin `ORC` to the reference bit for bit.  Here the chain is closed directly on the GPU box: `REF` is
oracle/_ref/liblora_ref.so (src/phy/*.cpp of the reference, unmodified) and `POTHOS` is
oracle/_ref/liblora_pothos_ref.so (the four LoRa-SDR blocks, unmodified, on the header-only Pothos stand-in).  Both
libraries are built by oracle/Makefile in the container and travel to the GPU box as prebuilt files; /root/reference is
not read here.  Everything is bit for bit.
"""
import numpy as np
import pytest

from _oracle import POTHOS, REF, Params, chain_cfg

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(REF is None or POTHOS is None, reason="oracle/_ref was not built (needs /root/reference)")]


@pytest.fixture(scope="module")
def env():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import lora_phy_b200 as L
    return {"torch": torch, "L": L}


def bits(a):
    return np.ascontiguousarray(a).view(np.uint8).tobytes()


def unfused_dechirp(iq, down):
    x = iq.reshape(-1, len(down))
    re = (x.real * down.real[None]).astype(np.float32) - (x.imag * down.imag[None]).astype(np.float32)
    im = (x.real * down.imag[None]).astype(np.float32) + (x.imag * down.real[None]).astype(np.float32)
    return (re + 1j * im).astype(np.complex64).reshape(-1)


@pytest.mark.parametrize("sf", [7, 8, 9, 10, 11, 12])
def test_modulate_demodulate_and_decode_equal_the_compiled_reference(env, sf):
    """lora_modulate, the caller-dechirp -> lora_demodulate pipeline of tests/performance_test.cpp:103-116, demodulate(),
    and LoRaEncoder / LoRaDecoder::work, packet by packet."""
    torch, L = env["torch"], env["L"]
    n = 1 << sf
    ctx = L.Context(sf)
    cfg = L.chain_cfg(sf, 1)
    ocfg = chain_cfg(sf, 1)
    packets = 24 if sf <= 10 else 10
    gen = torch.Generator(device="cuda").manual_seed(4000 + sf)
    pay = torch.randint(0, 256, (packets, 16), dtype=torch.uint8, device="cuda", generator=gen)
    sym = ctx.chain_encode_batch(cfg, pay)
    pay_h, sym_h = pay.cpu().numpy(), sym.cpu().numpy().view(np.uint16)
    for p in range(0, packets, 5):                      # LoRaEncoder::work
        assert bits(sym_h[p]) == bits(POTHOS.chain_encode(ocfg, bytes(pay_h[p])))
    clean = ctx.modulate_batch(sym)
    clean_h = clean.cpu().numpy()
    for p in (0, packets - 1):                          # lora_modulate
        assert bits(clean_h[p]) == bits(REF.lora_modulate(sym_h[p], sf))
    one = ctx.modulate_batch(sym[:1].contiguous()).cpu().numpy()   # the one-packet kernel
    assert bits(one[0]) == bits(clean_h[0])
    iq = clean.clone()
    snr = torch.tensor([-12.0, -6.0, 0.0, 10.0], device="cuda")[torch.randint(0, 4, (packets,), device="cuda", generator=gen)]
    ctx.awgn_batch(iq, snr.contiguous(), seed=sf)
    h_iq = iq.cpu().numpy()
    down = REF.gen_chirp(n, 1, n, 0.0, True, 1.0, 0.0, 1.0)[0]
    for precision in (L.PRECISION_FAST, L.PRECISION_EXACT):
        ctx.set_precision(precision)
        rx, met, _ = ctx.demod_batch(iq, mode=L.DEMOD_LEGACY_DECHIRP)
        out, res = ctx.chain_decode_batch(cfg, rx.contiguous())
        rx_h = rx.cpu().numpy().view(np.uint16)
        met_h = met.cpu().numpy().view(L.PKT_METRICS).reshape(-1)
        res_h = res.cpu().numpy().view(L.CHAIN_RESULT).reshape(-1)
        out_h = out.cpu().numpy()
        for p in range(packets):
            _, so, sync, mo = REF.lora_demodulate(sf, unfused_dechirp(h_iq[p], down))
            assert bits(rx_h[p]) == bits(so), (sf, precision, p)
            assert int(met_h[p]["sync_word"]) == sync
            assert bits(met_h[p]["cfo"]) == bits(np.float32(mo.cfo)) and bits(met_h[p]["time_offset"]) == bits(np.float32(mo.time_offset))
            ro, bo, _ = POTHOS.chain_decode(ocfg, so)   # LoRaDecoder::work on the reference's own symbols
            if ro == -100 or int(res_h[p]["n_bytes"]) == -3:   # undefined in the reference itself (LoRaDecoder.cpp:381-384)
                continue
            assert int(res_h[p]["n_bytes"]) == ro, (sf, precision, p)
            assert bits(out_h[p, :max(ro, 0)]) == bits(bo[:max(ro, 0)])
        sym_w, met_w, _ = ctx.demod_batch(iq, mode=L.DEMOD_WORKSPACE)          # demodulate() of the new API
        sym_w, met_w = sym_w.cpu().numpy().view(np.uint16), met_w.cpu().numpy().view(L.PKT_METRICS).reshape(-1)
        for p in range(0, packets, 3):
            _, so, mo = REF.demodulate(Params(sf, 125000, 1, 0, 0, 0x12), h_iq[p])
            assert bits(sym_w[p]) == bits(so) and bits(met_w[p]["cfo"]) == bits(np.float32(mo.cfo))
    ctx.set_precision(L.PRECISION_FAST)
    assert (res_h["crc_ok"][(snr >= 0).cpu().numpy()] == 1).all()


@pytest.mark.parametrize("sf", [7, 9, 10])
def test_frame_generator_and_stream_synchroniser_equal_the_compiled_blocks(env, sf):
    """LoRaMod::work and LoRaDemod::work of LoRa-SDR, unmodified, against the device kernels on the same stream."""
    torch, L = env["torch"], env["L"]
    n = 1 << sf
    ctx = L.Context(sf)
    rng = np.random.default_rng(70 + sf)
    frames, nsym, pad = 3, 18, 3
    parts = []
    for k in range(frames):
        sym = rng.integers(0, n, nsym, dtype=np.uint16)
        d_sym = torch.from_numpy(sym.astype(np.int16)).cuda().reshape(1, -1)
        fr = ctx.frame_modulate_batch(d_sym, padding=pad, amplitude=0.3)[0].cpu().numpy()
        ref = POTHOS.frame_modulate(sf, sym, padding=pad, ampl=0.3)
        assert len(fr) == len(ref) and bits(fr) == bits(ref), (sf, k)
        parts.append(np.concatenate([np.zeros(int(rng.integers(0, 2 * n)), np.complex64), ref]))
    x = np.concatenate(parts + [np.zeros(3 * n, np.complex64)])
    x = (x * np.exp(2j * np.pi * 1.3 * np.arange(len(x)) / n)).astype(np.complex64)
    x = (x + (rng.standard_normal(len(x)) + 1j * rng.standard_normal(len(x))) * 0.02).astype(np.complex64)
    want, w_start, w_err, w_consumed = POTHOS.stream_demod(sf, x, mtu=nsym, thresh=-10.0, pkt_cap=16)
    sym_d, pk, res = ctx.stream_detect_batch(torch.from_numpy(x[None]).cuda(), thresh_db=-10.0, mtu=nsym, max_packets=16,
                                             max_symbols=1024)
    torch.cuda.synchronize()
    sym_d = sym_d.cpu().numpy()
    pk, res = L.stream_records(pk, res)
    assert int(res[0]["consumed"]) == w_consumed and int(res[0]["n_packets"]) == len(want) and len(want) >= 2
    off = 0
    for k, w in enumerate(want):
        assert int(pk[0, k]["start"]) == int(w_start[k]) and int(pk[0, k]["freq_error"]) == int(w_err[k])
        assert int(pk[0, k]["n_symbols"]) == len(w) and bits(sym_d[0, off:off + len(w)]) == bits(w), (sf, k)
        off += len(w)
