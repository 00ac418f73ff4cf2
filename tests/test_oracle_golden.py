"""The oracle against the committed golden fixtures (tests/golden/*.npz, produced from the unmodified
reference by scripts/make_golden.py).  Needs neither /root/reference nor oracle/_ref, so it also pins the
oracle on the GPU box.  CPU only.
"""
import hashlib
import os

import numpy as np

from _oracle import ORC, Params

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def bits(a):
    return np.ascontiguousarray(a).view(np.uint8).tobytes()


def unfused_dechirp(iq, down):
    x = iq.reshape(-1, len(down))
    re = (x.real * down.real[None]).astype(np.float32) - (x.imag * down.imag[None]).astype(np.float32)
    im = (x.real * down.imag[None]).astype(np.float32) + (x.imag * down.real[None]).astype(np.float32)
    return (re + 1j * im).astype(np.complex64).reshape(-1)


def test_capture_rx_runner_path():
    g = np.load(os.path.join(G, "capture_sf7.npz"))
    ret, sym, m = ORC.demodulate(Params(7, 125000, 1, 0, 0, 0x12), g["iq"])
    assert ret == 2086 and bits(sym) == bits(g["symbols"])
    assert bits(np.float32(m.cfo)) == bits(g["cfo"]) and bits(np.float32(m.time_offset)) == bits(g["time_offset"])
    assert m.sync_word == int(g["sync_word"])
    payload, ok = ORC.decode(sym)
    assert bits(payload) == bits(g["payload"]) and ok == bool(g["crc_ok"])
    # the exact text runners/rx_runner.cpp:124-135 prints, rebuilt from the oracle's values
    text = "Payload: " + bytes(payload).hex() + "\n" + f"CRC OK: {'yes' if ok else 'no'}\n"
    text += f"CFO: {float(m.cfo):g}\nTime offset: {float(m.time_offset):g}\n"
    assert text.encode() == bytes(g["rx_runner_stdout"])
    assert hashlib.sha256(text.encode()).hexdigest() == \
        "91e55ba04aaff03239abd76229b6a6b13ec139f2e351439f92f87f726ac2cddb"


def test_interop_osr2():
    g = np.load(os.path.join(G, "interop_sf7_osr2.npz"))
    ret, sym, sync, m = ORC.lora_demodulate(7, g["iq"], osr=2)
    assert bits(sym) == bits(g["symbols"]) and sync == int(g["sync"]) == 0x29
    assert bytes(ORC.lora_decode(sym)) == bytes(g["payload"]) == bytes([0xBE, 0xE7, 0x82, 0x75, 0xE0])
    assert bits(np.float32(m.cfo)) == bits(g["cfo"])


def test_awgn_packets_and_tables():
    g = np.load(os.path.join(G, "awgn_packets.npz"))
    for sf in range(7, 13):
        n = 1 << sf
        assert bits(ORC.twiddles(n)) == bits(g[f"sf{sf}_twiddles"])
        down = ORC.gen_chirp(n, 1, n, 0.0, True, 1.0, 0.0, 1.0)[0]
        assert bits(down) == bits(g[f"sf{sf}_downchirp"])
        assert bits(ORC.lora_modulate(np.arange(3, dtype=np.uint16) * 37 % n, sf)) == bits(g[f"sf{sf}_modulated"])
        for k in range(3):
            t = f"sf{sf}_{k}"
            ret, sym, m = ORC.demodulate(Params(sf, 125000, 1, 0, 0, 0x12), g[f"{t}_iq"])
            assert bits(sym) == bits(g[f"{t}_demodulate"])
            assert bits(np.array([m.cfo, m.time_offset], np.float32)) == bits(g[f"{t}_demodulate_m"])
            assert m.sync_word == int(g[f"{t}_demodulate_sync"])
            ret, sym, sync, m = ORC.lora_demodulate(sf, unfused_dechirp(g[f"{t}_iq"], down))
            assert bits(sym) == bits(g[f"{t}_legacy"]) and sync == int(g[f"{t}_legacy_sync"])
            assert bits(np.array([m.cfo, m.time_offset], np.float32)) == bits(g[f"{t}_legacy_m"])
            if k == 0:  # 20 dB: the caller-dechirp path round-trips (tests/e2e_chain_test.cpp)
                assert list(sym) == list(g[f"{t}_tx"]) and sync == 0x12
