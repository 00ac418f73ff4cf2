#!/usr/bin/env python
"""Generate tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref, built from /root/reference).

Run in the build container only:  python scripts/make_golden.py
Each fixture stores the input and what the reference produced for it, so the GPU box (which has no
/root/reference) can check the CUDA path against the reference's own answers.  The two IQ files are
reference *data* fixtures (vectors_binary/..., test_output.iq), stored compressed; no source is copied.
"""
import hashlib
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _oracle import REF, Params, build_oracle  # noqa: E402

REF_ROOT = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
CAPTURE = f"{REF_ROOT}/vectors_binary/bw_125k_sf_7_cr_1_ldro_false_crc_true_implheader_false.unknown"


def awgn(x, snr_db, seed):
    sigma = 10.0 ** (-snr_db / 20.0)
    r = np.random.default_rng(seed)
    n = (r.standard_normal(len(x)) + 1j * r.standard_normal(len(x))) * (sigma / np.sqrt(2.0))
    return (x + n).astype(np.complex64)


def rx_runner_stdout():
    """Build the reference's own rx_runner (its sources, its flags) in /tmp and run it on the capture."""
    exe = "/tmp/ref_rx_runner"
    src = [f"{REF_ROOT}/runners/rx_runner.cpp"] + [f"{REF_ROOT}/src/phy/{f}" for f in
           ("phy.cpp", "LoRaDemod.cpp", "LoRaMod.cpp", "LoRaEncoder.cpp", "LoRaDecoder.cpp")]
    subprocess.run(["g++", "-std=c++11", "-O2", "-w", f"-I{REF_ROOT}/include", "-o", exe] + src, check=True)
    out = subprocess.run([exe, f"--in={CAPTURE}", "--sf=7", "--cr=1", "--bw=125000", "--report-offsets"],
                         check=True, capture_output=True).stdout
    return out


def main():
    if REF is None:
        build_oracle()
        raise SystemExit("oracle/_ref was missing; built it -- run again")
    os.makedirs(OUT, exist_ok=True)

    # config 1: the committed capture through demodulate() + decode()
    iq = np.fromfile(CAPTURE, np.complex64)
    p = Params(7, 125000, 1, 0, 0, 0x12)
    ret, sym, m = REF.demodulate(p, iq)
    payload, crc_ok = REF.decode(sym)
    stdout = rx_runner_stdout()
    sha = hashlib.sha256(stdout).hexdigest()
    assert sha == "91e55ba04aaff03239abd76229b6a6b13ec139f2e351439f92f87f726ac2cddb", sha
    np.savez_compressed(os.path.join(OUT, "capture_sf7.npz"), iq=iq, symbols=sym, payload=payload,
                        crc_ok=crc_ok, cfo=np.float32(m.cfo), time_offset=np.float32(m.time_offset),
                        sync_word=np.uint8(m.sync_word), rx_runner_stdout=np.frombuffer(stdout, np.uint8))

    # tests/gr_lora_sdr_interop.cpp: osr=2 capture through lora_demodulate()
    iq2 = np.fromfile(f"{REF_ROOT}/test_output.iq", np.complex64)
    ret, sym2, sync2, m2 = REF.lora_demodulate(7, iq2, osr=2)
    np.savez_compressed(os.path.join(OUT, "interop_sf7_osr2.npz"), iq=iq2, symbols=sym2, sync=np.uint8(sync2),
                        payload=REF.lora_decode(sym2), cfo=np.float32(m2.cfo),
                        time_offset=np.float32(m2.time_offset))

    # synthetic AWGN packets, both demod entry points, SF7..12
    pk = {}
    for sf in range(7, 13):
        r = np.random.default_rng(1000 + sf)
        nsym = 10 if sf <= 10 else 5
        n = 1 << sf
        for k, snr in enumerate((20.0, 0.0, -12.0)):
            sym_tx = r.integers(0, n, nsym, dtype=np.uint16)
            clean = REF.lora_modulate(sym_tx, sf)
            y = awgn(clean, snr, sf * 100 + k)
            tag = f"sf{sf}_{k}"
            pk[f"{tag}_tx"] = sym_tx
            pk[f"{tag}_iq"] = y
            _, s_new, m_new = REF.demodulate(Params(sf, 125000, 1, 0, 0, 0x12), y)
            pk[f"{tag}_demodulate"] = s_new
            pk[f"{tag}_demodulate_m"] = np.array([m_new.cfo, m_new.time_offset], np.float32)
            pk[f"{tag}_demodulate_sync"] = np.uint8(m_new.sync_word)
            # caller-side dechirp (tests/e2e_chain_test.cpp:79-93), un-fused complex product
            down, _ = REF.gen_chirp(n, 1, n, 0.0, True, 1.0, 0.0, 1.0)
            x = y.reshape(-1, n)
            re = (x.real * down.real[None]).astype(np.float32) - (x.imag * down.imag[None]).astype(np.float32)
            im = (x.real * down.imag[None]).astype(np.float32) + (x.imag * down.real[None]).astype(np.float32)
            de = (re + 1j * im).astype(np.complex64).reshape(-1)
            _, s_leg, sync_leg, m_leg = REF.lora_demodulate(sf, de)
            pk[f"{tag}_legacy"] = s_leg
            pk[f"{tag}_legacy_m"] = np.array([m_leg.cfo, m_leg.time_offset], np.float32)
            pk[f"{tag}_legacy_sync"] = np.uint8(sync_leg)
        pk[f"sf{sf}_modulated"] = REF.lora_modulate(np.arange(3, dtype=np.uint16) * 37 % n, sf)
        pk[f"sf{sf}_twiddles"] = REF.twiddles(n)
        pk[f"sf{sf}_downchirp"] = REF.gen_chirp(n, 1, n, 0.0, True, 1.0, 0.0, 1.0)[0]
    np.savez_compressed(os.path.join(OUT, "awgn_packets.npz"), **pk)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
