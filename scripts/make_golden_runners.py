#!/usr/bin/env python
"""tests/golden/runners_loopback.npz: the UNMODIFIED reference tx_runner / rx_runner (built in /tmp from
/root/reference, its own sources and flags) on a few payloads -- the drop-in target for bin/tx_runner and
bin/rx_runner of this repo.  Run in the build container only:  python scripts/make_golden_runners.py"""
import os
import subprocess

import numpy as np

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = [f"{REF}/src/phy/{f}" for f in ("phy.cpp", "LoRaDemod.cpp", "LoRaMod.cpp", "LoRaEncoder.cpp", "LoRaDecoder.cpp")]


def build(name):
    exe = f"/tmp/ref_{name}"
    subprocess.run(["g++", "-std=c++11", "-O2", "-w", f"-I{REF}/include", "-o", exe, f"{REF}/runners/{name}.cpp"] + LIB,
                   check=True)
    return exe


def main():
    tx, rx = build("tx_runner"), build("rx_runner")
    out = {}
    cases = [("deadbeef0102030405", 7, 125000), ("00ff10ef", 9, 250000), ("a5" * 12, 12, 500000)]
    for k, (payload, sf, bw) in enumerate(cases):
        path = f"/tmp/golden_tx_{k}.iq"
        subprocess.run([tx, f"--payload={payload}", f"--sf={sf}", f"--bw={bw}", f"--out={path}"], check=True)
        iq = np.fromfile(path, np.complex64)
        stdout = subprocess.run([rx, f"--in={path}", f"--sf={sf}", f"--bw={bw}", "--report-offsets"], check=True,
                                capture_output=True).stdout
        out[f"case{k}_args"] = np.array([payload, str(sf), str(bw)])
        out[f"case{k}_iq"] = iq
        out[f"case{k}_rx_stdout"] = np.frombuffer(stdout, np.uint8)
        print(payload, sf, bw, len(iq), stdout.decode().strip().replace("\n", " | "))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "runners_loopback.npz"), **out)


if __name__ == "__main__":
    main()
