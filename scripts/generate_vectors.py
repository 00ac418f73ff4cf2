#!/usr/bin/env python3
"""Stage-dump vectors from the device path, with a sha256 manifest.

Counterpart of the reference's scripts/generate_lora_phy_vectors.py:90-143 for this build: runs
`lora_phy_vector_dump` -- the reference's runners/lora_phy_vector_dump.cpp compiled UNMODIFIED against this repo's
include/lora_phy headers and liblora_phy.so, so encode / modulate / demodulate run on the GPU -- into
<root>/<out>/, optionally derives `iq_samples_offset.csv` with an injected carrier offset (per-symbol phase ramp of
--cfo-bins) and sample shift (--time-offset, zero fill), base64-encodes every dump file next to a sha256 of the
encoded file, and writes manifest.json {sf, seed, bytes, osr, bw, files:[{name, sha256}]} -- the same record the
reference script writes, so a manifest produced from the reference binary compares equal field by field.

    python scripts/generate_vectors.py --sf 7 --seed 0 --out sf7_seed0 [--binary PATH] [--root DIR]
"""
import argparse
import base64
import hashlib
import json
import math
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_BIN = os.path.join(os.path.dirname(HERE), "lora-sdr-lightweight-standalone-library-_b200", "bin", "lora_phy_vector_dump")


def inject_offsets(iq_csv, out_csv, n_per_symbol, cfo_bins, time_offset):
    samples = []
    with open(iq_csv) as f:
        for line in f:
            if line.strip():
                re, im = line.split(",")
                samples.append(complex(float(re), float(im)))
    if cfo_bins != 0.0:
        for k, s in enumerate(samples):
            ph = 2.0 * math.pi * cfo_bins * (k % n_per_symbol) / n_per_symbol
            samples[k] = s * complex(math.cos(ph), math.sin(ph))
    shift = int(round(time_offset))
    if shift > 0:
        samples = samples[shift:] + [0j] * shift
    elif shift < 0:
        samples = [0j] * (-shift) + samples[:shift]
    with open(out_csv, "w") as f:
        for s in samples:
            f.write(f"{s.real},{s.imag}\n")


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("--sf", type=int, required=True)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--bytes", type=int, default=16)
    ap.add_argument("--osr", type=int, default=1)
    ap.add_argument("--bw", type=int, default=125000)
    ap.add_argument("--out", required=True, help="sub-directory under --root")
    ap.add_argument("--root", default="vectors/lora_phy")
    ap.add_argument("--binary", default=os.environ.get("LORAPHY_VECTOR_BIN", DEFAULT_BIN))
    ap.add_argument("--cfo-bins", type=float, default=0.0)
    ap.add_argument("--time-offset", type=float, default=0.0)
    ap.add_argument("--window", default="none", choices=["none", "hann"])
    a = ap.parse_args(argv)
    if not os.path.isfile(a.binary):
        ap.error(f"vector dump binary not found: {a.binary}")
    out = os.path.join(a.root, a.out)
    os.makedirs(out, exist_ok=True)
    cmd = [os.path.abspath(a.binary), f"--sf={a.sf}", f"--seed={a.seed}", f"--bytes={a.bytes}", f"--out={out}",
           f"--osr={a.osr}", f"--bw={a.bw}"]
    if a.window != "none":
        cmd.append(f"--window={a.window}")
    print("[run]", " ".join(cmd), file=sys.stderr)
    subprocess.run(cmd, check=True)
    iq = os.path.join(out, "iq_samples.csv")
    if (a.cfo_bins != 0.0 or a.time_offset != 0.0) and os.path.isfile(iq):
        inject_offsets(iq, os.path.join(out, "iq_samples_offset.csv"), (1 << a.sf) * a.osr, a.cfo_bins, a.time_offset)
    files = []
    for name in sorted(os.listdir(out)):
        path = os.path.join(out, name)
        if name == "manifest.json" or name.endswith(".b64") or not os.path.isfile(path):
            continue
        with open(path, "rb") as src, open(path + ".b64", "wb") as dst:
            base64.encode(src, dst)
        os.unlink(path)
        with open(path + ".b64", "rb") as f:
            files.append({"name": name + ".b64", "sha256": hashlib.sha256(f.read()).hexdigest()})
    manifest = {"sf": a.sf, "seed": a.seed, "bytes": a.bytes, "osr": a.osr, "bw": a.bw, "files": files}
    with open(os.path.join(out, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=2)
    return manifest


if __name__ == "__main__":
    main()
