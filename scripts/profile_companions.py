#!/usr/bin/env python
"""Small fixed workloads for the companion kernels (ncu captures and CUDA-event timings):

    python scripts/profile_companions.py [--which modulate,frame,chain,stream7,awgn,scan] [--sf 12] [--reps 3]

  modulate   modulate_kernel            lora_modulate / genChirp (LoRaMod.cpp:8-43)        bytes: 8 per sample written
  frame      frame_modulate_kernel      LoRaMod::work (LoRa-SDR/LoRaMod.cpp:109-238)        bytes: 8 per sample written
  chain      chain_decode_kernel        LoRaDecoder::work (LoRa-SDR/LoRaDecoder.cpp:196-397) bytes: 2 per symbol + 24 per packet
  stream7    stream_detect_kernel<7>    LoRaDemod::work (LoRa-SDR/LoRaDemod.cpp:145-327)    bytes: 8 per sample scanned
  awgn       awgn_kernel                synthetic-input noise                                bytes: 16 per sample
Prints one line per kernel: best time, rate, fraction of the measured HBM peak (MEASURED_PEAKS.json).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-_b200"))


def best_ms(torch, fn, reps):
    fn()
    torch.cuda.synchronize()
    best = None
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--which", default="modulate,frame,chain,stream7,awgn")
    ap.add_argument("--sf", type=int, default=12)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--packets", type=int, default=0)
    ap.add_argument("--lib", default=None, help="variant library (same-box A/B)")
    args = ap.parse_args()
    import torch
    import lora_phy_b200 as L
    if args.lib:
        L.LIB_PATH = os.path.abspath(args.lib)
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    which = args.which.split(",")
    sf, n = args.sf, 1 << args.sf
    out = {}

    def report(name, ms, nbytes, units, unit):
        gbs = nbytes / ms / 1e6
        out[name] = {"ms": round(ms, 4), "GBps": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peak, 4),
                     "rate": round(units / ms / 1e3, 1), "rate_unit": unit}
        print(f"{name:10s} sf={sf}: {ms:.3f} ms  {gbs:.0f} GB/s = {gbs / peak:.3f} of {peak:.0f} GB/s   {units / ms / 1e3:.1f} {unit}")

    if "modulate" in which or "awgn" in which:
        packets = args.packets or min(65536, (24 << 30) // (35 * n * 8))
        ctx = L.Context(sf)
        g = torch.Generator(device="cuda").manual_seed(1)
        sym = torch.randint(0, n, (packets, 33), dtype=torch.int16, device="cuda", generator=g)
        iq = ctx.modulate_batch(sym)
        if "modulate" in which:
            ms = best_ms(torch, lambda: ctx.modulate_batch(sym, out=iq), args.reps)
            report("modulate", ms, iq.numel() * 8, iq.numel(), "Msamples/s")
        if "awgn" in which:
            snr = torch.zeros(packets, dtype=torch.float32, device="cuda")
            ms = best_ms(torch, lambda: ctx.awgn_batch(iq, snr, seed=1), args.reps)
            report("awgn", ms, iq.numel() * 16, iq.numel(), "Msamples/s")
        del iq
    if "frame" in which:
        packets = args.packets or min(65536, (24 << 30) // (50 * n * 8))
        ctx = L.Context(sf)
        g = torch.Generator(device="cuda").manual_seed(2)
        sym = torch.randint(0, n, (packets, 33), dtype=torch.int16, device="cuda", generator=g)
        fr = ctx.frame_modulate_batch(sym, amplitude=0.3, padding=2)
        ms = best_ms(torch, lambda: ctx.frame_modulate_batch(sym, amplitude=0.3, padding=2, out=fr), args.reps)
        report("frame", ms, fr.numel() * 8, fr.numel(), "Msamples/s")
        del fr
    if "chain" in which:
        packets = args.packets or 1 << 20
        ctx = L.Context(sf)
        cfg = L.chain_cfg(sf, 1)
        g = torch.Generator(device="cuda").manual_seed(3)
        pay = torch.randint(0, 256, (packets, 16), dtype=torch.uint8, device="cuda", generator=g)
        sym = ctx.chain_encode_batch(cfg, pay)
        o = torch.zeros((packets, L.PAYLOAD_STRIDE), dtype=torch.uint8, device="cuda")
        r = torch.zeros((packets, 16), dtype=torch.uint8, device="cuda")
        ms = best_ms(torch, lambda: ctx.chain_decode_batch(cfg, sym, payload=o, results=r), args.reps)
        report("chain", ms, packets * (2 * sym.shape[1] + 24), packets, "Mpackets/s")
    if "stream7" in which:
        ssf = 7
        sn = 1 << ssf
        streams = args.packets or 16384
        ctx = L.Context(ssf)
        g = torch.Generator(device="cuda").manual_seed(4)
        frames, nsym, pad = 4, 40, 4
        per = ctx.frame_samples(nsym, pad)
        total = 2 * sn + frames * (per + sn)
        iq = torch.zeros((streams, total), dtype=torch.complex64, device="cuda")
        for f in range(frames):
            sym = torch.randint(0, sn, (streams, nsym), dtype=torch.int16, device="cuda", generator=g)
            off = 2 * sn + f * (per + sn) + (37 * (f + 1)) % sn
            ctx.frame_modulate_batch(sym, amplitude=1.0, padding=pad, out=iq[:, off:off + per])
        ctx.awgn_batch(iq, torch.full((streams,), 10.0, dtype=torch.float32, device="cuda"), seed=7)
        ms = best_ms(torch, lambda: ctx.stream_detect_batch(iq, mtu=nsym, max_packets=frames + 2, max_symbols=(frames + 2) * nsym), args.reps)
        report("stream7", ms, iq.numel() * 8, iq.numel(), "Msamples/s")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
