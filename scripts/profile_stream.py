#!/usr/bin/env python
"""Throughput of the continuous-stream front end (BASELINE configs[4]-style input).

    python scripts/profile_stream.py --sf 9 --streams 4096 --frames 4 --symbols 40

Every stream is a run of generated frames (random payload symbols, random gap before each) with AWGN; the kernel runs
the frame synchroniser over all of them.  Prints CUDA-event time of the detector launch and the samples/s it implies.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=int, default=9)
    ap.add_argument("--streams", type=int, default=4096)
    ap.add_argument("--frames", type=int, default=4)
    ap.add_argument("--symbols", type=int, default=40)
    ap.add_argument("--snr", type=float, default=10.0)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    import torch
    import lora_phy_b200 as L

    sf, n = args.sf, 1 << args.sf
    ctx = L.Context(sf)
    gen = torch.Generator(device="cuda").manual_seed(sf)
    pad = 4
    per_frame = ctx.frame_samples(args.symbols, pad)
    total = 2 * n + args.frames * (per_frame + n)
    iq = torch.zeros((args.streams, total), dtype=torch.complex64, device="cuda")
    tx = []
    for f in range(args.frames):
        sym = torch.randint(0, n, (args.streams, args.symbols), dtype=torch.int16, device="cuda", generator=gen)
        tx.append(sym)
        # every stream gets the same frame slot but its own content; the slot start is shifted per frame
        off = 2 * n + f * (per_frame + n) + (37 * (f + 1)) % n
        ctx.frame_modulate_batch(sym, amplitude=1.0, padding=pad, out=iq[:, off:off + per_frame])
    snr = torch.full((args.streams,), args.snr, dtype=torch.float32, device="cuda")
    ctx.awgn_batch(iq, snr, seed=sf)
    torch.cuda.synchronize()
    ctx.set_kernel_timing(True)
    for _ in range(args.reps):
        sym, pk, res = ctx.stream_detect_batch(iq, mtu=args.symbols, max_packets=args.frames + 2,
                                               max_symbols=(args.frames + 2) * args.symbols)
        torch.cuda.synchronize()
        ms, k = ctx.kernel_time()
        print(f"sf={sf} streams={args.streams} samples/stream={total}: detector {ms:.3f} ms "
              f"-> {iq.numel() / ms / 1e6:.2f} Gsamples/s ({iq.numel() * 8 / ms / 1e6:.1f} GB/s)")
    pk, res = L.stream_records(pk, res)
    found = int(res["n_packets"].sum())
    sym = sym.cpu().numpy()
    good = 0
    for f in range(args.frames):
        t = tx[f].cpu().numpy()
        good += int(((sym[:, f * args.symbols:(f + 1) * args.symbols] == t).all(axis=1) & (res["n_packets"] == args.frames)).sum())
    print(f"packets found {found} of {args.streams * args.frames}; frames recovered symbol-exact {good}")


if __name__ == "__main__":
    main()
