#!/usr/bin/env python
"""Small, fixed workload for ncu: bench.py's configs[1] packets at one SF, demodulated a few times.

    python scripts/profile_demod.py --sf 12 --packets 2048 [--reps 3] [--precision fast|exact] [--mode legacy_dechirp]

Prints per-launch CUDA-event times of the demod kernels so a plain run can be compared with the profile.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=int, default=12)
    ap.add_argument("--packets", type=int, default=2048)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--precision", default="fast", choices=["fast", "exact"])
    ap.add_argument("--no-overlap", action="store_true")
    ap.add_argument("--overlap", type=int, default=None, help="lora_b200_set_overlap value (0, 1 or 2)")
    ap.add_argument("--mode", default="legacy_dechirp", choices=["legacy_dechirp", "workspace", "legacy"])
    ap.add_argument("--lib", default=None, help="variant library built by scripts/ab_build.sh (same-box A/B timing)")
    args = ap.parse_args()
    import torch
    import lora_phy_b200 as L
    if args.lib:
        L.LIB_PATH = os.path.abspath(args.lib)

    sf = args.sf
    ctx = L.Context(sf)
    ctx.set_precision(L.PRECISION_FAST if args.precision == "fast" else L.PRECISION_EXACT)
    ctx.set_overlap(args.overlap if args.overlap is not None else (0 if args.no_overlap else 1))
    cfg = L.chain_cfg(sf, 1)
    gen = torch.Generator(device="cuda").manual_seed(sf)
    pay = torch.randint(0, 256, (args.packets, 16), dtype=torch.uint8, device="cuda", generator=gen)
    sym = ctx.chain_encode_batch(cfg, pay)
    iq = ctx.modulate_batch(sym)
    snrs = torch.tensor([-20.0, -15.0, -10.0, -5.0, 0.0, 5.0, 10.0], device="cuda")
    snr = snrs[torch.randint(0, 7, (args.packets,), device="cuda", generator=gen)].contiguous()
    ctx.awgn_batch(iq, snr, seed=sf)
    mode = {"legacy_dechirp": L.DEMOD_LEGACY_DECHIRP, "workspace": L.DEMOD_WORKSPACE, "legacy": L.DEMOD_LEGACY}[args.mode]
    rx = torch.zeros((args.packets, sym.shape[1]), dtype=torch.int16, device="cuda")
    met = torch.zeros((args.packets, 16), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    ctx.set_kernel_timing(True)
    for _ in range(args.reps):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        ctx.demod_batch(iq, mode=mode, symbols=rx, metrics=met)
        ev1.record()
        torch.cuda.synchronize()
        ms, k = ctx.kernel_time()
        samples = iq.numel()
        print(f"sf={sf} packets={args.packets} whole demod_batch {ev0.elapsed_time(ev1):.3f} ms, demod kernels "
              f"{ms:.3f} ms ({k} timed regions): {samples / ms / 1e6:.1f} Gsamples/s, {samples * 8 / ms / 1e6:.1f} GB/s")
    w, r = ctx.recheck_stats()
    print(f"windows {w} rechecked {r} ({100.0 * r / max(w, 1):.3f} %)  symbols_eq_tx {(rx == sym).float().mean().item():.4f}")


if __name__ == "__main__":
    main()
