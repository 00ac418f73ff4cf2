#!/usr/bin/env python
"""Latency of the single-packet calls the reference's per-packet API makes (tests/performance_test.cpp:92-116):
lora_b200_modulate_batch and lora_b200_demod_batch on ONE packet, CUDA events, best of 20."""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-_b200"))
ap = argparse.ArgumentParser(); ap.add_argument("--lib", default=None); args = ap.parse_args()
import torch
import lora_phy_b200 as L
if args.lib: L.LIB_PATH = os.path.abspath(args.lib)
for sf in (7, 9, 12):
    ctx = L.Context(sf)
    sym = torch.randint(0, 1 << sf, (1, 64), dtype=torch.int16, device="cuda")   # 32 bytes -> 64 symbols, as performance_test
    best_m, best_d = 1e9, 1e9
    iq = ctx.modulate_batch(sym)
    rx = torch.zeros((1, 64), dtype=torch.int16, device="cuda"); met = torch.zeros((1, 16), dtype=torch.uint8, device="cuda")
    for _ in range(20):
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record(); iq = ctx.modulate_batch(sym); e1.record()
        ctx.demod_batch(iq, mode=L.DEMOD_LEGACY_DECHIRP, symbols=rx, metrics=met); e2.record()
        torch.cuda.synchronize()
        best_m = min(best_m, e0.elapsed_time(e1)); best_d = min(best_d, e1.elapsed_time(e2))
    print(f"sf={sf} one packet of 66 symbols: modulate {best_m*1e3:.0f} us, demodulate {best_d*1e3:.0f} us")
