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


def config3(args):
    """BASELINE configs[3]: 2^30 samples of continuous IQ (streams x samples), frames of SF7..SF10 mixed inside every
    stream, carrier offset uniform in +-0.25 BW, random start, SNR -10..+10 dB; one frame synchroniser per SF scans
    all of it.  Reports Msamples/s scanned per detector and overall, detection and false-alarm counts."""
    import numpy as np
    import torch
    import lora_phy_b200 as L

    sfs = (7, 8, 9, 10)
    n_streams, length = args.streams, args.length
    gen = torch.Generator(device="cuda").manual_seed(3)
    ctxs = {sf: L.Context(sf) for sf in sfs}
    max_sym = 60
    slot = ctxs[10].frame_samples(max_sym, 0) + 4096            # room for the longest frame plus a gap
    n_slots = (length - 4096) // slot
    iq = torch.zeros((n_streams, length), dtype=torch.complex64, device="cuda")
    truth = []                                                  # (slot, sf, stream indices, offsets, symbols, nsym)
    for j in range(n_slots):
        pick = torch.randint(0, len(sfs), (n_streams,), device="cuda", generator=gen)
        for k, sf in enumerate(sfs):
            idx = torch.nonzero(pick == k).flatten()
            if idx.numel() == 0:
                continue
            n = 1 << sf
            nsym = int(torch.randint(20, max_sym + 1, (1,), generator=gen, device="cuda").item())
            sym = torch.randint(0, n, (idx.numel(), nsym), dtype=torch.int16, device="cuda", generator=gen)
            fr = ctxs[sf].frame_modulate_batch(sym, amplitude=1.0, padding=0)
            off = torch.randint(0, 4096, (idx.numel(),), device="cuda", generator=gen) + j * slot
            cols = off[:, None] + torch.arange(fr.shape[1], device="cuda")[None, :]
            iq[idx[:, None], cols] = fr
            truth.append((j, sf, idx.cpu().numpy(), off.cpu().numpy(), sym.cpu().numpy(), nsym))
            del fr, cols
    f = (torch.rand((n_streams,), device="cuda", generator=gen) - 0.5) * 0.5
    for lo in range(0, n_streams, 256):                         # carrier offset, continuous phase over the stream
        k = torch.arange(length, device="cuda", dtype=torch.float64)[None, :]
        ph = (f[lo:lo + 256, None].double() * k) % 1.0
        iq[lo:lo + 256] *= torch.polar(torch.ones_like(ph, dtype=torch.float32), (2 * np.pi * ph).float())
        del k, ph
    snr = torch.rand((n_streams,), device="cuda", generator=gen) * 20.0 - 10.0
    ctxs[7].awgn_batch(iq, snr.float().contiguous(), seed=3)
    torch.cuda.synchronize()
    snr_h = snr.cpu().numpy()
    total_ms = 0.0
    report = {"workload": "configs[3]: continuous streams, SF7-SF10 frames mixed in every stream, CFO +-0.25 BW, SNR -10..+10 dB",
              "streams": n_streams, "samples_per_stream": length, "samples": int(iq.numel()), "per_sf": {}}
    for sf in sfs:
        ctx = ctxs[sf]
        ctx.set_kernel_timing(True)
        best = None
        for _ in range(args.reps):
            sym, pk, res = ctx.stream_detect_batch(iq, mtu=64, thresh_db=-6.0, max_packets=32, max_symbols=2048)
            torch.cuda.synchronize()
            ms, _ = ctx.kernel_time()
            best = ms if best is None else min(best, ms)
        total_ms += best
        pk, res = L.stream_records(pk, res)
        n = 1 << sf
        starts = [set() for _ in range(n_streams)]
        for i in range(n_streams):
            for k in range(min(int(res[i]["n_packets"]), pk.shape[1])):
                starts[i].add(int(pk[i, k]["start"]))
        sent = hit = 0
        claimed = set()
        for j, tsf, idx, off, _, nsym in truth:
            if tsf != sf:
                continue
            for i, o in zip(idx, off):
                want = int(o) + 14 * n + n // 4
                near = [s for s in starts[i] if abs(s - want) <= n // 2]
                claimed.update((i, s) for s in near)
                if snr_h[i] >= 0.0:
                    sent += 1
                    hit += 1 if near else 0
        found = sum(len(s) for s in starts)
        report["per_sf"][str(sf)] = {"detector_ms": round(best, 3), "Msamples_s": round(iq.numel() / best / 1e3, 1),
                                     "GBps": round(iq.numel() * 8 / best / 1e6, 1), "frames_snr_ge_0dB": sent,
                                     "detected": hit, "packets_reported": found, "not_at_a_sent_frame": found - len(claimed)}
        print(f"sf={sf}: detector {best:.3f} ms over {iq.numel()} samples -> {iq.numel() / best / 1e6:.1f} Gsamples/s "
              f"({iq.numel() * 8 / best / 1e6:.0f} GB/s); frames at SNR>=0 dB detected {hit}/{sent}; packets reported {found}, "
              f"of which not at a sent SF{sf} frame {found - len(claimed)}")
    print(f"all four detectors: {total_ms:.3f} ms -> {iq.numel() / total_ms / 1e6:.1f} Gsamples/s scanned by SF7..SF10 together")
    report["all_detectors_ms"] = round(total_ms, 3)
    report["value"] = iq.numel() / total_ms / 1e3
    report["unit"] = "Msamples/s scanned by the SF7, SF8, SF9 and SF10 synchronisers together"
    sent = sum(v["frames_snr_ge_0dB"] for v in report["per_sf"].values())
    report["p_detect_snr_ge_0dB"] = sum(v["detected"] for v in report["per_sf"].values()) / max(sent, 1)
    report["false_packets"] = sum(v["not_at_a_sent_frame"] for v in report["per_sf"].values())
    for c in ctxs.values():
        c.set_kernel_timing(False)
    del iq
    return report


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config3", action="store_true", help="BASELINE configs[3]: mixed-SF continuous streams")
    ap.add_argument("--length", type=int, default=1 << 18)
    ap.add_argument("--sf", type=int, default=9)
    ap.add_argument("--streams", type=int, default=4096)
    ap.add_argument("--frames", type=int, default=4)
    ap.add_argument("--symbols", type=int, default=40)
    ap.add_argument("--snr", type=float, default=10.0)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--lib", default=None, help="variant library built by scripts/ab_build.sh (same-box A/B timing)")
    args = ap.parse_args()
    if args.lib:
        import lora_phy_b200 as L0
        L0.LIB_PATH = os.path.abspath(args.lib)
    if args.config3:
        return config3(args)
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
