#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on its configs[1]: the SF7..SF12 sweep, 65,536 synthetic AWGN packets
per SF (SNR -20..+10 dB), demodulation + full decode.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--packets P] [--impl reference]

One "step" = one pass of the receive hot path over the whole sweep (six batches, one per SF): max-abs
prepass + offset estimate + fused dechirp/FFT/argmax + chain decode, with the reference's
`caller dechirp -> lora_demodulate` semantics (the pipeline its own tests/performance_test.cpp times), so
packets really decode.  Inputs are resident in HBM when the timed region starts; `e2e` repeats the same
metric through the host-buffer C-ABI call with the copies inside the timed region.  Prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-_b200"), os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

SFS = (7, 8, 9, 10, 11, 12)
PAYLOAD_LEN = 16           # bytes, uniform random
RDD = 1                    # CR 4/5, explicit header, CRC on
SNRS = (-20.0, -15.0, -10.0, -5.0, 0.0, 5.0, 10.0)
METRIC = "decoded Msamples/s (SF7-SF12 sweep, demod + full decode)"


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


def symbols_per_packet(sf):
    """LoRaEncoder.cpp:174-175 for a 16-byte payload + CRC, explicit header, CR 4/5, ppm = sf."""
    ncw = -(-((PAYLOAD_LEN + 2) * 2 + 5) // sf) * sf
    return 8 + (ncw // sf - 1) * (4 + RDD)


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: NVML polled every 10 ms from a thread (the
    timed region of a default run is ~0.3 s, shorter than the start-up of an `nvidia-smi -lms` child on an 8-GPU
    box); `nvidia-smi` is the fallback when the NVML binding is missing."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []
        self.nvml, self.handle, self.samples, self.mask, self.stop_flag = None, None, [], 0, False
        self.thread = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = self.index
            if visible:
                ids = [v.strip() for v in visible.split(",") if v.strip()]
                if self.index < len(ids) and ids[self.index].isdigit():
                    phys = int(ids[self.index])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nvml = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _poll(self):
        n = self.nvml
        while not self.stop_flag:
            try:
                self.samples.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
                self.mask |= int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
            except Exception:
                pass
            time.sleep(0.01)

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line)

    def stop(self):
        if self.nvml is not None:
            self.stop_flag = True
            self.thread.join(timeout=1.0)
            reasons = sorted(name for bit, name in self.REASONS if self.mask & bit)
            return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                    "samples": len(self.samples), "reasons": reasons, "source": "nvml"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons), "source": "nvidia-smi"}


class Workload:
    """Device-resident synthetic input for one rank: per SF, `packets` packets built with the library's own
    encoder -> modulator -> AWGN kernels (seeded; nothing is read from disk)."""

    def __init__(self, L, torch, packets, rank, sfs=SFS):
        self.L, self.torch = L, torch
        self.items = []
        for sf in sfs:
            ctx = L.Context(sf)
            cfg = L.chain_cfg(sf, RDD)
            gen = torch.Generator(device="cuda").manual_seed(1 + sf + 1000 * rank)  # mt19937 seed 1 analogue
            pay = torch.randint(0, 256, (packets, PAYLOAD_LEN), dtype=torch.uint8, device="cuda", generator=gen)
            sym = ctx.chain_encode_batch(cfg, pay)
            assert sym.shape[1] == symbols_per_packet(sf)
            iq = ctx.modulate_batch(sym)
            pick = torch.randint(0, len(SNRS), (packets,), device="cuda", generator=gen)
            snr = torch.tensor(SNRS, dtype=torch.float32, device="cuda")[pick]
            ctx.awgn_batch(iq, snr, seed=sf + 100 * rank)
            n, nsym = packets, sym.shape[1]
            self.items.append(dict(
                sf=sf, ctx=ctx, cfg=cfg, iq=iq, payload=pay, tx=sym, snr=snr,
                rx=torch.zeros((n, nsym), dtype=torch.int16, device="cuda"),
                metrics=torch.zeros((n, 16), dtype=torch.uint8, device="cuda"),
                out=torch.zeros((n, L.PAYLOAD_STRIDE), dtype=torch.uint8, device="cuda"),
                results=torch.zeros((n, 16), dtype=torch.uint8, device="cuda"),
                nsym=nsym, windows=n * (nsym + 2), samples=n * (nsym + 2) * (1 << sf)))
        torch.cuda.synchronize()
        self.samples = sum(i["samples"] for i in self.items)
        self.windows = sum(i["windows"] for i in self.items)
        self.data_symbols = sum(i["rx"].numel() for i in self.items)

    def step(self, mode):
        for i in self.items:
            i["ctx"].demod_batch(i["iq"], mode=mode, symbols=i["rx"], metrics=i["metrics"])
            i["ctx"].chain_decode_batch(i["cfg"], i["rx"], payload=i["out"], results=i["results"])

    def set_timing(self, on):
        for i in self.items:
            i["ctx"].set_kernel_timing(on)

    def set_schedule(self, overlap):
        """lora_b200_set_overlap: 2 = prepass and fused demod as items of one persistent launch (the default of the
        library), 0 = two launches back to back."""
        for i in self.items:
            i["ctx"].set_overlap(overlap)

    def kernel_time(self):
        per_sf, tot_ms, tot_k = {}, 0.0, 0
        for i in self.items:
            ms, k = i["ctx"].kernel_time()
            per_sf[i["sf"]] = (ms, k)
            tot_ms += ms; tot_k += k
        return tot_ms, tot_k, per_sf


# ---- the reference's CPU path, timed on the host cores ---------------------------------------------------
def cpu_reference_time(host_items, threads):
    """host_items: [(sf, iq complex64 [P, count], nsym)].  The reference pipeline of tests/performance_test.cpp:92-116
    -- workspace and down-chirp built once per thread OUTSIDE the timed loop, then per packet caller-dechirp ->
    lora_demodulate -- followed by the full chain decode, all in native code on `threads` host threads:
    oracle/_ref/liblora_ref.so (the unmodified src/phy/*.cpp; `ref_legacy_batch`) and oracle/_ref/liblora_pothos_ref.so
    (the unmodified LoRa-SDR/LoRaDecoder.cpp; `pothos_chain_decode_batch`), kind "reference".  Where those were never
    built the C restatement runs instead (kind "port", python thread pool over the thread-safe ctypes doors).
    -> (seconds, samples, kind, crc-ok packets)."""
    import ctypes as C
    from _oracle import ORC, REF, POTHOS, ChainCfg, chain_cfg as orc_cfg, c64p, u16p, u8p

    samples = sum(iq.size for _, iq, _ in host_items)
    if REF is not None and REF.has("legacy_batch"):
        demod = REF.fn("legacy_batch", C.c_double, [C.c_uint, C.c_uint, c64p, C.c_size_t, C.c_size_t, C.c_uint, u16p,
                                                    C.c_size_t, C.c_void_p, C.c_void_p])
        decode = None
        if POTHOS is not None and POTHOS.has("chain_decode_batch"):
            decode = POTHOS.fn("chain_decode_batch", C.c_double,
                               [C.POINTER(ChainCfg), u16p, C.c_size_t, C.c_size_t, C.c_size_t, u8p, C.c_size_t,
                                np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS"), C.c_uint])
        dt, ok = 0.0, 0
        cpu_reference_time.last = []       # per SF: (bytes posted per packet, payload bytes) for the live cross-check
        for sf, iq, nsym in host_items:
            iq = np.ascontiguousarray(iq)
            packets, count = iq.shape
            sym = np.zeros((packets, nsym), np.uint16)
            t = demod(sf, 125000, iq, packets, count, threads, sym, nsym, None, None)
            assert t >= 0.0
            dt += t
            cfg = orc_cfg(sf, RDD)
            if decode is not None:
                out = np.zeros((packets, 64), np.uint8)
                nb = np.zeros(packets, np.int64)
                t = decode(C.byref(cfg), sym, packets, nsym, nsym, out, 64, nb, threads)
                assert t >= 0.0
                dt += t
                ok += int((nb == PAYLOAD_LEN).sum())     # crcc on: a packet whose CRC fails is dropped (LoRaDecoder.cpp:371)
                cpu_reference_time.last.append((nb, out))
            else:
                t0 = time.perf_counter()
                ok += sum(ORC.chain_decode(cfg, sym[p])[2].crc_ok for p in range(packets))
                dt += time.perf_counter() - t0
        return dt, samples, "reference", ok

    from concurrent.futures import ThreadPoolExecutor

    def work(task):
        sf, iq, a, b = task
        n = 1 << sf
        down = ORC.gen_chirp(n, 1, n, 0.0, True, 1.0, 0.0, 1.0)[0]
        cfg = orc_cfg(sf, RDD)
        ok = 0
        for p in range(a, b):
            x = iq[p].reshape(-1, n)
            re = (x.real * down.real[None]).astype(np.float32) - (x.imag * down.imag[None]).astype(np.float32)
            im = (x.real * down.imag[None]).astype(np.float32) + (x.imag * down.real[None]).astype(np.float32)
            _, sym, _, _ = ORC.lora_demodulate(sf, (re + 1j * im).astype(np.complex64).reshape(-1))
            ok += ORC.chain_decode(cfg, sym)[2].crc_ok
        return ok

    tasks = []
    for sf, iq, _ in host_items:
        chunk = max(1, len(iq) // (threads * 4))
        tasks += [(sf, iq, a, min(len(iq), a + chunk)) for a in range(0, len(iq), chunk)]
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        ok = sum(ex.map(work, tasks))
    return time.perf_counter() - t0, samples, "port", ok


def performance_test_runs():
    """The reference's own benchmark, tests/performance_test.cpp UNMODIFIED (1000 packets x 32 bytes, modulate + caller
    dechirp + lora_demodulate, one thread), with tests/profiles.yaml extended to SF7-SF12 in a scratch directory:
    once linked against the reference itself (oracle/_ref/performance_test) and once against this repo's drop-in
    (bin/ref_performance_test, every call a single-packet GPU call).  Also `rx_runner` wall time on the golden capture
    for both.  -> dict, or None where the binaries were never built."""
    import tempfile
    out = {}
    cand = {"reference": os.path.join(ROOT, "oracle", "_ref", "performance_test"),
            "b200_dropin": os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-_b200", "bin", "ref_performance_test")}
    with tempfile.TemporaryDirectory() as tmp:
        os.makedirs(os.path.join(tmp, "tests"))
        with open(os.path.join(tmp, "tests", "profiles.yaml"), "w") as f:
            for sf in SFS:
                f.write(f"-\n  name: sf{sf}_bw125_cr45\n  sf: {sf}\n  bw: 125000\n  cr: 4/5\n")
        for who, exe in cand.items():
            if not os.path.exists(exe):
                continue
            try:
                subprocess.run([exe], cwd=tmp, env=dict(os.environ, RUN_ID=who), check=True, timeout=240,
                               stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                rows = open(os.path.join(tmp, "logs", f"performance_{who}.csv")).read().strip().splitlines()[1:]
                out[who] = {r.split(",")[1]: {"pps": float(r.split(",")[4]), "cycles_per_symbol": float(r.split(",")[5])}
                            for r in rows}
            except Exception as e:  # a failed side program must never cost the bench line
                out[who] = {"error": repr(e)[:200]}
    cap = os.path.join(ROOT, "tests", "golden", "capture_sf7.npz")
    runners = {"reference": os.path.join(ROOT, "oracle", "_ref", "rx_runner"),
               "b200_dropin": os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-_b200", "bin", "rx_runner")}
    if os.path.exists(cap):
        with tempfile.TemporaryDirectory() as tmp:
            path = os.path.join(tmp, "capture.iq")
            np.load(cap)["iq"].astype(np.complex64).tofile(path)
            wall = {}
            for who, exe in runners.items():
                if not os.path.exists(exe):
                    continue
                try:
                    best = None
                    for _ in range(3):
                        t0 = time.perf_counter()
                        subprocess.run([exe, f"--in={path}", "--sf=7", "--cr=1", "--bw=125000", "--report-offsets"],
                                       check=True, timeout=120, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                        dt = time.perf_counter() - t0
                        best = dt if best is None else min(best, dt)
                    wall[who] = round(best, 4)
                except Exception as e:
                    wall[who] = repr(e)[:200]
            out["rx_runner_wall_s_golden_capture"] = wall
    return out or None


def cpu_synth(packets, threads, seed=0):
    """CPU-side synthesis for the reference arm (none of this repo's kernels on that path): the chain encoder
    restatement (pinned to LoRaEncoder.cpp by tests/test_oracle_vs_pothos.py), the reference's own lora_modulate where
    oracle/_ref exists, numpy AWGN."""
    import ctypes as C
    from _oracle import ORC, REF, chain_cfg as orc_cfg, c64p, u16p
    rng = np.random.default_rng(seed)
    items = []
    for sf in SFS:
        nsym = symbols_per_packet(sf)
        count = (nsym + 2) << sf
        sym = np.zeros((packets, nsym), np.uint16)
        cfg = orc_cfg(sf, RDD)
        for p in range(packets):
            sym[p] = ORC.chain_encode(cfg, bytes(rng.integers(0, 256, PAYLOAD_LEN, dtype=np.uint8)))
        iq = np.zeros((packets, count), np.complex64)
        if REF is not None and REF.has("modulate_batch"):
            REF.fn("modulate_batch", None, [u16p, C.c_size_t, C.c_size_t, C.c_uint, C.c_uint, C.c_float, C.c_uint8, c64p,
                                            C.c_uint])(sym, packets, nsym, sf, 125000, 1.0, 0x12, iq, threads)
        else:
            for p in range(packets):
                iq[p] = ORC.lora_modulate(sym[p], sf)
        sigma = (10.0 ** (-rng.choice(SNRS, packets) / 20.0) / np.sqrt(2.0)).astype(np.float32)
        for a in range(0, packets, 256):        # in slabs: the SF12 batch is 1.7 GB
            sl = slice(a, min(packets, a + 256))
            noise = rng.standard_normal((sl.stop - sl.start, count, 2), dtype=np.float32)
            iq[sl] += (noise[..., 0] + 1j * noise[..., 1]) * sigma[sl, None]
        items.append((sf, iq, nsym))
    return items


def run_reference(args, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    packets = args.ref_packets
    items = cpu_synth(packets, threads)
    for _ in range(args.warmup):
        cpu_reference_time(items, threads)
    t, samples, kind, ok = 0.0, 0, "port", 0
    for _ in range(args.steps):
        dt, s, kind, ok = cpu_reference_time(items, threads)
        t += dt; samples += s
    v = samples / t / 1e6
    sample = f"{packets} packets per SF (SF7-12), {samples // args.steps} samples per step, {ok} CRC-ok packets"
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1]: SF7-SF12 sweep, synthetic AWGN -20..+10 dB, demod + full decode",
                   "payload_bytes": PAYLOAD_LEN, "cr": "4/5", "header": "explicit", "crc": True, "bw_hz": 125000, "osr": 1,
                   "demod_semantics": "legacy_dechirp", "sample": sample},
        "cpu_baseline": {"value": v, "unit": "Msamples/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})


def bind_to_gpu_numa_node(local):
    """Pin this rank's host threads to the CPUs of the NUMA node its GPU hangs off (sysfs: the GPU's PCI device ->
    numa_node -> cpulist), so the pinned staging buffers it allocates next are first-touched on that node and the H2D
    copies of eight ranks do not all pull from node 0 (round 1: per-GPU end-to-end rate fell 2.3x from N=2 to N=8).
    -> a small record for the JSON line; does nothing (and says so) where sysfs has no answer or there is one node."""
    info = {"numa_node": None, "cpus": None, "bound": False}
    try:
        import pynvml
        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = local
        if visible:
            ids = [v.strip() for v in visible.split(",") if v.strip()]
            if local < len(ids) and ids[local].isdigit():
                phys = int(ids[local])
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(phys)).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if len(bus.split(":")[0]) == 8:
            bus = bus[4:]                       # sysfs uses a 4-digit PCI domain
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        info["numa_node"] = node
        nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()]
        if node < 0 or len(nodes) < 2:
            return info
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            info.update(cpus=len(cpus), bound=True)
    except Exception as e:  # best effort: a box without sysfs / NVML just runs unbound
        info["error"] = repr(e)[:120]
    return info


def run_extras(args, torch, L, dist, rank, world):
    """configs[3] (continuous-stream scan) and configs[4] (multi-channel gateway) as short legs after the headline, so
    the driver's record carries them at every N.  Both are strong scaling: the input set is fixed and split over ranks."""
    import contextlib
    import types
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    out = {}
    try:
        import profile_stream
        n_streams = 4096 // world
        with contextlib.redirect_stdout(sys.stderr):   # the script narrates; this program prints ONE JSON line
            rep = profile_stream.config3(types.SimpleNamespace(streams=n_streams, length=1 << 18, reps=2))
        ms = torch.tensor([rep["all_detectors_ms"]], dtype=torch.float64, device="cuda")
        cnt = torch.tensor([sum(v["frames_snr_ge_0dB"] for v in rep["per_sf"].values()),
                            sum(v["detected"] for v in rep["per_sf"].values()), rep["false_packets"]],
                           dtype=torch.int64, device="cuda")
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.all_reduce(cnt)
        pk, _ = peaks()
        samples = rep["samples"] * world
        t = float(ms.item())
        rep.update(streams=n_streams * world, samples=samples, all_detectors_ms=round(t, 3), value=samples / t / 1e3,
                   p_detect_snr_ge_0dB=int(cnt[1]) / max(int(cnt[0]), 1), false_packets=int(cnt[2]), n_gpus=world,
                   scaling="strong",
                   roofline={"bound": "hbm", "unit": "GB/s", "peak": pk["hbm_gbs"],
                             "algorithmic_bytes": 8 * samples * 4, "note": "each of the four synchronisers reads every sample once",
                             "achieved": 8 * samples * 4 / world / t / 1e6, "frac": 8 * samples * 4 / world / t / 1e6 / pk["hbm_gbs"]})
        rep["per_sf"] = {k: v for k, v in rep["per_sf"].items()} if rank == 0 else {}
        out["stream"] = rep
    except Exception as e:
        out["stream"] = {"error": repr(e)[:300]}
    torch.cuda.empty_cache()
    try:
        import gateway
        g = gateway.run(types.SimpleNamespace(channels=4096, sf12_packets=4, steps=max(args.steps, 3), warmup=3, graph=False),
                        rank, world, dist)
        out["gateway"] = g
    except Exception as e:
        out["gateway"] = {"error": repr(e)[:300]}
    torch.cuda.empty_cache()
    return out


# ---- this repo's arm ---------------------------------------------------------------------------------------
_JSON_FD = None


def emit(line):
    """The ONE JSON line, on the process's real stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    # Libraries write to file descriptor 1 behind Python's back (NCCL prints its version banner there when the box sets
    # NCCL_DEBUG).  Keep the real stdout for the JSON line and send everything else to stderr.
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--packets", type=int, default=65536, help="packets per SF per GPU")
    ap.add_argument("--e2e-packets", type=int, default=2048)
    ap.add_argument("--cpu-packets", type=int, default=2048)
    ap.add_argument("--ref-packets", type=int, default=2048)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="legacy_dechirp", choices=["legacy_dechirp", "workspace"])
    ap.add_argument("--schedule", default="merged", choices=["merged", "two_launch"],
                    help="merged = prepass and demod items of one persistent launch (library default); two_launch = back to back")
    ap.add_argument("--no-int16", action="store_true", help="skip the extra int16-ingest end-to-end leg")
    ap.add_argument("--no-perf-test", action="store_true", help="skip the reference's own performance_test / rx_runner runs")
    ap.add_argument("--no-extras", action="store_true", help="skip the configs[3] stream scan and configs[4] gateway legs")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    args.warmup = max(args.warmup, 3)

    numa = bind_to_gpu_numa_node(local)
    import torch
    import lora_phy_b200 as L
    from lora_phy_b200.shard import gather_records_to_root
    assert torch.cuda.is_available(), "bench.py needs a CUDA device; there is no CPU fallback"
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        # NCCL writes its banner ("NCCL version ...") to stdout when the box sets NCCL_DEBUG: stdout is the JSON line's
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    mode = L.DEMOD_LEGACY_DECHIRP if args.mode == "legacy_dechirp" else L.DEMOD_WORKSPACE

    # size the batch to the HBM that is actually free (111.7 GB for 65,536 packets per SF)
    free, _ = torch.cuda.mem_get_info()
    per_packet = sum((symbols_per_packet(sf) + 2) * (1 << sf) * 8 + 1024 for sf in SFS)
    packets = args.packets
    while packets * per_packet > 0.85 * free and packets > 256:
        packets //= 2
    wl = Workload(L, torch, packets, rank)
    merged = args.schedule == "merged"
    wl.set_schedule(2 if merged else 0)
    gather_buf = None
    res_cat = None
    # decoded record of a packet as a gateway forwards it: chain result (length, CRC and header
    # flags, 16 B) + demod metrics (cfo, time offset, sync word, 16 B) + payload (240 B) = 272 B
    REC = 272
    if dist is not None:
        res_cat = torch.zeros((len(SFS) * packets, REC), dtype=torch.uint8, device="cuda")
        gather_buf = torch.zeros((world * len(SFS) * packets, REC), dtype=torch.uint8, device="cuda") if rank == 0 else None

    def one_step():
        wl.step(mode)
        if dist is not None:  # the only NVLink traffic: the decoded records travel once, to the forwarding rank
            for k, i in enumerate(wl.items):
                rows = res_cat[k * packets:(k + 1) * packets]
                rows[:, 0:16] = i["results"]
                rows[:, 16:32] = i["metrics"]
                rows[:, 32:REC] = i["out"][:, :REC - 32]
            gather_records_to_root(res_cat, dist, root=0, out=gather_buf)

    for _ in range(args.warmup):
        one_step()
    torch.cuda.synchronize()
    wl.set_timing(True)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(local)
    clocks.start()
    launches0 = L.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        one_step()
    ev1.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    clock_info = clocks.stop()
    launches = L.launch_count() - launches0
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    k_ms, k_n, per_sf = wl.kernel_time()
    # Outside the timed region, for continuity with the round-1 figure: the fused demod kernel on its own (two-launch
    # schedule, 8*N + 6 algorithmic bytes per window), two steps.
    alone = None
    if merged:
        wl.set_schedule(0)
        wl.step(mode)
        torch.cuda.synchronize()
        wl.kernel_time()
        for _ in range(2):
            wl.step(mode)
        torch.cuda.synchronize()
        a_ms, a_n, a_per_sf = wl.kernel_time()
        alone = (a_ms, a_n, a_per_sf, 2)
        wl.set_schedule(2)
    wl.set_timing(False)

    # sanity of what was decoded (not timed): CRC-ok rate and symbol agreement with what was sent
    crc_ok, sym_ok = {}, {}
    for i in wl.items:
        r = i["results"].cpu().numpy().view(L.CHAIN_RESULT).reshape(-1)
        crc_ok[i["sf"]] = round(float((r["crc_ok"] == 1).mean()), 4)
        sym_ok[i["sf"]] = round(float((i["rx"] == i["tx"]).float().mean().item()), 4)

    # e2e: the same metric through the host-buffer C-ABI call (lora_b200_receive_host), pinned HOST buffers in,
    # payloads / results / metrics out, H2D and D2H copies inside the timed region; every rank does its own share,
    # the time is the max over ranks.
    pe = min(args.e2e_packets, packets)
    pinned = []
    for i in wl.items:
        h = torch.empty((pe, i["iq"].shape[1]), dtype=torch.complex64, pin_memory=True)
        h.copy_(i["iq"][:pe])
        pinned.append((i, h.numpy(), np.zeros((pe, L.PAYLOAD_STRIDE), np.uint8), np.zeros(pe, L.CHAIN_RESULT),
                       np.zeros(pe, L.PKT_METRICS), h))
    torch.cuda.synchronize()

    def e2e_step():
        for i, h_iq, pay, res, met, _ in pinned:
            i["ctx"].receive_host(i["cfg"], h_iq, mode=mode, payload=pay, results=res, metrics=met)

    for _ in range(2):
        e2e_step()
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    dt_t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(dt_t, op=dist.ReduceOp.MAX)
    dt = float(dt_t.item())
    e_samples = sum(h.size for _, h, _, _, _, _ in pinned)
    e2e = {"value": e_samples * world * args.steps / dt / 1e6, "unit": "Msamples/s",
           "h2d_bytes_per_step": int(e_samples * 8) * world,
           "d2h_bytes_per_step": int(sum(p.nbytes + r.nbytes + m.nbytes for _, _, p, r, m, _ in pinned)) * world,
           "h2d_GBps_per_gpu": e_samples * 8 * args.steps / dt / 1e9,
           "sample": f"{pe} packets per SF per GPU through lora_b200_receive_host, pinned host buffers"}

    # The host limit of that leg, measured: the same pinned buffers copied to the device with NO kernels running, all
    # ranks at once.  `e2e` cannot beat this; at N > 2 it is the host's memory system, not PCIe per GPU, that sets it.
    for _, _, _, _, _, h in pinned[:1]:
        wl.items[0]["iq"][:pe].copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        for (i, _, _, _, _, h) in pinned:
            i["iq"][:pe].copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    dc_t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(dc_t, op=dist.ReduceOp.MAX)
    copy_only = e_samples * 8 * args.steps / float(dc_t.item()) / 1e9
    e2e["h2d_copy_only_GBps_per_gpu"] = copy_only
    e2e["frac_of_copy_only"] = e2e["h2d_GBps_per_gpu"] / copy_only if copy_only > 0 else None

    # the same end-to-end leg with int16 I/Q samples: half the PCIe bytes.  Reported
    # beside `e2e`, never instead of it -- the reference-facing call takes complex<float>.
    e2e_i16 = None
    if not args.no_int16:
        q_scale = 2048.0                                          # +-16 full scale: the -20 dB packets peak near 10
        i16 = []
        for i, h_iq, pay, res, met, h in pinned:
            q = torch.empty((pe, i["iq"].shape[1], 2), dtype=torch.int16, pin_memory=True)
            q.copy_(torch.view_as_real(i["iq"][:pe]).mul(q_scale).round_().clamp_(-32768, 32767).to(torch.int16))
            i16.append((i, q.numpy(), pay, res, met, q))
        torch.cuda.synchronize()

        def i16_step():
            for i, q, pay, res, met, _ in i16:
                i["ctx"].receive_host_i16(i["cfg"], q, scale=1.0 / q_scale, mode=mode, payload=pay, results=res, metrics=met)

        i16_step()
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            i16_step()
        dq = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(dq, op=dist.ReduceOp.MAX)
        ok16 = int(sum((res["crc_ok"] == 1).sum() for _, _, _, res, _, _ in i16))
        e2e_i16 = {"value": e_samples * world * args.steps / float(dq.item()) / 1e6, "unit": "Msamples/s",
                   "h2d_bytes_per_step": int(e_samples * 4) * world, "crc_ok_packets_rank0": ok16,
                   "call": "lora_b200_receive_host_i16 (int16 I/Q, scale 2^-11), same packets quantised"}
        del i16
    # ... and with int8 I/Q (RTL-SDR class front ends): a quarter of the PCIe bytes.  Full scale +-16 again, so the
    # quantisation step is 1/8 -- coarse next to unit-amplitude chirps: the CRC-ok count shows what it costs.
    e2e_i8 = None
    if not args.no_int16:
        q8_scale = 8.0
        i8 = []
        for i, h_iq, pay, res, met, h in pinned:
            q = torch.empty((pe, i["iq"].shape[1], 2), dtype=torch.int8, pin_memory=True)
            q.copy_(torch.view_as_real(i["iq"][:pe]).mul(q8_scale).round_().clamp_(-128, 127).to(torch.int8))
            i8.append((i, q.numpy(), pay, res, met, q))
        torch.cuda.synchronize()

        def i8_step():
            for i, q, pay, res, met, _ in i8:
                i["ctx"].receive_host_i8(i["cfg"], q, scale=1.0 / q8_scale, mode=mode, payload=pay, results=res, metrics=met)

        i8_step()
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            i8_step()
        d8 = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(d8, op=dist.ReduceOp.MAX)
        ok8 = int(sum((res["crc_ok"] == 1).sum() for _, _, _, res, _, _ in i8))
        e2e_i8 = {"value": e_samples * world * args.steps / float(d8.item()) / 1e6, "unit": "Msamples/s",
                  "h2d_bytes_per_step": int(e_samples * 2) * world, "crc_ok_packets_rank0": ok8,
                  "call": "lora_b200_receive_host_i8 (int8 I/Q, scale 1/8), same packets quantised"}
        del i8

    # what the CPU cross-check needs from the device, then the 100+ GB of inputs make room for the extra legs
    for i in wl.items:
        i["results_h"] = i["results"][:pe].cpu().numpy().view(L.CHAIN_RESULT).reshape(-1)
        i["out_h"] = i["out"][:pe, :PAYLOAD_LEN].cpu().numpy()
        for k in ("iq", "rx", "tx_full", "out", "results", "metrics", "payload", "snr"):
            i.pop(k, None)
    torch.cuda.empty_cache()
    extras = None if args.no_extras else run_extras(args, torch, L, dist, rank, world)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    pk, pk_src = peaks()
    samples_step = wl.samples * world
    value = samples_step * args.steps / (ms_total * 1e-3) / 1e6
    # Roofline.  Algorithmic bytes: 8*N*osr + 6 per window.  The dominant kernel is the demod
    # kernel of the largest SF (half of all samples of the sweep); `achieved` = its algorithmic bytes per launch over
    # its average launch duration (CUDA events recorded by the library around the launch, on the launching stream).
    # With the merged schedule the timed launch does BOTH passes of the legacy semantics (normalisation scan + estimate
    # items and fused demod items in one persistent kernel), so its algorithmic bytes are SURVEY 8(d)'s two-pass figure,
    # 16*N*osr + 6 per window; ncu shows the DRAM traffic to match (profiles/r2_traffic_merged.json).
    passes = 2 if args.mode in ("legacy", "legacy_dechirp") else 1
    kpass = passes if merged else 1
    win_bytes = lambda i: 8 * kpass * (1 << i["sf"]) + 6
    alg_bytes = sum(i["windows"] * win_bytes(i) for i in wl.items) * args.steps
    achieved_all = alg_bytes / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
    dom = max(wl.items, key=lambda i: per_sf[i["sf"]][0])
    dom_ms, dom_k = per_sf[dom["sf"]]
    dom_bytes = dom["windows"] * win_bytes(dom)
    achieved = dom_bytes / (dom_ms / max(dom_k, 1) * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    pipeline_bytes = sum(i["windows"] * (8 * passes * (1 << i["sf"]) + 6) for i in wl.items)
    pipeline_gbs = pipeline_bytes * args.steps / (ms_total * 1e-3) / 1e9    # per GPU: ms_total is the max over ranks
    # DRAM traffic of that kernel from the committed ncu --set full capture (bytes moved per algorithmic byte),
    # scaled to this launch's size; null when no capture of this SF's kernel is on file.
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "r2_traffic_merged.json" if merged else "r2_traffic.json")) as f:
            tj = json.load(f)
        if int(tj["sf"]) == dom["sf"]:
            traffic = (tj["dram_bytes_read"] + tj["dram_bytes_write"]) / tj["algorithmic_bytes"] * dom_bytes
            traffic_src = tj["source"]
    except (OSError, KeyError, ValueError):
        pass
    per_sf_out = {}
    for i in wl.items:
        ms_sf, k_sf = per_sf[i["sf"]]
        b = i["windows"] * win_bytes(i) * max(k_sf, 1)
        per_sf_out[str(i["sf"])] = {"kernel_ms": round(ms_sf / max(k_sf, 1), 4),
                                    "GBps": round(b / (ms_sf * 1e-3) / 1e9, 1) if ms_sf > 0 else 0.0,
                                    "Msamples_s": round(i["samples"] * max(k_sf, 1) / (ms_sf * 1e-3) / 1e6, 1) if ms_sf > 0 else 0.0,
                                    "crc_ok": crc_ok[i["sf"]], "symbols_eq_tx": sym_ok[i["sf"]]}

    # cpu_baseline: the same packets as the e2e sample, the reference's CPU path on this box's host cores.  Whatever
    # happens in this leg, the GPU measurement above is printed: a failure becomes "cpu_baseline": null + "cpu_error".
    cpu, cpu_error = None, None
    if world == 1:
        try:
            pc = min(args.cpu_packets, pe)
            threads = os.cpu_count() or 1
            host_items = [(i["sf"], h_iq[:pc], i["nsym"]) for i, h_iq, _, _, _, _ in pinned]
            cpu_reference_time([(sf, iq[:64], n) for sf, iq, n in host_items], threads)          # warm the pages and threads
            dt_c, s_c, kind, ok_cpu = cpu_reference_time(host_items, threads)
            # live parity check on this sample: the packets the reference's CPU code (unmodified lora_demodulate +
            # LoRaDecoder::work) delivers, and their payload bytes, against what the GPU delivered for the same packets
            ok_gpu, differ = 0, 0
            last = getattr(cpu_reference_time, "last", [])
            for k, i in enumerate(wl.items):
                r = i["results_h"][:pc]
                ok_gpu += int((r["n_bytes"] == PAYLOAD_LEN).sum())
                if k < len(last):
                    nb, out = last[k]
                    g_out = i["out_h"][:pc]
                    both = (nb >= 0) & (r["n_bytes"] >= 0)
                    differ += int(((nb >= 0) != (r["n_bytes"] >= 0)).sum()) + int((nb[both] != r["n_bytes"][both]).sum())
                    same_len = both & (nb == PAYLOAD_LEN) & (r["n_bytes"] == PAYLOAD_LEN)
                    differ += int((out[same_len, :PAYLOAD_LEN] != g_out[same_len]).any(axis=1).sum())
            cpu = {"value": s_c / dt_c / 1e6, "unit": "Msamples/s", "cores": threads, "kind": kind,
                   "sample": f"first {pc} packets per SF of the same workload ({s_c} samples, {dt_c:.1f} s in the timed loops)",
                   "pipeline": "tests/performance_test.cpp:92-116 per packet (workspace + down-chirp once per thread, outside "
                               "the timed loop) + LoRaDecoder::work, native threads",
                   "delivered_packets_cpu": int(ok_cpu), "delivered_packets_gpu_same_sample": ok_gpu,
                   "packets_that_differ": differ if last else None}
            if not args.no_perf_test:
                cpu["performance_test"] = performance_test_runs()
        except Exception as e:
            cpu_error = repr(e)[:300]

    line = {
        "metric": METRIC, "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "msymbols_per_s": wl.data_symbols * world * args.steps / (ms_total * 1e-3) / 1e6,
        "config": {"workload": "configs[1]: SF7-SF12 sweep, synthetic AWGN -20..+10 dB, demod + full decode",
                   "packets_per_sf_per_gpu": packets, "payload_bytes": PAYLOAD_LEN, "cr": "4/5",
                   "header": "explicit", "crc": True, "bw_hz": 125000, "osr": 1,
                   "demod_semantics": args.mode, "parallelism": f"{world} independent shards, no data-path collective; one gather of 272-byte decoded records to rank 0 per step",
                   "l2": "inputs (>= 2.3 GB per batch) are larger than the 126 MB L2; no flush needed",
                   "per_sf": per_sf_out},
        "clocks": clock_info, "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
                     "frac": achieved / pk["hbm_gbs"], "traffic": traffic, "peak_source": pk_src,
                     "kernel": (f"receive_merged_kernel<SF={dom['sf']}> (scan + estimate items and fused dechirp + rotate + FFT + "
                                "argmax items of one persistent launch), timed with its margin-recheck launch" if merged else
                                f"demod_blocked_kernel<SF={dom['sf']}, fused flavour> (dechirp + rotate + FFT + argmax), "
                                "timed with its margin-recheck launch"),
                     "algorithmic_bytes_per_window": f"{8 * kpass}*N + 6 (SURVEY.md 8d, {'two-pass legacy semantics' if kpass == 2 else 'one pass'})",
                     "algorithmic_bytes_per_launch": dom_bytes, "kernel_ms_per_launch": dom_ms / max(dom_k, 1),
                     "traffic_source": traffic_src,
                     "all_sf": {"achieved": achieved_all, "frac": achieved_all / pk["hbm_gbs"],
                                "kernel_ms_per_step": k_ms / args.steps, "kernel_launches": k_n},
                     # the whole step against the same peak: legacy semantics read every sample twice (normalisation
                     # scan, then demod), 16*N*osr + 6 algorithmic bytes per window
                     "whole_step": {"achieved": pipeline_gbs, "frac": pipeline_gbs / pk["hbm_gbs"],
                                    "algorithmic_bytes_per_step": pipeline_bytes, "ms_per_step": ms_total / args.steps / 1.0}},
    }
    if alone is not None:
        a_ms, a_n, a_per_sf, a_steps = alone
        a_bytes = sum(i["windows"] * (8 * (1 << i["sf"]) + 6) for i in wl.items) * a_steps
        d_ms, d_k = a_per_sf[dom["sf"]]
        d_bytes = dom["windows"] * (8 * (1 << dom["sf"]) + 6)
        d_ach = d_bytes / (d_ms / max(d_k, 1) * 1e-3) / 1e9 if d_ms > 0 else 0.0
        line["roofline"]["demod_kernel_alone"] = {
            "what": "the fused demod kernel of the two-launch schedule (lora_b200_set_overlap(0)), 8*N + 6 bytes per window, "
                    "measured over two extra steps outside the timed region: the round-1 figure's successor",
            "achieved": d_ach, "frac": d_ach / pk["hbm_gbs"], "kernel_ms_per_launch": d_ms / max(d_k, 1),
            "all_sf": {"achieved": a_bytes / (a_ms * 1e-3) / 1e9 if a_ms > 0 else 0.0,
                       "frac": (a_bytes / (a_ms * 1e-3) / 1e9 if a_ms > 0 else 0.0) / pk["hbm_gbs"],
                       "kernel_ms_per_step": a_ms / a_steps}}
    line["config"]["schedule"] = args.schedule
    if e2e:
        line["e2e"] = e2e
    if e2e_i16:
        line["e2e_int16"] = e2e_i16
    if e2e_i8:
        line["e2e_int8"] = e2e_i8
    line["cpu_baseline"] = cpu
    line["host"] = {"numa": numa, "cpus": os.cpu_count()}
    if extras is not None:
        line["extra"] = extras
    if cpu_error:
        line["cpu_error"] = cpu_error
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
