"""ctypes doors into the two CPU checkers.  TEST INFRASTRUCTURE ONLY.

`ORC` is oracle/liblora_oracle.so (the plain-C restatement, prefix ``orc_``); `REF` is
oracle/_ref/liblora_ref.so (the unmodified reference compiled from /root/reference, prefix ``ref_``)
or None where it was never built.  Both expose the same signatures, so `Lib.call("fft", ...)` style
helpers below work on either.  Nothing in the product imports this module.
"""
import ctypes as C
import os
import subprocess
import threading

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

c64p = np.ctypeslib.ndpointer(np.complex64, flags="C_CONTIGUOUS")
u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


class Params(C.Structure):
    _fields_ = [("sf", C.c_uint), ("bw_hz", C.c_uint), ("osr", C.c_uint), ("hann", C.c_int),
                ("has_window", C.c_int), ("sync_word", C.c_uint8)]


class Metrics(C.Structure):
    _fields_ = [("crc_ok", C.c_int), ("cfo", C.c_float), ("time_offset", C.c_float),
                ("sync_word", C.c_uint8)]


class ChainCfg(C.Structure):
    _fields_ = [("sf", C.c_uint), ("ppm", C.c_uint), ("rdd", C.c_uint), ("explicit_header", C.c_int),
                ("crc", C.c_int), ("whitening", C.c_int), ("interleaving", C.c_int),
                ("error_check", C.c_int), ("hdr_out", C.c_int), ("data_length", C.c_uint)]


class StdCfg(C.Structure):
    """The deployed LoRa coding chain (oracle/std_chain.c; same layout as lora_b200_std_cfg)."""
    _fields_ = [("sf", C.c_uint), ("cr", C.c_uint), ("ldro", C.c_int), ("implicit_header", C.c_int),
                ("has_crc", C.c_int), ("payload_len", C.c_uint)]


class StdInfo(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("n_bytes", "hdr_ok", "hdr_length", "hdr_cr", "has_crc", "crc_ok", "corrected",
                                        "uncorrectable", "symbols_used")]


class ChainInfo(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("drop_line", "parity_error", "uncorrectable", "hdr_length",
                                        "hdr_has_crc", "hdr_rdd", "hdr_checksum", "crc_present", "crc_ok")]


def build_oracle():
    """(Re)build the checkers with oracle/Makefile; _ref only where /root/reference exists."""
    subprocess.run(["make", "-C", ORACLE_DIR, "--no-print-directory"], check=True,
                   stdout=subprocess.DEVNULL)


class Lib:
    """One checker library; `prefix` is 'orc' or 'ref'."""

    def __init__(self, path, prefix):
        self.lib = C.CDLL(path)
        self.prefix = prefix
        self._bound = {}
        self._lock = threading.Lock()

    def fn(self, name, restype, argtypes):
        """The exported function `<prefix>_<name>` with its prototype bound ONCE.  The returned object is a private
        ctypes function pointer per (name, prototype): `restype`/`argtypes` are never assigned on a call path, so
        worker threads calling the same door cannot rebuild each other's converters (the round-1 bench crash)."""
        key = (name, restype, tuple(argtypes))
        f = self._bound.get(key)
        if f is None:
            with self._lock:
                f = self._bound.get(key)
                if f is None:
                    proto = C.CFUNCTYPE(restype, *argtypes)
                    f = proto((f"{self.prefix}_{name}", self.lib))
                    self._bound[key] = f
        return f

    def has(self, name):
        return hasattr(self.lib, f"{self.prefix}_{name}")

    # ---- DSP ------------------------------------------------------------------------------------
    def gen_chirp(self, N, osr, NN, f0, down, ampl, phase, bw_scale=1.0):
        out = np.zeros(NN, np.complex64)
        ph = C.c_float(phase)
        self.fn("gen_chirp", C.c_int, [c64p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_float,
                                       C.POINTER(C.c_float), C.c_float])(
            out, N, osr, NN, f0, int(down), ampl, C.byref(ph), bw_scale)
        return out, np.float32(ph.value)

    def twiddles(self, n, inverse=False):
        tw = np.zeros(n, np.complex64)
        self.fn("fft_twiddles", None, [c64p, C.c_int, C.c_int])(tw, n, int(inverse))
        return tw

    def factor(self, n):
        r, m = np.zeros(32, np.int32), np.zeros(32, np.int32)
        k = self.fn("fft_factor", C.c_int, [C.c_int, i32p, i32p])(n, r, m)
        return list(r[:k]), list(m[:k])

    def fft(self, x, inverse=False):
        x = np.ascontiguousarray(x, np.complex64)
        y = np.zeros_like(x)
        if self.prefix == "orc":
            tw = self.twiddles(len(x), inverse)
            self.fn("fft", None, [c64p, C.c_int, C.c_int, c64p, c64p])(tw, len(x), int(inverse), x, y)
        else:
            self.fn("fft", None, [C.c_int, C.c_int, c64p, c64p])(len(x), int(inverse), x, y)
        return y

    def detect(self, x):
        """-> (index, power, power_avg, f_index, spectrum)"""
        x = np.ascontiguousarray(x, np.complex64)
        spec = np.zeros_like(x)
        p, pa, fi = C.c_float(), C.c_float(), C.c_float()
        fp = C.POINTER(C.c_float)
        if self.prefix == "orc":
            tw = self.twiddles(len(x))
            idx = self.fn("detect", C.c_size_t, [c64p, C.c_size_t, c64p, c64p, fp, fp, fp, fp])(
                tw, len(x), x, spec, C.byref(p), C.byref(pa), C.byref(fi), None)
        else:
            idx = self.fn("detect", C.c_size_t, [C.c_size_t, c64p, c64p, fp, fp, fp])(
                len(x), x, spec, C.byref(p), C.byref(pa), C.byref(fi))
        return int(idx), np.float32(p.value), np.float32(pa.value), np.float32(fi.value), spec

    # ---- lightweight PHY ------------------------------------------------------------------------
    def lora_encode(self, data):
        data = np.frombuffer(bytes(data), np.uint8).copy()
        sym = np.zeros(2 * len(data), np.uint16)
        n = self.fn("lora_encode", C.c_size_t, [u8p, C.c_size_t, u16p])(data, len(data), sym)
        return sym[:n]

    def lora_decode(self, sym):
        sym = np.ascontiguousarray(sym, np.uint16)
        out = np.zeros(len(sym) // 2 + 1, np.uint8)
        n = self.fn("lora_decode", C.c_size_t, [u16p, C.c_size_t, u8p])(sym, len(sym), out)
        return out[:n]

    def decode(self, sym):
        """phy.cpp decode(): -> (bytes, crc_ok)"""
        sym = np.ascontiguousarray(sym, np.uint16)
        out = np.zeros(len(sym) // 2 + 1, np.uint8)
        ok = C.c_int()
        n = self.fn("decode", C.c_long, [u16p, C.c_size_t, u8p, C.c_size_t, C.POINTER(C.c_int)])(
            sym, len(sym), out, len(out), C.byref(ok))
        return out[:max(n, 0)], bool(ok.value)

    def lora_modulate(self, sym, sf, osr=1, bw_hz=125000, amplitude=1.0, sync=0x12):
        sym = np.ascontiguousarray(sym, np.uint16)
        out = np.zeros((len(sym) + 2) * (1 << sf) * osr, np.complex64)
        # a zero-length array still needs a valid pointer for the ndpointer check
        arg = sym if len(sym) else np.zeros(1, np.uint16)
        self.fn("lora_modulate", C.c_size_t, [u16p, C.c_size_t, c64p, C.c_uint, C.c_uint, C.c_uint,
                                              C.c_float, C.c_uint8])(
            arg, len(sym), out, sf, osr, bw_hz, amplitude, sync)
        return out

    def estimate_offsets(self, params, samples):
        samples = np.ascontiguousarray(samples, np.complex64)
        cfo, to = C.c_float(0), C.c_float(0)
        fp = C.POINTER(C.c_float)
        self.fn("estimate_offsets", None, [C.POINTER(Params), c64p, C.c_size_t, fp, fp])(
            C.byref(params), samples, len(samples), C.byref(cfo), C.byref(to))
        return np.float32(cfo.value), np.float32(to.value)

    def compensate_offsets(self, params, cfo, time_offset, samples):
        s = np.array(samples, np.complex64)
        self.fn("compensate_offsets", None, [C.POINTER(Params), C.c_float, C.c_float, c64p, C.c_size_t])(
            C.byref(params), cfo, time_offset, s, len(s))
        return s

    def demodulate(self, params, iq, want_peaks=False):
        """phy.cpp demodulate(): -> (ret, symbols, Metrics[, peaks])"""
        iq = np.ascontiguousarray(iq, np.complex64)
        step = (1 << params.sf) * max(params.osr, 1)
        total = len(iq) // step
        sym = np.zeros(max(total, 2), np.uint16)
        m = Metrics()
        if self.prefix == "orc":
            peaks = np.zeros(max(total, 1), np.float32)
            r = self.fn("demodulate", C.c_long, [C.POINTER(Params), c64p, C.c_size_t, u16p, C.c_size_t,
                                                 C.POINTER(Metrics), f32p])(
                C.byref(params), iq, len(iq), sym, len(sym), C.byref(m), peaks)
        else:
            peaks = None
            r = self.fn("demodulate", C.c_long, [C.POINTER(Params), c64p, C.c_size_t, u16p, C.c_size_t,
                                                 C.POINTER(Metrics)])(
                C.byref(params), iq, len(iq), sym, len(sym), C.byref(m))
        out = (int(r), sym[:max(r, 0)], m)
        return out + (peaks[:total],) if want_peaks else out

    def lora_demodulate(self, sf, samples, osr=1, hann=False, scratch=True, want_peaks=False):
        """LoRaDemod.cpp lora_demodulate(): -> (ret, symbols, sync, Metrics[, peaks])"""
        samples = np.ascontiguousarray(samples, np.complex64)
        total = len(samples) // ((1 << sf) * osr)
        sym = np.zeros(max(total, 1), np.uint16)
        sync = C.c_uint8()
        m = Metrics()
        scr = np.zeros(max(len(samples), 1), np.complex64) if scratch else None
        scr_arg = scr.ctypes.data_as(C.c_void_p) if scratch else None
        scr_len = len(samples) if scratch else 0
        base = [C.c_uint, C.c_int, c64p, C.c_size_t, u16p, C.c_uint, C.POINTER(C.c_uint8), C.c_void_p,
                C.c_size_t, C.POINTER(Metrics)]
        if self.prefix == "orc":
            peaks = np.zeros(max(total, 1), np.float32)
            r = self.fn("lora_demodulate", C.c_size_t, base + [f32p])(
                sf, int(hann), samples, len(samples), sym, osr, C.byref(sync), scr_arg, scr_len,
                C.byref(m), peaks)
        else:
            peaks = None
            r = self.fn("lora_demodulate", C.c_size_t, base)(
                sf, int(hann), samples, len(samples), sym, osr, C.byref(sync), scr_arg, scr_len, C.byref(m))
        out = (int(r), sym[:r], int(sync.value), m)
        return out + (peaks[:total],) if want_peaks else out


    # ---- full coding chain (oracle only: the Pothos blocks do not compile here) ---------------------
    def chain_encode(self, cfg, payload):
        payload = np.frombuffer(bytes(payload), np.uint8).copy()
        arg = payload if len(payload) else np.zeros(1, np.uint8)
        f = self.fn("chain_encode", C.c_size_t, [C.POINTER(ChainCfg), u8p, C.c_size_t, C.c_void_p, C.c_size_t])
        need = f(C.byref(cfg), arg, len(payload), None, 0)
        sym = np.zeros(max(need, 1), np.uint16)
        n = f(C.byref(cfg), arg, len(payload), sym.ctypes.data_as(C.c_void_p), len(sym))
        return sym[:n]

    def chain_decode(self, cfg, symbols):
        """-> (ret, bytes, ChainInfo); ret < 0 means dropped / no output (see lora_oracle.h)."""
        symbols = np.ascontiguousarray(symbols, np.uint16)
        arg = symbols if len(symbols) else np.zeros(1, np.uint16)
        out = np.zeros(4 * len(symbols) + 64, np.uint8)
        info = ChainInfo()
        r = self.fn("chain_decode", C.c_long, [C.POINTER(ChainCfg), u16p, C.c_size_t, u8p, C.c_size_t,
                                               C.POINTER(ChainInfo)])(
            C.byref(cfg), arg, len(symbols), out, len(out), C.byref(info))
        return int(r), out[:max(r, 0)], info


    # ---- the deployed LoRa coding chain (oracle/std_chain.c) ------------------------------------------
    def std_symbol_count(self, cfg, length):
        return self.fn("std_symbol_count", C.c_size_t, [C.POINTER(StdCfg), C.c_size_t])(C.byref(cfg), length)

    def std_encode(self, cfg, payload):
        payload = np.ascontiguousarray(payload, np.uint8)
        n = self.std_symbol_count(cfg, len(payload))
        out = np.zeros(max(n, 1), np.uint16)
        arg = payload if len(payload) else np.zeros(1, np.uint8)
        r = self.fn("std_encode", C.c_size_t, [C.POINTER(StdCfg), u8p, C.c_size_t, u16p, C.c_size_t])(
            C.byref(cfg), arg, len(payload), out, n)
        return out[:r]

    def std_decode(self, cfg, symbols):
        """-> (ret, bytes, StdInfo)"""
        symbols = np.ascontiguousarray(symbols, np.uint16)
        arg = symbols if len(symbols) else np.zeros(1, np.uint16)
        out = np.zeros(256, np.uint8)
        info = StdInfo()
        r = self.fn("std_decode", C.c_long, [C.POINTER(StdCfg), u16p, C.c_size_t, u8p, C.c_size_t, C.POINTER(StdInfo)])(
            C.byref(cfg), arg, len(symbols), out, len(out), C.byref(info))
        return int(r), out[:max(r, 0)], info

    def std_header_checksum(self, n0, n1, n2):
        return int(self.fn("std_header_checksum", C.c_ubyte, [C.c_uint, C.c_uint, C.c_uint])(n0, n1, n2))

    def std_crc(self, payload):
        payload = np.ascontiguousarray(payload, np.uint8)
        arg = payload if len(payload) else np.zeros(1, np.uint8)
        return int(self.fn("std_crc", C.c_ushort, [u8p, C.c_size_t])(arg, len(payload)))

    # ---- upstream Pothos blocks restated (oracle/stream.c) -------------------------------------------
    def frame_modulate(self, sf, symbols, sync=0x12, padding=1, ampl=0.3, ovs=1):
        symbols = np.ascontiguousarray(symbols, np.uint16)
        f = self.fn("frame_modulate", C.c_size_t, [C.c_uint, C.c_uint, u16p, C.c_size_t, C.c_ubyte, C.c_uint,
                                                   C.c_float, C.c_void_p])
        n = f(sf, ovs, symbols, len(symbols), sync, padding, ampl, None)
        out = np.zeros(n, np.complex64)
        f(sf, ovs, symbols, len(symbols), sync, padding, ampl, out.ctypes.data_as(C.c_void_p))
        return out

    def stream_tables(self, sf):
        n = 1 << sf
        up, down, fine = np.zeros(n, np.complex64), np.zeros(n, np.complex64), np.zeros(n * 128, np.complex64)
        self.fn("stream_tables", None, [C.c_uint, c64p, c64p, c64p])(sf, up, down, fine)
        return up, down, fine

    def stream_demod(self, sf, samples, sync=0x12, thresh=-30.0, mtu=256, pkt_cap=64, frame_header=0):
        """-> (list of int16 symbol arrays, starts, freq errors, samples consumed)"""
        samples = np.ascontiguousarray(samples, np.complex64)
        sym = np.zeros(pkt_cap * mtu, np.int16)
        plen, pstart, perr = np.zeros(pkt_cap, np.uint32), np.zeros(pkt_cap, np.uint64), np.zeros(pkt_cap, np.int32)
        consumed = C.c_uint64()
        tail = [c64p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                C.POINTER(C.c_uint64)]
        rest = (samples, len(samples), sym.ctypes.data, len(sym), plen.ctypes.data, pstart.ctypes.data, perr.ctypes.data,
                pkt_cap, C.byref(consumed))
        if self.has("stream_demod_hdr"):
            n = self.fn("stream_demod_hdr", C.c_size_t, [C.c_uint, C.c_ubyte, C.c_float, C.c_size_t, C.c_int] + tail)(
                sf, sync, thresh, mtu, frame_header, *rest)
        else:   # the compiled Pothos block: the reference's behaviour only
            assert frame_header == 0
            n = self.fn("stream_demod", C.c_size_t, [C.c_uint, C.c_ubyte, C.c_float, C.c_size_t] + tail)(
                sf, sync, thresh, mtu, *rest)
        assert n != C.c_size_t(-1).value, "the block threw"
        n = min(n, pkt_cap)
        pk, o = [], 0
        for k in range(n):
            pk.append(sym[o:o + plen[k]].copy())
            o += int(plen[k])
        return pk, pstart[:n].copy(), perr[:n].copy(), int(consumed.value)


def chain_cfg(sf, rdd, ppm=0, explicit=True, crc=True, whitening=True, interleaving=True,
              error_check=False, hdr_out=False, data_length=8):
    return ChainCfg(sf, ppm, rdd, int(explicit), int(crc), int(whitening), int(interleaving),
                    int(error_check), int(hdr_out), data_length)


def std_cfg(sf, cr=1, ldro=False, implicit=False, crc=True, payload_len=0):
    return StdCfg(sf, cr, int(ldro), int(implicit), int(crc), payload_len)


def _load(path, prefix):
    return Lib(path, prefix) if os.path.exists(path) else None


_orc_path = os.path.join(ORACLE_DIR, "liblora_oracle.so")
if not os.path.exists(_orc_path):
    build_oracle()
ORC = Lib(_orc_path, "orc")
REF = _load(os.path.join(ORACLE_DIR, "_ref", "liblora_ref.so"), "ref")
# the four upstream Pothos blocks, unmodified, behind oracle/pothos_stub (doors: chain_encode, chain_decode,
# frame_modulate, stream_demod -- same signatures as the orc_ ones)
POTHOS = _load(os.path.join(ORACLE_DIR, "_ref", "liblora_pothos_ref.so"), "pothos")
