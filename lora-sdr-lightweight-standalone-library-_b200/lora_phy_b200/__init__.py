"""lora_phy_b200 -- Python host layer over the C ABI of liblora_phy_b200.so (include/lora_phy_b200.h).

Nothing here computes: every method forwards to a CUDA entry point.  torch is used only to own device
memory and streams.  There is no CPU fallback -- if the shared library or a CUDA device is missing,
loading or context creation raises.

The method names follow the reference API they stand in for (include/lora_phy/phy.hpp:102-215):
`demodulate`, `lora_demodulate`, `estimate_offsets`, `decode`, `lora_encode`, `lora_modulate`, plus the
full coding chain of LoRa-SDR/LoRaEncoder.cpp / LoRaDecoder.cpp as `chain_encode` / `chain_decode`.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "liblora_phy_b200.so")

DEMOD_WORKSPACE, DEMOD_LEGACY, DEMOD_LEGACY_DECHIRP = 0, 1, 2
PRECISION_FAST, PRECISION_EXACT = 0, 1
PAYLOAD_STRIDE = 264


class Config(C.Structure):
    _fields_ = [("sf", C.c_uint), ("bw_hz", C.c_uint), ("osr", C.c_uint), ("window", C.c_int),
                ("device", C.c_int)]


class ChainCfg(C.Structure):
    _fields_ = [("sf", C.c_uint), ("ppm", C.c_uint), ("rdd", C.c_uint), ("explicit_header", C.c_int),
                ("crc", C.c_int), ("whitening", C.c_int), ("interleaving", C.c_int),
                ("error_check", C.c_int), ("hdr_out", C.c_int), ("data_length", C.c_uint)]


PKT_METRICS = np.dtype([("cfo", "<f4"), ("time_offset", "<f4"), ("max_amp", "<f4"), ("sync_word", "u1"),
                        ("normalised", "u1"), ("n_symbols", "<u2")])
CHAIN_RESULT = np.dtype([("n_bytes", "<i2"), ("drop_line", "<i2"), ("parity_error", "u1"),
                         ("uncorrectable", "u1"), ("hdr_length", "u1"), ("hdr_has_crc", "u1"),
                         ("hdr_rdd", "u1"), ("hdr_checksum", "u1"), ("crc_present", "u1"), ("crc_ok", "u1"),
                         ("reserved", "u1", 4)])
class StdCfg(C.Structure):
    """lora_b200_std_cfg: the LoRa coding chain as deployed on the air (csrc/std_kernels.cu)."""
    _fields_ = [("sf", C.c_uint), ("cr", C.c_uint), ("ldro", C.c_int), ("implicit_header", C.c_int),
                ("has_crc", C.c_int), ("payload_len", C.c_uint)]


STD_RESULT = np.dtype([("n_bytes", "<i2"), ("symbols_used", "<u2"), ("hdr_ok", "u1"), ("hdr_length", "u1"),
                       ("hdr_cr", "u1"), ("has_crc", "u1"), ("crc_ok", "u1"), ("corrected", "u1"),
                       ("uncorrectable", "u1"), ("reserved", "u1", (5,))])
assert PKT_METRICS.itemsize == 16 and CHAIN_RESULT.itemsize == 16 and STD_RESULT.itemsize == 16


class StreamCfg(C.Structure):  # lora_b200_stream_cfg
    _fields_ = [("sync", C.c_uint8), ("thresh_db", C.c_float), ("mtu", C.c_uint32), ("max_packets", C.c_uint32),
                ("max_symbols", C.c_uint32), ("frame_header", C.c_uint32)]


# lora_b200_stream_packet / lora_b200_stream_result as numpy record layouts
STREAM_PACKET = np.dtype([("start", "<u8"), ("first_symbol", "<u4"), ("n_symbols", "<u4"), ("freq_error", "<i4"),
                          ("reserved", "<u4")])
STREAM_RESULT = np.dtype([("consumed", "<u8"), ("n_packets", "<u4"), ("n_symbols", "<u4")])


DETECT_RESULT = np.dtype([("index", "<u4"), ("power", "<f4"), ("power_avg", "<f4"), ("f_index", "<f4")])


class LoraB200Error(RuntimeError):
    pass


def std_cfg(sf, cr=1, ldro=False, implicit=False, crc=True, payload_len=0):
    return StdCfg(sf, cr, int(ldro), int(implicit), int(crc), payload_len)


def chain_cfg(sf, rdd, ppm=0, explicit=True, crc=True, whitening=True, interleaving=True,
              error_check=False, hdr_out=False, data_length=8):
    return ChainCfg(sf, ppm, rdd, int(explicit), int(crc), int(whitening), int(interleaving),
                    int(error_check), int(hdr_out), data_length)


_lib = None

_SIGNATURES = {
    "lora_b200_ctx_create": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "lora_b200_ctx_destroy": (None, [C.c_void_p]),
    "lora_b200_last_error": (C.c_char_p, []),
    "lora_b200_launch_count": (C.c_ulonglong, []),
    "lora_b200_set_kernel_timing": (C.c_int, [C.c_void_p, C.c_int]),
    "lora_b200_kernel_time": (C.c_int, [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_ulonglong)]),
    "lora_b200_set_precision": (C.c_int, [C.c_void_p, C.c_int]),
    "lora_b200_set_overlap": (C.c_int, [C.c_void_p, C.c_int]),
    "lora_b200_std_decode_batch": (C.c_int, [C.c_void_p, C.POINTER(StdCfg), C.c_void_p, C.c_size_t, C.c_size_t,
                                             C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lora_b200_std_encode_batch": (C.c_int, [C.c_void_p, C.POINTER(StdCfg), C.c_void_p, C.c_size_t, C.c_size_t,
                                             C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t), C.c_void_p]),
    "lora_b200_std_stream_decode_batch": (C.c_int, [C.c_void_p, C.POINTER(StdCfg), C.c_void_p, C.c_void_p, C.c_void_p,
                                                    C.c_size_t, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p,
                                                    C.c_void_p]),
    "lora_b200_std_receive_stream_host": (C.c_int, [C.c_void_p, C.POINTER(StdCfg), C.c_void_p, C.c_void_p, C.c_size_t,
                                                    C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32)]),
    "lora_b200_recheck_stats": (C.c_int, [C.c_void_p, C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]),
    "lora_b200_ctx_table": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]),
    "lora_b200_demod_batch": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                        C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lora_b200_demod_batch_static": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                               C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lora_b200_estimate_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                           C.c_void_p, C.c_void_p]),
    "lora_b200_decode_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                         C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]),
    "lora_b200_chain_decode_batch": (C.c_int, [C.c_void_p, C.POINTER(ChainCfg), C.c_void_p, C.c_size_t,
                                               C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lora_b200_encode_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                         C.c_void_p, C.c_size_t, C.c_void_p]),
    "lora_b200_chain_encode_batch": (C.c_int, [C.c_void_p, C.POINTER(ChainCfg), C.c_void_p, C.c_size_t,
                                               C.c_size_t, C.c_size_t, C.c_void_p, C.c_size_t,
                                               C.POINTER(C.c_size_t), C.c_void_p]),
    "lora_b200_modulate_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                           C.c_float, C.c_uint8, C.c_void_p, C.c_size_t, C.c_void_p]),
    "lora_b200_awgn_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                       C.c_void_p, C.c_uint64, C.c_void_p]),
    "lora_b200_encode_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "lora_b200_decode_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t),
                                        C.POINTER(C.c_int)]),
    "lora_b200_compensate_host": (C.c_int, [C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_size_t]),
    "lora_b200_demodulate_host": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p,
                                            C.c_size_t, C.POINTER(C.c_size_t), C.c_void_p]),
    "lora_b200_estimate_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_float),
                                          C.POINTER(C.c_float)]),
    "lora_b200_modulate_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_float, C.c_uint8,
                                          C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "lora_b200_receive_host": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(ChainCfg), C.c_void_p, C.c_size_t,
                                         C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lora_b200_receive_host_i16": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(ChainCfg), C.c_void_p, C.c_size_t,
                                             C.c_size_t, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lora_b200_receive_host_i8": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(ChainCfg), C.c_void_p, C.c_size_t,
                                            C.c_size_t, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lora_b200_fft_detect_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lora_b200_fft_detect_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p]),
    "lora_b200_frame_samples": (C.c_size_t, [C.c_void_p, C.c_size_t, C.c_uint]),
    "lora_b200_frame_modulate_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_float,
                                                 C.c_uint8, C.c_uint, C.c_void_p, C.c_size_t, C.c_void_p]),
    "lora_b200_stream_detect_batch": (C.c_int, [C.c_void_p, C.POINTER(StreamCfg), C.c_void_p, C.c_size_t, C.c_size_t,
                                                C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lora_b200_stream_decode_batch": (C.c_int, [C.c_void_p, C.POINTER(ChainCfg), C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_size_t, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
}
EXPORTS = tuple(_SIGNATURES)


def load_library():
    """Load liblora_phy_b200.so (built in tree by __graft_entry__.build / make).  Raises if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise LoraB200Error(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; "
                                "g.build()'`; there is no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            f = getattr(lib, name)
            f.restype, f.argtypes = res, args
        _lib = lib
    return _lib


def launch_count():
    return int(load_library().lora_b200_launch_count())


def _check(rc):
    if rc != 0:
        raise LoraB200Error(f"lora_b200 error {rc}: {load_library().lora_b200_last_error().decode()}")


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class Context:
    """One (sf, bw, osr, window) configuration on one GPU; owns the uploaded tables."""

    def __init__(self, sf, bw_hz=125000, osr=1, window=False, device=-1):
        self.lib = load_library()
        self.sf, self.bw_hz, self.osr, self.window = sf, bw_hz, max(osr, 1), bool(window)
        self.n = 1 << sf
        self.step = self.n * self.osr
        cfg = Config(sf, bw_hz, osr, int(bool(window)), device)
        h = C.c_void_p()
        _check(self.lib.lora_b200_ctx_create(C.byref(cfg), C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.lora_b200_ctx_destroy(self.h)
            self.h = None

    __del__ = close

    def set_precision(self, precision):
        """PRECISION_FAST (default): fused kernel + bit-faithful recheck of near-tie windows; PRECISION_EXACT:
        bit-faithful kernel on every window (peak values bit-identical too)."""
        _check(self.lib.lora_b200_set_precision(self.h, int(precision)))

    def set_overlap(self, enable):
        """lora_b200_set_overlap: how prepass and demod are scheduled (results never depend on it).  0 two launches,
        1 scan of sub-batch k+1 on a second stream, 2 (default) one persistent launch for calls of >= 600 MB of
        samples, 3 the same without the size threshold."""
        _check(self.lib.lora_b200_set_overlap(self.h, int(enable)))

    def recheck_stats(self):
        """-> (windows demodulated in FAST precision, windows rechecked) since the last call."""
        w, r = C.c_ulonglong(), C.c_ulonglong()
        _check(self.lib.lora_b200_recheck_stats(self.h, C.byref(w), C.byref(r)))
        return w.value, r.value

    def set_kernel_timing(self, enable):
        _check(self.lib.lora_b200_set_kernel_timing(self.h, int(enable)))

    def kernel_time(self):
        """-> (summed ms, launches) of the fused demod kernel since the last call."""
        ms, k = C.c_double(), C.c_ulonglong()
        _check(self.lib.lora_b200_kernel_time(self.h, C.byref(ms), C.byref(k)))
        return ms.value, int(k.value)

    def table(self, which):
        out = np.zeros(self.n, np.float32 if which == 2 else np.complex64)
        _check(self.lib.lora_b200_ctx_table(self.h, which, out.ctypes.data_as(C.c_void_p), self.n))
        return out

    # ---- device-tensor batch API ------------------------------------------------------------------
    def demod_batch(self, iq, mode=DEMOD_WORKSPACE, want_peaks=False, symbols=None, metrics=None, peaks=None):
        """iq: CUDA complex64 tensor [n_packets, samples].  -> (symbols u16 [n, total-2], metrics, peaks)."""
        import torch
        assert iq.is_cuda and iq.dtype == torch.complex64 and iq.dim() == 2 and iq.stride(1) == 1
        n, count = iq.shape
        total = count // self.step
        n_out = total - 2 if total >= 2 else total
        if symbols is None:
            symbols = torch.zeros((n, max(n_out, 1)), dtype=torch.int16, device=iq.device)
        if metrics is None:
            metrics = torch.zeros((n, 16), dtype=torch.uint8, device=iq.device)
        if want_peaks and peaks is None:
            peaks = torch.zeros((n, max(total, 1)), dtype=torch.float32, device=iq.device)
        _check(self.lib.lora_b200_demod_batch(self.h, mode, _ptr(iq), n, count, iq.stride(0), _ptr(symbols),
                                              symbols.stride(0), _ptr(peaks), _ptr(metrics), _stream()))
        return symbols[:, :n_out], metrics, peaks

    def demod_batch_static(self, iq, symbols, metrics=None, peaks=None, mode=DEMOD_WORKSPACE):
        """lora_b200_demod_batch_static: the caller passes every buffer; nothing is allocated, here or in the library."""
        n, count = iq.shape
        _check(self.lib.lora_b200_demod_batch_static(self.h, mode, _ptr(iq), n, count, iq.stride(0), _ptr(symbols),
                                                     symbols.stride(0), _ptr(peaks), _ptr(metrics), _stream()))

    def estimate_batch(self, iq):
        import torch
        n, count = iq.shape
        metrics = torch.zeros((n, 16), dtype=torch.uint8, device=iq.device)
        _check(self.lib.lora_b200_estimate_batch(self.h, _ptr(iq), n, count, iq.stride(0), _ptr(metrics), _stream()))
        return metrics

    def decode_batch(self, symbols):
        """lora_decode + CRC flag.  symbols: CUDA int16 [n, nsym] -> (bytes u8 [n, nsym//2], crc_ok u8 [n])."""
        import torch
        n, nsym = symbols.shape
        out = torch.zeros((n, max(nsym // 2, 1)), dtype=torch.uint8, device=symbols.device)
        ok = torch.zeros(n, dtype=torch.uint8, device=symbols.device)
        _check(self.lib.lora_b200_decode_batch(self.h, _ptr(symbols), n, nsym, symbols.stride(0), _ptr(out),
                                               out.stride(0), _ptr(ok), _stream()))
        return out[:, : nsym // 2], ok

    def encode_batch(self, data):
        import torch
        n, nb = data.shape
        sym = torch.zeros((n, max(2 * nb, 1)), dtype=torch.int16, device=data.device)
        _check(self.lib.lora_b200_encode_batch(self.h, _ptr(data), n, nb, data.stride(0), _ptr(sym),
                                               sym.stride(0), _stream()))
        return sym[:, : 2 * nb]

    def chain_decode_batch(self, cfg, symbols, payload=None, results=None):
        import torch
        n, nsym = symbols.shape
        if payload is None:
            payload = torch.zeros((n, PAYLOAD_STRIDE), dtype=torch.uint8, device=symbols.device)
        if results is None:
            results = torch.zeros((n, 16), dtype=torch.uint8, device=symbols.device)
        _check(self.lib.lora_b200_chain_decode_batch(self.h, C.byref(cfg), _ptr(symbols), n, nsym,
                                                     symbols.stride(0), _ptr(payload), _ptr(results), _stream()))
        return payload, results

    def chain_symbol_count(self, cfg, payload_len):
        k = C.c_size_t()
        _check(self.lib.lora_b200_chain_encode_batch(self.h, C.byref(cfg), None, 0, payload_len, payload_len,
                                                     None, 0, C.byref(k), None))
        return k.value

    def chain_encode_batch(self, cfg, payload):
        import torch
        n, plen = payload.shape
        nsym = self.chain_symbol_count(cfg, plen)
        sym = torch.zeros((n, nsym), dtype=torch.int16, device=payload.device)
        k = C.c_size_t()
        _check(self.lib.lora_b200_chain_encode_batch(self.h, C.byref(cfg), _ptr(payload), n, plen,
                                                     payload.stride(0), _ptr(sym), sym.stride(0), C.byref(k), _stream()))
        return sym

    def modulate_batch(self, symbols, amplitude=1.0, sync=0x12, out=None):
        import torch
        n, nsym = symbols.shape
        count = (nsym + 2) * self.step
        if out is None:
            out = torch.zeros((n, count), dtype=torch.complex64, device=symbols.device)
        _check(self.lib.lora_b200_modulate_batch(self.h, _ptr(symbols), n, nsym, symbols.stride(0), amplitude,
                                                 sync, _ptr(out), out.stride(0), _stream()))
        return out

    # ---- continuous streams and full frames (LoRa-SDR/LoRaMod.cpp:109-238, LoRaDemod.cpp:145-327) ----------
    def frame_samples(self, n_symbols, padding=0):
        return self.lib.lora_b200_frame_samples(self.h, n_symbols, padding)

    def frame_modulate_batch(self, symbols, amplitude=1.0, sync=0x12, padding=0, out=None):
        """One full frame per row of `symbols`: preamble, sync word, down-chirps, data, zero padding."""
        import torch
        n, nsym = symbols.shape
        count = self.frame_samples(nsym, padding)
        if out is None:
            out = torch.zeros((n, count), dtype=torch.complex64, device=symbols.device)
        _check(self.lib.lora_b200_frame_modulate_batch(self.h, _ptr(symbols), n, nsym, symbols.stride(0), amplitude, sync,
                                                       padding, _ptr(out), out.stride(0), _stream()))
        return out

    def stream_detect_batch(self, iq, sync=0x12, thresh_db=-30.0, mtu=256, max_packets=8, max_symbols=1024,
                            frame_header=0):
        """LoRaDemod::work over every row of `iq` (streams x samples).  -> (symbols int16 [streams, max_symbols],
        packets uint8 [streams, max_packets, 24] (view with STREAM_PACKET), results uint8 [streams, 16] (STREAM_RESULT))."""
        import torch
        n, count = iq.shape
        sym = torch.zeros((n, max_symbols), dtype=torch.int16, device=iq.device)
        pk = torch.zeros((n, max_packets, STREAM_PACKET.itemsize), dtype=torch.uint8, device=iq.device)
        res = torch.zeros((n, STREAM_RESULT.itemsize), dtype=torch.uint8, device=iq.device)
        cfg = StreamCfg(sync, thresh_db, mtu, max_packets, max_symbols, frame_header)
        stride = iq.stride(0) if n > 1 else max(iq.stride(0), count)  # a size-1 dimension may report any stride
        _check(self.lib.lora_b200_stream_detect_batch(self.h, C.byref(cfg), _ptr(iq), n, count, stride, _ptr(sym),
                                                      _ptr(pk), _ptr(res), _stream()))
        return sym, pk, res

    def stream_decode_batch(self, cfg, symbols, packets, results):
        """LoRaDecoder::work on every packet of stream_detect_batch's output -> (payload u8 [streams, max_packets, 264],
        chain results u8 [streams, max_packets, 16]); empty slots have n_bytes = -1."""
        import torch
        n, max_packets, max_symbols = packets.shape[0], packets.shape[1], symbols.shape[1]
        payload = torch.zeros((n, max_packets, PAYLOAD_STRIDE), dtype=torch.uint8, device=symbols.device)
        res = torch.zeros((n, max_packets, 16), dtype=torch.uint8, device=symbols.device)
        _check(self.lib.lora_b200_stream_decode_batch(self.h, C.byref(cfg), _ptr(symbols), _ptr(packets), _ptr(results), n,
                                                      max_packets, max_symbols, _ptr(payload), _ptr(res), _stream()))
        return payload, res

    # ---- the deployed LoRa coding chain (csrc/std_kernels.cu) -----------------------
    def std_symbol_count(self, cfg, payload_len):
        k = C.c_size_t()
        _check(self.lib.lora_b200_std_encode_batch(self.h, C.byref(cfg), None, 0, payload_len, payload_len, None, 0,
                                                   C.byref(k), None))
        return k.value

    def std_encode_batch(self, cfg, payload):
        import torch
        n, plen = payload.shape
        sym = torch.zeros((n, self.std_symbol_count(cfg, plen)), dtype=torch.int16, device=payload.device)
        _check(self.lib.lora_b200_std_encode_batch(self.h, C.byref(cfg), _ptr(payload), n, plen, payload.stride(0),
                                                   _ptr(sym), sym.stride(0), None, _stream()))
        return sym

    def std_decode_batch(self, cfg, symbols):
        """-> (payload u8 [n, 264], results u8 [n, 16] viewed as STD_RESULT)"""
        import torch
        n, nsym = symbols.shape
        payload = torch.zeros((n, PAYLOAD_STRIDE), dtype=torch.uint8, device=symbols.device)
        res = torch.zeros((n, 16), dtype=torch.uint8, device=symbols.device)
        _check(self.lib.lora_b200_std_decode_batch(self.h, C.byref(cfg), _ptr(symbols), n, nsym, symbols.stride(0),
                                                   _ptr(payload), _ptr(res), _stream()))
        return payload, res

    def std_stream_decode_batch(self, cfg, symbols, packets, results):
        """The deployed chain on every packet of stream_detect_batch's output (layout as stream_decode_batch)."""
        import torch
        n, max_packets, max_symbols = packets.shape[0], packets.shape[1], symbols.shape[1]
        payload = torch.zeros((n, max_packets, PAYLOAD_STRIDE), dtype=torch.uint8, device=symbols.device)
        res = torch.zeros((n, max_packets, 16), dtype=torch.uint8, device=symbols.device)
        _check(self.lib.lora_b200_std_stream_decode_batch(self.h, C.byref(cfg), _ptr(symbols), _ptr(packets),
                                                          _ptr(results), n, max_packets, max_symbols, _ptr(payload),
                                                          _ptr(res), _stream()))
        return payload, res

    def std_receive_stream_host(self, cfg, iq, sync=0x12, thresh_db=-30.0, mtu=0, max_packets=64):
        """One host capture -> (list of payload bytes, STD_RESULT records, STREAM_PACKET records); mtu 0 = from the headers."""
        iq = np.ascontiguousarray(iq, np.complex64)
        sc = StreamCfg(sync, thresh_db, mtu, max_packets, len(iq) // self.step + 8)
        payload = np.zeros((max_packets, PAYLOAD_STRIDE), np.uint8)
        res = np.zeros(max_packets, STD_RESULT)
        pk = np.zeros(max_packets, STREAM_PACKET)
        n = C.c_uint32()
        _check(self.lib.lora_b200_std_receive_stream_host(self.h, C.byref(cfg), C.byref(sc), iq.ctypes.data_as(C.c_void_p),
                                                          len(iq), payload.ctypes.data_as(C.c_void_p),
                                                          res.ctypes.data_as(C.c_void_p), pk.ctypes.data_as(C.c_void_p),
                                                          C.byref(n)))
        k = n.value
        return [bytes(payload[i, :max(int(res[i]["n_bytes"]), 0)]) for i in range(k)], res[:k], pk[:k]

    def awgn_batch(self, iq, snr_db, seed=0):
        n, count = iq.shape
        _check(self.lib.lora_b200_awgn_batch(self.h, _ptr(iq), n, count, iq.stride(0), _ptr(snr_db), seed, _stream()))
        return iq

    # ---- host-buffer API (what include/lora_phy/phy.hpp forwards to) ------------------------------------
    def _demod_host(self, mode, iq):
        iq = np.ascontiguousarray(iq, np.complex64)
        total = len(iq) // self.step
        sym = np.zeros(max(total, 1), np.uint16)
        m = np.zeros(1, PKT_METRICS)
        k = C.c_size_t()
        rc = self.lib.lora_b200_demodulate_host(self.h, mode, iq.ctypes.data_as(C.c_void_p), len(iq),
                                                sym.ctypes.data_as(C.c_void_p), len(sym), C.byref(k),
                                                m.ctypes.data_as(C.c_void_p))
        return rc, sym[: k.value], m[0]

    def demodulate(self, iq):
        """lora_phy::demodulate (phy.cpp:178-239): -> (ret, symbols, metrics); ret = -1 on the reference's guards."""
        rc, sym, m = self._demod_host(DEMOD_WORKSPACE, iq)
        if rc == -1:
            return -1, sym[:0], m
        _check(rc)
        return len(sym), sym, m

    def lora_demodulate(self, samples, dechirp=False, scratch=True):
        """lora_phy::lora_demodulate (LoRaDemod.cpp:49-195): -> (ret, symbols, sync, metrics).
        dechirp=True fuses the caller-side dechirp of tests/e2e_chain_test.cpp:79-93 in front."""
        rc, sym, m = self._demod_host(DEMOD_LEGACY_DECHIRP if dechirp else DEMOD_LEGACY, samples)
        _check(rc)
        if m["normalised"] and not scratch:  # LoRaDemod.cpp:69-71: no scratch buffer -> 0 symbols
            return 0, sym[:0], 0, m
        total = len(samples) // self.step
        return (len(sym) if total >= 2 else total), sym, int(m["sync_word"]), m

    def estimate_offsets(self, samples):
        samples = np.ascontiguousarray(samples, np.complex64)
        cfo, to = C.c_float(0), C.c_float(0)
        _check(self.lib.lora_b200_estimate_host(self.h, samples.ctypes.data_as(C.c_void_p), len(samples),
                                                C.byref(cfo), C.byref(to)))
        return np.float32(cfo.value), np.float32(to.value)

    def fft(self, windows, inverse=False):
        """kissfft<float>::transform (kissfft.hh:100-185) of every row of `windows` (host array, N columns)."""
        x = np.ascontiguousarray(windows, np.complex64).reshape(-1, self.n)
        out = np.zeros_like(x)
        _check(self.lib.lora_b200_fft_detect_host(self.h, x.ctypes.data_as(C.c_void_p), len(x), int(inverse),
                                                  out.ctypes.data_as(C.c_void_p), None))
        return out

    def detect(self, windows, want_spectrum=False):
        """LoRaDetector<float>::detect (LoRaDetector.hpp:39-74) per row -> DETECT_RESULT records [, spectrum]."""
        x = np.ascontiguousarray(windows, np.complex64).reshape(-1, self.n)
        res = np.zeros(len(x), DETECT_RESULT)
        spec = np.zeros_like(x) if want_spectrum else None
        _check(self.lib.lora_b200_fft_detect_host(self.h, x.ctypes.data_as(C.c_void_p), len(x), 0,
                                                  spec.ctypes.data_as(C.c_void_p) if want_spectrum else None,
                                                  res.ctypes.data_as(C.c_void_p)))
        return (res, spec) if want_spectrum else res

    def compensate_offsets(self, cfo, time_offset, samples):
        """lora_phy::compensate_offsets (phy.cpp:147-176) on a copy of `samples`."""
        out = np.array(samples, np.complex64)
        _check(self.lib.lora_b200_compensate_host(self.h, float(cfo), float(time_offset), out.ctypes.data_as(C.c_void_p),
                                                  len(out)))
        return out

    def encode(self, payload):
        """lora_phy::encode / lora_encode with host buffers -> uint16 symbols."""
        payload = np.ascontiguousarray(payload, np.uint8)
        out = np.zeros(2 * len(payload), np.uint16)
        k = C.c_size_t()
        _check(self.lib.lora_b200_encode_host(self.h, payload.ctypes.data_as(C.c_void_p), len(payload),
                                              out.ctypes.data_as(C.c_void_p), len(out), C.byref(k)))
        return out[:k.value]

    def decode(self, symbols):
        """lora_phy::decode with host buffers -> (bytes, crc_ok)."""
        symbols = np.ascontiguousarray(symbols, np.uint16)
        out = np.zeros(len(symbols) // 2, np.uint8)
        k, ok = C.c_size_t(), C.c_int()
        _check(self.lib.lora_b200_decode_host(self.h, symbols.ctypes.data_as(C.c_void_p), len(symbols),
                                              out.ctypes.data_as(C.c_void_p), len(out), C.byref(k), C.byref(ok)))
        return out[:k.value], bool(ok.value)

    def lora_modulate(self, symbols, amplitude=1.0, sync=0x12):
        symbols = np.ascontiguousarray(symbols, np.uint16)
        out = np.zeros((len(symbols) + 2) * self.step, np.complex64)
        k = C.c_size_t()
        _check(self.lib.lora_b200_modulate_host(self.h, symbols.ctypes.data_as(C.c_void_p) if len(symbols) else None,
                                                len(symbols), amplitude, sync, out.ctypes.data_as(C.c_void_p),
                                                len(out), C.byref(k)))
        return out[: k.value]

    def receive_host(self, cfg, iq, mode=DEMOD_LEGACY_DECHIRP, payload=None, results=None, metrics=None):
        """Host IQ [n, samples] -> (payload u8 [n, 264], results, metrics): demod + chain decode end to end."""
        iq = np.ascontiguousarray(iq, np.complex64)
        n, count = iq.shape
        payload = np.zeros((n, PAYLOAD_STRIDE), np.uint8) if payload is None else payload
        results = np.zeros(n, CHAIN_RESULT) if results is None else results
        metrics = np.zeros(n, PKT_METRICS) if metrics is None else metrics
        _check(self.lib.lora_b200_receive_host(self.h, mode, C.byref(cfg), iq.ctypes.data_as(C.c_void_p), n, count,
                                               payload.ctypes.data_as(C.c_void_p), results.ctypes.data_as(C.c_void_p),
                                               metrics.ctypes.data_as(C.c_void_p)))
        return payload, results, metrics

    def receive_host_i16(self, cfg, iq_i16, scale=1.0 / 32768.0, mode=DEMOD_LEGACY_DECHIRP, payload=None, results=None,
                         metrics=None):
        """Host int16 IQ [n, samples, 2] (I, Q interleaved) -> the same outputs as receive_host on float32(iq) * scale."""
        iq_i16 = np.ascontiguousarray(iq_i16, np.int16)
        n, count = iq_i16.shape[0], iq_i16.shape[1]
        payload = np.zeros((n, PAYLOAD_STRIDE), np.uint8) if payload is None else payload
        results = np.zeros(n, CHAIN_RESULT) if results is None else results
        metrics = np.zeros(n, PKT_METRICS) if metrics is None else metrics
        _check(self.lib.lora_b200_receive_host_i16(self.h, mode, C.byref(cfg), iq_i16.ctypes.data_as(C.c_void_p), n, count,
                                                   scale, payload.ctypes.data_as(C.c_void_p),
                                                   results.ctypes.data_as(C.c_void_p), metrics.ctypes.data_as(C.c_void_p)))
        return payload, results, metrics


    def receive_host_i8(self, cfg, iq_i8, scale=1.0 / 128.0, mode=DEMOD_LEGACY_DECHIRP, payload=None, results=None,
                        metrics=None):
        """Host int8 IQ [n, samples, 2] (I, Q interleaved) -> the same outputs as receive_host on float32(iq) * scale."""
        iq_i8 = np.ascontiguousarray(iq_i8, np.int8)
        n, count = iq_i8.shape[0], iq_i8.shape[1]
        payload = np.zeros((n, PAYLOAD_STRIDE), np.uint8) if payload is None else payload
        results = np.zeros(n, CHAIN_RESULT) if results is None else results
        metrics = np.zeros(n, PKT_METRICS) if metrics is None else metrics
        _check(self.lib.lora_b200_receive_host_i8(self.h, mode, C.byref(cfg), iq_i8.ctypes.data_as(C.c_void_p), n, count,
                                                  scale, payload.ctypes.data_as(C.c_void_p),
                                                  results.ctypes.data_as(C.c_void_p), metrics.ctypes.data_as(C.c_void_p)))
        return payload, results, metrics


def stream_records(packets, results):
    """Device tensors of stream_detect_batch -> numpy record arrays (packets [streams, max_packets], results [streams])."""
    p = packets.cpu().numpy()
    r = results.cpu().numpy()
    return (p.reshape(p.shape[0], -1).view(STREAM_PACKET).reshape(p.shape[0], p.shape[1]),
            r.reshape(r.shape[0], -1).view(STREAM_RESULT).reshape(r.shape[0]))


def metrics_to_numpy(t):
    """CUDA uint8 [n,16] metrics / results tensor -> structured numpy array (pass dtype via .view)."""
    return t.cpu().numpy()
